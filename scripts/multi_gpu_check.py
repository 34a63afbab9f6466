#!/usr/bin/env python
"""Multi-GPU correctness + P2P probe (launch with torchrun, one rank per GPU):
  * every rank addresses every peer's shard through CUDA IPC (pull = P2P loads),
  * remote pushes land (red.add over NVLink),
  * the fused MF kernel trains with items sharded over all GPUs,
  * NVLink rooflines for the three access types (load / store / red.add) -> gpurun_out/p2p_probe.json
"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fps_b200.models.mf.device import DeviceMF  # noqa: E402
from fps_b200.ops import _native as N  # noqa: E402
from fps_b200.ops.table import ShardedTable, new_flow, read_flow  # noqa: E402
from fps_b200.parallel.device_context import init_context  # noqa: E402


def main():
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    ok = True

    # 1. peer-mapped pull
    t = ShardedTable(10_000, 64, init_range=(-1.0, 1.0), seed=5, ctx=ctx)
    ids = torch.arange(10_000)
    got = t.gather(ids).cpu().numpy()
    want = t.expected_init(np.arange(10_000))
    ok &= bool(np.allclose(got, want, atol=1e-7))
    print(f"[rank {rank}] peer pull of all shards matches init: {np.allclose(got, want, atol=1e-7)}", flush=True)

    # 2. remote push: every rank adds (rank+1) to every row
    ctx.barrier()
    t.scatter(ids, torch.full((10_000, 64), float(rank + 1)), "add")
    torch.cuda.synchronize()
    ctx.barrier()
    got2 = t.gather(ids).cpu().numpy()
    inc = world * (world + 1) / 2
    ok2 = bool(np.allclose(got2, want + inc, atol=1e-4))
    ok &= ok2
    print(f"[rank {rank}] remote red.add from all ranks landed: {ok2}", flush=True)

    # 3. fused MF across ranks
    n_users, n_items, dim = 200_000, 50_000, 64
    mf = DeviceMF(n_users, n_items, dim, 0.0, 0.2, 0.05, pull_limit=1600, seed=1, ctx=ctx)
    g = torch.Generator().manual_seed(10 + rank)
    n = 400_000
    users = (torch.randint(0, n_users // world, (n,), generator=g) * world + rank).int().cuda()
    items = torch.randint(0, n_items, (n,), generator=g).int().cuda()
    truth_u = torch.randn(n_users, 4, generator=torch.Generator().manual_seed(1))
    truth_i = torch.randn(n_items, 4, generator=torch.Generator().manual_seed(2))
    ratings = (truth_u[users.cpu().long()] * truth_i[items.cpu().long()]).sum(1).mul(0.3).add(0.5).float().cuda()
    r0 = mf.rmse(users, items, ratings)
    for _ in range(15):
        mf.step(users, items, ratings)
    torch.cuda.synchronize()
    ctx.barrier()
    r1 = mf.rmse(users, items, ratings)
    s = mf.stats()
    frac = s["remote_pulls"] / max(s["combined_rows"] or s["pulls"], 1)   # physical rows when the combine variant ran
    ok3 = r1 < 0.6 * r0 and abs(frac - (world - 1) / world) < 0.02 and s["bad_ids"] == 0 and s["max_in_flight"] <= 1600
    ok &= ok3
    print(f"[rank {rank}] MF rmse {r0:.4f} -> {r1:.4f}, remote fraction {frac:.3f}, max_in_flight {s['max_in_flight']}: {ok3}", flush=True)
    # every rank sees the same item rows
    probe = torch.arange(0, n_items, 97)
    mine = mf.items.gather(probe)
    gathered = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)
    ok4 = all(torch.equal(gathered[0], x) for x in gathered)
    ok &= ok4
    print(f"[rank {rank}] all ranks read identical item rows: {ok4}", flush=True)

    # 4. P2P probe (rank 0 drives; peer = shard 1)
    res = {}
    if world > 1:
        big = ShardedTable(64 * 1024 * 1024 // 64 * world, 64, init_range=(0.0, 0.0), ctx=ctx)   # 256 MiB / shard
        nbytes = big.rows_per_shard * big.stride * 4
        ctx.barrier()
        if rank == 0:
            local = big._view.rows[0]
            peer = big._view.rows[1]
            sink = torch.zeros(64, device="cuda")
            ms = C.c_float()
            for name, dst, src, mode in [("load_local", N.ptr(sink), local, 0), ("load_peer", N.ptr(sink), peer, 0),
                                         ("store_local", local, local, 1), ("store_peer", peer, peer, 1),
                                         ("red_local", local, local, 2), ("red_peer", peer, peer, 2)]:
                N.check(N.lib().fps_p2p_probe(dst, src, nbytes, mode, 5, C.byref(ms), N.current_stream()), name)
                res[name + "_gbs"] = nbytes * 5 / (ms.value * 1e-3) / 1e9
            print("[p2p probe GB/s]", json.dumps(res), flush=True)
            os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
            json.dump({"world": world, "bytes": nbytes, **res}, open(os.path.join(ROOT, "gpurun_out", f"p2p_probe_{world}gpu.json"), "w"))
        ctx.barrier()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("MULTI_GPU_CHECK", "PASS" if flag.item() == 1 else "FAIL", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if flag.item() == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
