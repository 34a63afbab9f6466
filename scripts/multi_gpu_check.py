#!/usr/bin/env python
"""Multi-GPU correctness + P2P probe (launch with torchrun, one rank per GPU):
  * every rank addresses every peer's shard through its VMM / IPC mapping (pull = P2P loads),
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

    # 2b. locked read-modify-write from every rank on the same few keys (N13: lock word in the row, system-scope
    #     CAS): non-atomic read - add 1 - write cycles; every cycle of every rank must count
    lt = ShardedTable(29, 6, init_range=(0.0, 0.0), aux_cols=1, ctx=ctx)
    n_cyc = 50_000
    lids = torch.randint(0, 29, (n_cyc,), generator=torch.Generator().manual_seed(100 + rank))
    all_counts = sum(torch.bincount(torch.randint(0, 29, (n_cyc,), generator=torch.Generator().manual_seed(100 + r)),
                                    minlength=29) for r in range(world)).float()
    ctx.barrier()
    lt.locked_axpby(lids, 1.0, torch.ones((n_cyc, 6)))
    torch.cuda.synchronize()
    ctx.barrier()
    lrows = lt.gather(torch.arange(29)).cpu()
    ok2b = bool(torch.equal(lrows[:, :6], all_counts[:, None].expand(29, 6))) and bool(torch.all(lrows[:, 6] == 0))
    ok &= ok2b
    print(f"[rank {rank}] locked read-modify-write cycles of all ranks all counted: {ok2b}", flush=True)

    # 3. fused MF across ranks
    n_users, n_items, dim = 200_000, 50_000, 64
    mf = DeviceMF(n_users, n_items, dim, 0.0, 0.2, 0.05, pull_limit=1600, seed=1, ctx=ctx, replicate_items=False)
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

    # 3c. rank 128 = the TMA-staged tile path reading PEER rows with cp.async.bulk: one conflict-free step
    #     whose items all live on the next rank, replayed in fp64 from rows gathered before it
    mt = DeviceMF(40_000, 40_000, 128, 0.0, 0.2, 0.05, seed=3, ctx=ctx, replicate_items=False, variant="tiles")
    k = torch.arange(3000, device="cuda", dtype=torch.int64)
    pu, pi = k * world + rank, k * world + (rank + 1) % world
    pr = torch.full((3000,), 0.5, device="cuda")
    ctx.barrier()
    U0, V0 = mt.users.gather(pu).double(), mt.items.gather(pi).double()
    torch.cuda.synchronize(); ctx.barrier()
    mt.step(pu.int(), pi.int(), pr)
    torch.cuda.synchronize(); ctx.barrier()
    U1, V1 = mt.users.gather(pu).double(), mt.items.gather(pi).double()
    e = (0.5 - (U0 * V0).sum(1))[:, None]
    err = max(float((U0 + 0.05 * e * V0 - U1).abs().max()), float((V0 + 0.05 * e * U0 - V1).abs().max()))
    st = mt.stats()
    ok3c = err < 1e-6 and st["records"] == 3000 and (world == 1 or st["remote_pulls"] == 3000) and mt.variant == "tiles"
    ok &= ok3c
    print(f"[rank {rank}] tile path (TMA rings over peer rows), fp64 replay max err {err:.2e}, remote pulls "
          f"{st['remote_pulls']}: {ok3c}", flush=True)

    # 3a. the same job with the (12.8 MB) item table replicated on every GPU: local pulls, deferred delta exchange
    if world > 1:
        mfr = DeviceMF(n_users, n_items, dim, 0.0, 0.2, 0.05, pull_limit=1600, seed=1, ctx=ctx, replicate_items=True)
        for _ in range(15):
            mfr.step(users, items, ratings)
        torch.cuda.synchronize()
        ctx.barrier()
        r2 = mfr.rmse(users, items, ratings)
        sr = mfr.stats()
        mine_r = mfr.items.rows[0][::97].clone()
        gathered_r = [torch.empty_like(mine_r) for _ in range(world)]
        dist.all_gather(gathered_r, mine_r)
        drift_r = max(float((x - gathered_r[0]).abs().max()) for x in gathered_r)
        ok3a = mfr.items.replicated and r2 < 0.6 * r0 and sr["remote_pulls"] == 0 and drift_r < 1e-3 and sr["max_in_flight"] <= 1600
        ok &= ok3a
        print(f"[rank {rank}] MF with replicated items: rmse {r0:.4f} -> {r2:.4f} (sharded {r1:.4f}), replica drift "
              f"{drift_r:.2e}, remote pulls {sr['remote_pulls']}: {ok3a}", flush=True)

    # 3b. replicated table: pulls are local, one multimem.red (or a red.add per replica) updates every copy
    if world > 1:
        for scalar, dim in ((True, 1), (False, 12)):
            rt = ShardedTable(5_000, dim, init_range=(0.0, 0.0), ctx=ctx, replicated=True, scalar=scalar, lazy=True)
            rid = torch.arange(0, 5_000, 3)
            ctx.barrier()
            rt.scatter(rid, torch.full((rid.numel(), dim), float(rank + 1)), "add")
            rt.scatter(rid[:50], torch.full((50, dim), 1.0), "add")
            torch.cuda.synchronize()
            ctx.barrier()
            local = rt.rows[0][:, :dim].cpu().numpy()
            want_r = np.zeros((5_000, dim), np.float32)
            want_r[rid.numpy()] = inc
            want_r[rid[:50].numpy()] += world
            ok5 = bool(np.array_equal(local, want_r))
            rt.load(torch.tensor([1, 2]), torch.full((2, dim), 7.0 + rank)) if rank == world - 1 else None
            torch.cuda.synchronize()
            ctx.barrier()
            ok5 &= bool(np.all(rt.rows[0][1:3, :dim].cpu().numpy() == 7.0 + world - 1))
            d_ids, d_rows = rt.dump()
            ok5 &= (d_ids.numel() == (rid.numel() + 2 if rank == 0 else 0))
            ok &= ok5
            print(f"[rank {rank}] replicated table (scalar={scalar}, multicast={rt.multicast}): every replica holds "
                  f"the sum of all pushes, load reaches all replicas, dump on rank 0 only: {ok5}", flush=True)
        from fps_b200.models.pa.device import DevicePA
        from fps_b200.models.pa.sparse import SparseVector
        rng = np.random.default_rng(100 + rank)
        w_true = np.random.default_rng(7).normal(size=300)
        vecs, labs = [], []
        for _ in range(4000):
            ix = np.sort(rng.choice(300, 12, replace=False))
            vv = rng.normal(size=12)
            vecs.append(SparseVector(ix, vv, 300))
            labs.append(bool(vv @ w_true[ix] > 0))
        pa = DevicePA(300, algorithm="binary", variant="pa_i", aggressiveness=0.5, ctx=ctx)
        for b0 in range(0, 4000, 500):
            pa.step_vectors(vecs[b0:b0 + 500], labs[b0:b0 + 500])
        torch.cuda.synchronize()
        ctx.barrier()
        pred = pa.step_vectors(vecs[:1000], None).cpu().numpy() > 0
        acc = float(np.mean(pred == np.asarray(labs[:1000])))
        wl = pa.table.rows[0].flatten().clone()
        ws = [torch.empty_like(wl) for _ in range(world)]
        dist.all_gather(ws, wl)
        drift = max(float((x - ws[0]).abs().max()) for x in ws)
        ok6 = pa.table.replicated and acc > 0.9 and drift < 1e-3 and pa.stats()["remote_pulls"] == 0
        ok &= ok6
        print(f"[rank {rank}] replicated binary PA: accuracy {acc:.3f}, replica drift {drift:.2e}, remote pulls "
              f"{pa.stats()['remote_pulls']}: {ok6}", flush=True)

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
