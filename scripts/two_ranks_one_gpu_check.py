"""Two ranks on ONE GPU (gloo control plane, CUDA-IPC peer mapping): the cross-process data plane - peer
pulls, remote red.add pushes, the fused MF step over a table sharded across the two processes - on a box that
has a single GPU.  Launched by tests/test_multi_gpu.py under torchrun.  Nothing here makes one process's
kernel wait for the other's (the two contexts are time-sliced on the GPU)."""
import os
import sys

os.environ["FPS_PEER_MAPPING"] = "ipc"          # VMM symmetric memory needs one GPU per rank (NCCL rendezvous)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np
import torch
import torch.distributed as dist

from fps_b200.models.mf.device import DeviceMF
from fps_b200.ops.table import ShardedTable
from fps_b200.parallel.device_context import init_context


def main():
    ctx = init_context(backend="gloo")
    rank, world = ctx.rank, ctx.world_size
    assert world == 2 and torch.cuda.current_device() == 0
    ok = True
    # 1. peer-mapped pull of both shards
    t = ShardedTable(4_000, 32, init_range=(-1.0, 1.0), seed=5, ctx=ctx)
    ids = torch.arange(4_000)
    want = t.expected_init(np.arange(4_000))
    got = t.gather(ids).cpu().numpy()
    ok1 = bool(np.allclose(got, want, atol=1e-7))
    print(f"[rank {rank}] pull of both processes' shards matches init: {ok1}", flush=True)
    # 2. pushes from both processes land on both shards
    torch.cuda.synchronize()
    ctx.barrier()
    t.scatter(ids, torch.full((4_000, 32), float(rank + 1)), "add")
    torch.cuda.synchronize()
    ctx.barrier()
    got2 = t.gather(ids).cpu().numpy()
    ok2 = bool(np.allclose(got2, want + 3.0, atol=1e-4))
    print(f"[rank {rank}] red.add from both processes landed: {ok2}", flush=True)
    # 3. fused MF over items sharded across the two processes
    n_users, n_items, dim = 20_000, 5_000, 32
    mf = DeviceMF(n_users, n_items, dim, 0.0, 0.2, 0.05, pull_limit=1600, seed=1, ctx=ctx, replicate_items=False)
    g = torch.Generator().manual_seed(10 + rank)
    n = 100_000
    users = (torch.randint(0, n_users // world, (n,), generator=g) * world + rank).int().cuda()
    items = torch.randint(0, n_items, (n,), generator=g).int().cuda()
    tu = torch.randn(n_users, 4, generator=torch.Generator().manual_seed(1))
    ti = torch.randn(n_items, 4, generator=torch.Generator().manual_seed(2))
    ratings = (tu[users.cpu().long()] * ti[items.cpu().long()]).sum(1).mul(0.3).add(0.5).float().cuda()
    r0 = mf.rmse(users, items, ratings)
    for _ in range(12):
        mf.step(users, items, ratings)
        torch.cuda.synchronize()
        ctx.barrier()
    r1 = mf.rmse(users, items, ratings)
    st = mf.stats()
    ok3 = r1 < 0.7 * r0 and st["nan_count"] == 0 and st["bad_ids"] == 0 and st["remote_pulls"] > 0
    print(f"[rank {rank}] fused MF across processes: rmse {r0:.4f} -> {r1:.4f}, remote pulls {st['remote_pulls']}: {ok3}", flush=True)
    flag = torch.tensor([1 if (ok1 and ok2 and ok3) else 0])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("TWO_RANKS_ONE_GPU", "PASS" if flag.item() == 1 else "FAIL", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if flag.item() == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
