"""BL1 (all-to-all PS) host-side logic on CPU: two gloo ranks must produce exactly what a single
process produces on the union of their batches when no key is shared between the ranks' batches."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fps_b200.parallel.nccl_baseline import AllToAllPS
    ps = AllToAllPS(40, 30, 6, 0.05, init_range=(0.1, 0.3), seed=3)
    g = torch.Generator().manual_seed(100 + rank)
    # rank r owns users with user % world == r; items are disjoint between the ranks' batches
    users = torch.randint(0, 20, (64,), generator=g) * world + rank
    items = torch.randperm(15, generator=g)[:8].repeat(8) * world + ((rank + 1) % world)
    ratings = torch.rand(64, generator=g)
    items0, users0 = ps.items.clone(), ps.users.clone()
    ps.step(users, items, ratings)
    torch.save({"users": users, "items": items, "ratings": ratings, "items0": items0, "users0": users0,
                "items1": ps.items, "users1": ps.users, "loss": ps.loss_sum}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_all_to_all_ps_two_gloo_ranks(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(tmp_path / f"r{k}.pt") for k in range(world)]
    lr = 0.05
    # item shard s (owned by rank s) received the deltas of the *other* rank's batch
    for rank in range(world):
        d = r[rank]
        owner = (rank + 1) % world
        items_ref = r[owner]["items0"].clone()
        users_ref = d["users0"].clone()
        u_slot, i_slot = d["users"] // world, d["items"] // world
        u = users_ref[u_slot]
        v = items_ref[i_slot]
        e = d["ratings"] - (u * v).sum(1)
        users_ref.index_add_(0, u_slot, (lr * e)[:, None] * v)
        items_ref.index_add_(0, i_slot, (lr * e)[:, None] * u)
        torch.testing.assert_close(d["users1"], users_ref)
        torch.testing.assert_close(r[owner]["items1"], items_ref)
        assert float(d["loss"]) == pytest.approx(float((e.double() ** 2).sum()))
