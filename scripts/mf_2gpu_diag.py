#!/usr/bin/env python
"""Diagnose the fused MF kernel on 2+ GPUs: time it with only local items, only remote items and
the natural mix, with one rank active or all ranks active (torchrun)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fps_b200.models.mf.device import DeviceMF  # noqa: E402
from fps_b200.parallel.device_context import init_context  # noqa: E402


def main():
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    n_users = int(os.environ.get("DIAG_USERS", 100_000_000))
    n_items = int(os.environ.get("DIAG_ITEMS", 10_000_000))
    B = 1 << 22
    mf = DeviceMF(n_users, n_items, 128, -0.01, 0.01, 0.01, ctx=ctx)
    g = torch.Generator().manual_seed(5 + rank)
    users = (torch.randint(0, n_users // world, (B,), generator=g) * world + rank).int().cuda()
    base = torch.randint(0, n_items // world, (B,), generator=g)
    batches = {
        "all_local": (base * world + rank).int().cuda(),
        "all_remote": (base * world + (rank + 1) % world).int().cuda(),
        "mix": torch.randint(0, n_items, (B,), generator=g).int().cuda(),
    }
    ratings = torch.rand(B, generator=g).cuda()
    res = {}

    def timeit(items, active):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if active:
            for _ in range(2):
                mf.step(users, items, ratings)
            e0.record()
            for _ in range(5):
                mf.step(users, items, ratings)
            e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        return e0.elapsed_time(e1) / 5 if active else 0.0

    for name, items in batches.items():
        for who in ("rank0_only", "all_ranks"):
            active = (who == "all_ranks") or rank == 0
            ms = timeit(items, active)
            if rank == 0:
                res[f"{name}/{who}"] = round(ms, 3)
                print(f"{name}/{who}: {ms:.3f} ms/step  ({B / ms / 1e6:.2f} G ratings/s/GPU)", flush=True)
    # tuning knobs on the worst case
    for bps in (2, 4, 8, 12):
        for unroll in (1, 2, 4):
            mf.params.blocks_per_sm, mf.params.unroll = bps, unroll
            ms = timeit(batches["mix"], True)
            if rank == 0:
                res[f"mix/all_ranks/bps{bps}/unroll{unroll}"] = round(ms, 3)
                print(f"mix/all_ranks bps={bps} unroll={unroll}: {ms:.3f} ms", flush=True)
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"mf_diag_{world}gpu.json"), "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
