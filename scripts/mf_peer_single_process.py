#!/usr/bin/env python
"""Single-process, two-GPU diagnosis of the fused MF kernel's remote path (needs 2 visible GPUs):
item shard 1 lives on cuda:1 and is reached through cudaDeviceEnablePeerAccess, so the kernel can
be timed - and profiled with ncu - without a second rank."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fps_b200.models.mf.device import DeviceMF  # noqa: E402
from fps_b200.ops import _native as N  # noqa: E402
from fps_b200.parallel.device_context import DeviceContext  # noqa: E402


def main():
    torch.cuda.set_device(0)
    ctx = DeviceContext(0, 1, 0, torch.device("cuda", 0), None)
    n_users, n_items, B = 4_000_000, 10_000_000, 1 << 21
    mf = DeviceMF(n_users, n_items, 128, -0.01, 0.01, 0.01, ctx=ctx, n_shards=2)
    # move item shard 1 to GPU 1
    N.check(N.lib().fps_enable_peer_access(1), "peer access")
    remote = torch.empty((mf.items.rows_per_shard, mf.items.stride), device="cuda:1", dtype=torch.float32)
    remote.copy_(mf.items.rows[1])
    torch.cuda.synchronize()
    mf.items._view.rows[1] = remote.data_ptr()
    g = torch.Generator().manual_seed(5)
    users = torch.randint(0, n_users, (B,), generator=g).int().cuda()
    base = torch.randint(0, n_items // 2, (B,), generator=g)
    sets = {"all_local": (base * 2).int().cuda(), "all_remote": (base * 2 + 1).int().cuda()}
    ratings = torch.rand(B, generator=g).cuda()
    only = os.environ.get("ONLY")
    for name, items in sets.items():
        if only and name != only:
            continue
        for _ in range(2):
            mf.step(users, items, ratings)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            mf.step(users, items, ratings)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        print(f"{name}: {ms:.3f} ms per {B} ratings ({B / ms / 1e6:.3f} G/s)", flush=True)
    s = mf.stats()
    print("records", s["records"], "bad", s["bad_ids"], flush=True)


if __name__ == "__main__":
    main()
