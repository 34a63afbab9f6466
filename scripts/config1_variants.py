#!/usr/bin/env python
"""Config 1 (ML-100K-shaped, rank 10, 10 epochs) with the direct and the combine kernel on one GPU."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "benchmarks"))
from run_configs import timed, zipf_items
from fps_b200.models.mf.device import DeviceMF
nu, ni, n, dim, epochs = 943, 1682, 100_000, 10, 10
g = torch.Generator().manual_seed(1)
u = torch.randint(0, nu, (n,), generator=g).int()
i = zipf_items(n, ni, g, 2.0).int()
tu, ti = torch.randn(nu, 4, generator=g), torch.randn(ni, 4, generator=g)
r = ((tu[u.long()] * ti[i.long()]).sum(1) * 0.1 + 0.6).clamp(0.2, 1.0).float()
du, di, dr = u.cuda(), i.cuda(), r.cuda()
for variant in ("direct", "combine"):
    mf = DeviceMF(nu, ni, dim, -0.01, 0.01, 0.01, seed=1, variant=variant)
    def job():
        for _ in range(epochs):
            mf.step(du, di, dr)
    ms = timed(job, 2, 5)
    fresh = DeviceMF(nu, ni, dim, -0.01, 0.01, 0.01, seed=1, variant=variant)
    for _ in range(epochs):
        fresh.step(du, di, dr)
    print(variant, f"{ms:.3f} ms per job, {n * epochs / ms / 1e6:.2f} G ratings/s, rmse {fresh.rmse(du, di, dr):.4f}", flush=True)
