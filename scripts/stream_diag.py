#!/usr/bin/env python
"""Drain rate of the persistent ring consumer in isolation: fill the ring once, then watch ring_done."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.models.mf.device import DeviceMF
from fps_b200.models.mf.stream import OnlineStream

nu, ni, dim = 138_493, 26_744, int(sys.argv[1]) if len(sys.argv) > 1 else 64
n = 2_000_000
mf = DeviceMF(nu, ni, dim, -0.01, 0.01, 0.01, pull_limit=0, seed=2, variant="direct")
g = torch.Generator().manual_seed(8)
u = torch.randint(0, nu, (n,), generator=g, dtype=torch.int32).numpy()
i = (torch.rand(n, generator=g) ** 3 * ni).int().numpy()
r = torch.rand(n, generator=g).numpy()
st = OnlineStream(mf, capacity=1 << 22, iteration_wait_ms=500, max_lifetime_ms=60_000)
t0 = time.time()
st.feed(u, i, r)
t1 = time.time()
print(f"feed of {n} records returned after {t1 - t0:.3f} s", flush=True)
last = 0
while True:
    d = st._retired()
    t = time.time() - t1
    if d != last:
        print(f"  t={t * 1e3:8.1f} ms  retired {d}  ({d / max(t, 1e-9) / 1e6:.2f} M/s)", flush=True)
        last = d
    if d >= n or t > 20:
        break
    time.sleep(0.02)
print("exit reason", st.join(), mf.stats()["records"])
