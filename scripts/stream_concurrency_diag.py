#!/usr/bin/env python
"""Which producer-side torch operations stall while the persistent ring consumer is resident?"""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.models.mf.device import DeviceMF
from fps_b200.models.mf.stream import OnlineStream

lazy = len(sys.argv) > 1 and sys.argv[1] == "lazy"
mf = DeviceMF(2000, 1000, 8, 0.0, 0.2, 0.05, pull_limit=1600, lazy=lazy, variant="direct")
x = torch.arange(100000, device="cuda")
seen = torch.zeros(1000, dtype=torch.bool, device="cuda")
gen = torch.Generator(device="cuda").manual_seed(1)
def ops():
    return {
        "add+stream_sync": lambda: (x + 1, torch.cuda.current_stream().synchronize()),
        "nonzero": lambda: torch.nonzero(seen),
        "randint": lambda: (torch.randint(0, 10, (100, 3), device="cuda", generator=gen), torch.cuda.current_stream().synchronize()),
        "index_put": lambda: (seen.__setitem__(x[:10].long() % 1000, True), torch.cuda.current_stream().synchronize()),
        "cat": lambda: (torch.cat([x[:10], x[:10]]), torch.cuda.current_stream().synchronize()),
        ".cpu()": lambda: x[:100].cpu(),
        "new alloc 64MB": lambda: (torch.empty(16 << 20, device="cuda"), torch.cuda.current_stream().synchronize()),
    }
for name, f in ops().items():      # first use (module loads) without the kernel
    f()
torch.cuda.synchronize()
st = OnlineStream(mf, capacity=1 << 14, iteration_wait_ms=1500, max_lifetime_ms=30_000)
import numpy as np
st.feed(np.zeros(64, np.int32), np.zeros(64, np.int32), np.zeros(64, np.float32))   # kernel is now resident
time.sleep(0.05)
for name, f in ops().items():
    t0 = time.time()
    f()
    dt = (time.time() - t0) * 1e3
    st.feed(np.zeros(64, np.int32), np.zeros(64, np.int32), np.zeros(64, np.float32))   # keeps it alive / re-arms
    print(f"{name:18s} {dt:9.2f} ms   (relaunches so far {st.relaunches})", flush=True)
print("exit", st.join())
