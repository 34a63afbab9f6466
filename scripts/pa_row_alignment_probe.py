"""Does row alignment matter for the multiclass PA kernel?  Config 4 shape with 100 classes (400-byte rows that
straddle 128-byte lines) against 128 classes (512-byte rows, line-aligned) and 96 (384 bytes).  python scripts/pa_row_alignment_probe.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.models.pa.device import DevicePA
from fps_b200.parallel.device_context import init_context
ctx = init_context()
F, nnz, B = 1_000_000, 100, 1 << 15
g = torch.Generator().manual_seed(4)
idx = torch.sort(torch.randint(0, F, (B, nnz), generator=g), dim=1).values.reshape(-1).int().cuda()
val = torch.randn(B * nnz, generator=g).cuda()
indptr = torch.arange(0, (B + 1) * nnz, nnz, dtype=torch.int64).cuda()
for L in (96, 100, 128, 64):
    pa = DevicePA(F, label_count=L, algorithm="ova", variant="pa_i", aggressiveness=0.1, range_partitioning=True, pull_limit=0, lazy=False, ctx=ctx)
    lab = torch.randint(0, L, (B,), generator=g).int().cuda()
    for _ in range(3):
        pa.step(indptr, idx, val, lab)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        pa.step(indptr, idx, val, lab)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(f"{L} classes ({L * 4} B rows): {ms:.3f} ms/step, {B * nnz * L * 4 * 2 / ms / 1e9:.2f} TB/s of row payload (pull + push)", flush=True)
    del pa
