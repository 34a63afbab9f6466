"""Cycles per tcgen05.mma kind::tf32 (M128 x N x K8) instruction for the issue patterns of fps_umma_probe, on one
CTA and on one CTA per SM.  python scripts/umma_probe.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.ops import _native as N
names = {0: "TS N128, 1 accumulator", 1: "TS N128, 2 accumulators", 2: "SS N128, 1 accumulator", 3: "SS N128, 2 accumulators",
         4: "TS N256, 1 accumulator", 5: "TS N64, 4 accumulators", 6: "TS N128, 2 issuing warps"}
sms = torch.cuda.get_device_properties(0).multi_processor_count
for ctas in (1, sms):
    for mode in range(7):
        out = torch.zeros(ctas, device="cuda")
        for _ in range(2):
            N.check(N.lib().fps_umma_probe(mode, 20000, ctas, N.ptr(out), N.current_stream()), "probe")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        N.check(N.lib().fps_umma_probe(mode, 20000, ctas, N.ptr(out), N.current_stream()), "probe")
        e1.record(); torch.cuda.synchronize()
        n = {4: 256, 5: 64}.get(mode, 128)
        cyc = out.mean().item()
        flop = 2.0 * 128 * n * 8
        print(f"ctas {ctas:3d}  {names[mode]:26s}  {cyc:7.1f} cycles/MMA  {flop / cyc:7.0f} FLOP/clk/SM  kernel {e0.elapsed_time(e1):.3f} ms"
              f" -> {flop * 20000 * ctas / (e0.elapsed_time(e1) * 1e-3) / 1e12:7.1f} TFLOP/s", flush=True)
