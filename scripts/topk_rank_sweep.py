"""All-tiles time of the packed top-K pipeline at several ranks (4096 queries x 1M items, k = 100).
python scripts/topk_rank_sweep.py"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.ops.topk import PackedItems, lemp_topk_packed, sort_items_by_length
g = torch.Generator(device="cuda").manual_seed(6)
for dim in (128, 64, 32):
    q = torch.randn(4096, dim, device="cuda", generator=g)
    items = torch.randn(1_000_000, dim, device="cuda", generator=g)
    it_s, len_s, ids_s = sort_items_by_length(items)
    packed = PackedItems(it_s, len_s, ids_s)
    for _ in range(2):
        lemp_topk_packed(q, packed, 100, use_pruning=False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        lemp_topk_packed(q, packed, 100, use_pruning=False)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(f"rank {dim}: {ms:.3f} ms  ({2.0 * 4096 * 1e6 * dim / ms / 1e9:.0f} TFLOP/s fp32-equivalent)", flush=True)
