#!/usr/bin/env python
"""Packed tcgen05 top-K launches for ncu / quick timing.

    python scripts/topk_once.py            # 2048 x 262144 random items (worst case for the selection)
    python scripts/topk_once.py config6    # BASELINE config 6: 4096 x 1M, long-tailed lengths, length-sorted
"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.ops.topk import PackedItems, lemp_topk_packed, sort_items_by_length

mode = sys.argv[1] if len(sys.argv) > 1 else "random"
g = torch.Generator(device="cuda").manual_seed(6)
if mode == "config6":
    n_q, n_items = 4096, 1_000_000
    q = torch.randn(n_q, 128, device="cuda", generator=g)
    lens = torch.rand(n_items, 1, device="cuda", generator=g) ** 4 * 5 + 0.05
    items = torch.nn.functional.normalize(torch.randn(n_items, 128, device="cuda", generator=g), dim=1) * lens
    it_s, len_s, ids_s = sort_items_by_length(items)
    packed = PackedItems(it_s, len_s, ids_s)
else:
    n_q, n_items = 2048, 262144
    q = torch.randn(n_q, 128, device="cuda", generator=g)
    packed = PackedItems(torch.randn(n_items, 128, device="cuda", generator=g))
stats = torch.zeros(4, device="cuda", dtype=torch.int64)
for prune in ([False, True] if mode == "config6" else [False]):
    for _ in range(2):
        lemp_topk_packed(q, packed, 100, use_pruning=prune)
    stats.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    s, i = lemp_topk_packed(q, packed, 100, use_pruning=prune, stats=stats)
    e1.record()
    torch.cuda.synchronize()
    t, tt, acc, pr = stats.tolist()
    print(f"{mode} prune={prune}: {e0.elapsed_time(e1):.3f} ms  tiles {t}/{tt}  accepted/query {acc / n_q:.0f}  "
          f"prunes/query {pr / n_q:.1f}  top {float(s[0, 0]):.4f}")
