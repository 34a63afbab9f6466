#!/usr/bin/env python
"""One packed tcgen05 top-K launch (for ncu): 2048 queries x 262144 items, rank 128, k = 100, no pruning."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.ops.topk import PackedItems, lemp_topk_packed
g = torch.Generator(device="cuda").manual_seed(0)
q = torch.randn(2048, 128, device="cuda", generator=g)
items = torch.randn(262144, 128, device="cuda", generator=g)
packed = PackedItems(items)
for _ in range(3):
    s, i = lemp_topk_packed(q, packed, 100, use_pruning=False)
torch.cuda.synchronize()
print("ok", float(s[0, 0]))
