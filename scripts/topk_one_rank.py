"""One packed all-tiles top-K problem (4096 x 1M, k = 100) at a given rank, for ncu.  python scripts/topk_one_rank.py [rank]"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.ops.topk import PackedItems, lemp_topk_packed, sort_items_by_length
dim = int(sys.argv[1]) if len(sys.argv) > 1 else 32
g = torch.Generator(device="cuda").manual_seed(6)
q = torch.randn(4096, dim, device="cuda", generator=g)
items = torch.randn(1_000_000, dim, device="cuda", generator=g)
packed = PackedItems(*sort_items_by_length(items))
for _ in range(3):
    lemp_topk_packed(q, packed, 100, use_pruning=False)
torch.cuda.synchronize()
print("done")
