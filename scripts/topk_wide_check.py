#!/usr/bin/env python
"""Wide top-K kernel: a small exactness check against torch, then config-6 timings of both layouts."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fps_b200.ops.topk import PackedItems, lemp_topk_packed, sort_items_by_length

g = torch.Generator(device="cuda").manual_seed(1)
for dim, n_q, n_items, k in ((128, 300, 20011, 100), (10, 37, 1000, 10), (96, 129, 70, 5)):
    q = torch.randn(n_q, dim, device="cuda", generator=g)
    items = torch.randn(n_items, dim, device="cuda", generator=g) * (0.2 + 2.0 * torch.rand(n_items, 1, device="cuda", generator=g) ** 3)
    it_s, len_s, ids_s = sort_items_by_length(items)
    ref = (q.double() @ items.double().T).topk(k, dim=1).values
    for wide in (False, True):
        packed = PackedItems(it_s, len_s, ids_s, wide=wide)
        for prune in (False, True):
            s, i = lemp_topk_packed(q, packed, k, use_pruning=prune)
            torch.cuda.synchronize()
            err = float(((s.double() - ref).abs() / (1.0 + ref.abs())).max())
            print(f"dim {dim} n_q {n_q} items {n_items} k {k} wide={wide} prune={prune}: max rel err vs fp64 = {err:.2e}", flush=True)
            assert err < 2e-5, err
if len(sys.argv) > 1 and sys.argv[1] == "config6":
    n_q, n_items = 4096, 1_000_000
    q = torch.randn(n_q, 128, device="cuda", generator=g)
    lens = torch.rand(n_items, 1, device="cuda", generator=g) ** 4 * 5 + 0.05
    items = torch.nn.functional.normalize(torch.randn(n_items, 128, device="cuda", generator=g), dim=1) * lens
    it_s, len_s, ids_s = sort_items_by_length(items)
    for wide in (False, True):
        packed = PackedItems(it_s, len_s, ids_s, wide=wide)
        stats = torch.zeros(4, device="cuda", dtype=torch.int64)
        for prune in (False, True):
            for _ in range(2):
                lemp_topk_packed(q, packed, 100, use_pruning=prune)
            stats.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(3):
                s, i = lemp_topk_packed(q, packed, 100, use_pruning=prune, stats=stats)
            e1.record()
            torch.cuda.synchronize()
            t, tt, acc, pr = stats.tolist()
            print(f"config6 wide={wide} prune={prune}: {e0.elapsed_time(e1) / 3:.3f} ms  tiles {t // 3}/{tt // 3}  "
                  f"accepted/query {acc / n_q / 3:.0f}  prunes/query {pr / n_q / 3:.1f}  top {float(s[0, 0]):.4f}", flush=True)
