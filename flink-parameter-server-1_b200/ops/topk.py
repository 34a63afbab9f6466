"""Python front-end of the LEMP top-K kernels (``csrc/topk.cu``).

``lemp_topk`` = tensor-core path (tcgen05 / TMEM, 3xTF32), ``topk_simt`` = CUDA-core fp32 path with
the same tiling and selection (baseline + on-device oracle), ``topk_merge`` = merge of partial
lists with an optional seen-set filter (``CollectTopKFromEachWorker.scala:40-48``).
"""
from __future__ import annotations

from typing import Optional, Tuple

from . import _native as N
from ..utils.observability import nvtx_range


def sort_items_by_length(items, ids=None):
    """Length-descending order (the LEMP bucket order, ``PSTopKGeneratorWorker.scala:19``).
    Returns ``(items_sorted, lengths_sorted, ids_sorted int32)``."""
    import torch
    lens = torch.linalg.vector_norm(items.float(), dim=1)
    order = torch.argsort(lens, descending=True)
    ids = torch.arange(items.shape[0], device=items.device, dtype=torch.int32) if ids is None else ids.int()
    return items[order].contiguous(), lens[order].contiguous(), ids[order].contiguous()


def _pad4(t):
    import torch
    if t.shape[1] % 4 == 0 and t.is_contiguous():
        return t
    pad = (-t.shape[1]) % 4
    return torch.nn.functional.pad(t, (0, pad)).contiguous()


def scratch_bytes(n_q: int, k: int, n_items: int, full: bool = False) -> int:
    """Scratch of ``fps_lemp_topk`` / ``fps_topk_simt``: the minimum (<= 16 item splits per query tile), or
    with ``full`` what lets a small query batch spread over all SMs (<= 64 splits; the kernel takes as many
    splits as the buffer it is given allows)."""
    L = N.lib()
    return int((L.fps_topk_scratch_bytes_full if full else L.fps_topk_scratch_bytes)(n_q, k, n_items))


def _run(fn_name: str, queries, items, k: int, item_len, item_ids, use_pruning: bool, scratch, stats):
    import torch
    dim = queries.shape[1]
    assert items.shape[1] == dim, "queries and items must have the same rank"
    q = _pad4(queries.float())
    it = _pad4(items.float())
    n_q, n_items = q.shape[0], it.shape[0]
    need = scratch_bytes(n_q, k, n_items)
    if scratch is None:
        scratch = torch.empty(scratch_bytes(n_q, k, n_items, full=True), device="cuda", dtype=torch.uint8)
    assert scratch.numel() >= need
    out_s = torch.empty((n_q, k), device="cuda", dtype=torch.float32)
    out_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
    L = N.lib()
    if fn_name == "tc":
        N.check(L.fps_lemp_topk(N.ptr(q), q.shape[1], n_q, N.ptr(it), it.shape[1], N.ptr(item_len),
                                N.ptr(item_ids), n_items, dim, k, int(use_pruning), N.ptr(scratch),
                                scratch.numel(), N.ptr(out_s), N.ptr(out_i), N.ptr(stats), N.current_stream()),
                "fps_lemp_topk")
    else:
        N.check(L.fps_topk_simt(N.ptr(q), q.shape[1], n_q, N.ptr(it), it.shape[1], N.ptr(item_ids), n_items,
                                dim, k, N.ptr(scratch), scratch.numel(), N.ptr(out_s), N.ptr(out_i),
                                N.current_stream()), "fps_topk_simt")
    return out_s, out_i


def lemp_topk(queries, items, k: int, *, item_len=None, item_ids=None, use_pruning: bool = True,
              scratch=None, stats=None):
    """Top-``k`` inner products of every query against ``items`` (already sorted by length
    descending if ``use_pruning``).  ``stats`` (uint64[4], optional): tiles scored, tiles total,
    candidates accepted, prunes."""
    import torch
    if use_pruning and item_len is None:
        item_len = torch.linalg.vector_norm(items.float(), dim=1).contiguous()
    return _run("tc", queries, items, k, item_len, item_ids, use_pruning, scratch, stats)


class PackedItems:
    """Item matrix pre-split (tf32 hi / lo) and pre-swizzled into 64-item tiles - the image the
    tensor cores read.  Built once per model refresh (``fps_topk_pack_items``), then streamed by
    the TMA engine on every query batch (``lemp_topk_packed``).  Costs 2x the fp32 item bytes."""

    def __init__(self, items_sorted, item_len=None, item_ids=None, wide=None):
        """``wide``: 128-item tiles for the wide kernel (default; ``FPS_TOPK_WIDE=0`` or ``wide=False`` =
        the 64-item layout of ``fps_lemp_topk_packed``)."""
        import os
        import torch
        it = _pad4(items_sorted.float())
        self.n_items, self.dim = it.shape[0], items_sorted.shape[1]
        self.wide = (os.environ.get("FPS_TOPK_WIDE", "1") != "0") if wide is None else bool(wide)
        L = N.lib()
        nbytes = int((L.fps_topk_wide_packed_bytes if self.wide else L.fps_topk_packed_bytes)(self.n_items, self.dim))
        if nbytes == 0:
            raise ValueError("packed top-K supports ranks up to 128")
        self.data = torch.empty(nbytes, device="cuda", dtype=torch.uint8)
        N.check((L.fps_topk_pack_items_wide if self.wide else L.fps_topk_pack_items)(
            N.ptr(it), it.shape[1], self.n_items, self.dim, N.ptr(self.data), N.current_stream()), "fps_topk_pack_items")
        self.item_len = (torch.linalg.vector_norm(items_sorted.float(), dim=1) if item_len is None else item_len).contiguous()
        self.item_ids = None if item_ids is None else item_ids.int().contiguous()


def lemp_topk_packed(queries, packed: PackedItems, k: int, *, use_pruning: bool = True, scratch=None, stats=None):
    """Same result as ``lemp_topk`` on the items ``packed`` was built from (sorted by length
    descending when ``use_pruning``)."""
    import torch
    assert queries.shape[1] == packed.dim
    q = _pad4(queries.float())
    n_q = q.shape[0]
    wide = getattr(packed, "wide", False)
    need = int((N.lib().fps_topk_wide_scratch_bytes if wide else N.lib().fps_topk_packed_scratch_bytes)(n_q, k, packed.n_items))
    if scratch is None or (wide and scratch.numel() < need):
        # room for every selection group's candidate lists (a smaller caller buffer = fewer groups, 64-item path)
        scratch = torch.empty(need, device="cuda", dtype=torch.uint8)
    out_s = torch.empty((n_q, k), device="cuda", dtype=torch.float32)
    out_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
    with nvtx_range("fps.topk.packed"):
        N.check((N.lib().fps_lemp_topk_wide if wide else N.lib().fps_lemp_topk_packed)(N.ptr(q), q.shape[1], n_q, N.ptr(packed.data), N.ptr(packed.item_len),
                                             N.ptr(packed.item_ids), packed.n_items, packed.dim, k, int(use_pruning),
                                             N.ptr(scratch), scratch.numel(), N.ptr(out_s), N.ptr(out_i),
                                             N.ptr(stats), N.current_stream()), "fps_lemp_topk_packed")
    return out_s, out_i


def topk_simt(queries, items, k: int, *, item_ids=None, scratch=None):
    return _run("simt", queries, items, k, None, item_ids, False, scratch, None)


def _strategy_code(strategy):
    """``LEMPPruningStrategy`` (or its string form) -> ``(code, n_focus, switch_threshold)`` of the C ABI."""
    from ..models.mf import pruning as P
    if isinstance(strategy, str):
        strategy = P.from_string(strategy)
    if isinstance(strategy, P.LENGTH):
        return 0, 0, 0.0
    if isinstance(strategy, P.COORD):
        return 1, 0, 0.0
    if isinstance(strategy, P.INCR):
        return 2, int(strategy.num_focus_coordinates), 0.0
    if isinstance(strategy, P.LC):
        return 3, 0, float(strategy.algorithm_switch_threshold)
    if isinstance(strategy, P.LI):
        return 4, int(strategy.num_focus_coordinates), float(strategy.algorithm_switch_threshold)
    raise TypeError(f"not a LEMP pruning strategy: {strategy!r}")


def lemp_keep_mask(queries, items_sorted, item_len, theta, strategy, bucket_size: int):
    """``keep[q, i]`` (bool): does item ``i`` pass the predicate its bucket gets under ``strategy`` for query
    ``q`` at threshold ``theta[q]`` - the device twins of ``lengthPruning`` / ``coordPruning`` /
    ``incrPruning`` (``pruning/LEMPPruningFunctions.scala:10-90``) chosen per bucket as in
    ``PSTopKGeneratorWorker.scala:82-96``.  Items sorted by length descending."""
    import torch
    code, n_focus, thr = _strategy_code(strategy)
    q = _pad4(queries.float())
    it = _pad4(items_sorted.float())
    n_q, n_items, dim = q.shape[0], it.shape[0], queries.shape[1]
    keep = torch.empty((n_q, n_items), device="cuda", dtype=torch.uint8)
    th = theta.float().contiguous()
    ln = item_len.float().contiguous()
    N.check(N.lib().fps_lemp_keep_mask(N.ptr(q), q.shape[1], n_q, N.ptr(it), it.shape[1], N.ptr(ln), n_items, dim,
                                       N.ptr(th), int(bucket_size), code, n_focus, thr, N.ptr(keep),
                                       N.current_stream()), "fps_lemp_keep_mask")
    return keep.bool()


def lemp_topk_filtered(queries, items_sorted, item_len, k: int, strategy, bucket_size: int, *, item_ids=None,
                       stats=None):
    """The reference's bucketed LEMP retrieval with its candidate filters executed on the device
    (``PSTopKGeneratorWorker.scala:49-111``; one warp per query, CUDA cores).  Exact like every other path;
    ``stats`` (uint64[4], accumulated): candidates seen, pruned, scored, buckets skipped - the numbers the
    host engine's ``lemp_top_k(..., stats=)`` counts.  The parity path: ``lemp_topk_packed`` is the fast one."""
    import torch
    code, n_focus, thr = _strategy_code(strategy)
    q = _pad4(queries.float())
    it = _pad4(items_sorted.float())
    n_q, n_items, dim = q.shape[0], it.shape[0], queries.shape[1]
    ln = item_len.float().contiguous()
    part_s = torch.empty((n_q, k), device="cuda", dtype=torch.float32)
    part_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
    out_s = torch.empty((n_q, k), device="cuda", dtype=torch.float32)
    out_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
    ids = None if item_ids is None else item_ids.int().contiguous()
    with nvtx_range("fps.topk.filtered"):
        N.check(N.lib().fps_lemp_topk_filtered(N.ptr(q), q.shape[1], n_q, N.ptr(it), it.shape[1], N.ptr(ln), N.ptr(ids),
                                               n_items, dim, int(k), int(bucket_size), code, n_focus, thr,
                                               N.ptr(part_s), N.ptr(part_i), N.ptr(out_s), N.ptr(out_i),
                                               N.ptr(stats), N.current_stream()), "fps_lemp_topk_filtered")
    return out_s, out_i


def topk_merge(part_scores, part_ids, k_out: int, seen=None):
    """``part_scores`` / ``part_ids``: ``[n_q, n_parts, k_in]``.  ``seen``: ``[n_q, m]`` int32 item ids
    to drop (-1 padded)."""
    import torch
    n_q, n_parts, k_in = part_scores.shape
    ps = part_scores.float().contiguous()
    pi = part_ids.int().contiguous()
    out_s = torch.empty((n_q, k_out), device="cuda", dtype=torch.float32)
    out_i = torch.empty((n_q, k_out), device="cuda", dtype=torch.int32)
    seen_stride = 0
    if seen is not None:
        seen = seen.int().contiguous()
        seen_stride = seen.shape[1]
    N.check(N.lib().fps_topk_merge(N.ptr(ps), N.ptr(pi), n_parts, n_q, k_in, k_out, N.ptr(seen), seen_stride,
                                   N.ptr(out_s), N.ptr(out_i), N.current_stream()), "fps_topk_merge")
    return out_s, out_i


class InPlaceScorer:
    """Top-``k`` against a worker-resident table scored in place through the TMA + tcgen05 pipeline
    (``fps_lemp_topk_wide``): every call re-packs the table (one pass over it) into a persistent image whose
    extra column carries the live mask (0 for a live slot, -1e30 for a dead one; the queries get a 1 there), so
    the scoring kernel needs no mask logic and live slots score exactly their inner product.  Rank < 128.
    Replaces the in-kernel-staging kernel on this path: 256 queries x 19K items, rank 10: 0.2 ms -> see
    ``profiles/README.md`` R2.5."""

    def __init__(self, n_items: int, dim: int):
        import torch
        if dim >= 128:
            raise ValueError("InPlaceScorer needs rank < 128 (one more column carries the live mask)")
        self.n_items, self.dim = int(n_items), int(dim)
        L = N.lib()
        self.packed = torch.empty(int(L.fps_topk_wide_packed_bytes(self.n_items, self.dim + 1)), device="cuda",
                                  dtype=torch.uint8)
        self.scratch = None

    def topk(self, queries, table_rows, k: int, live_ids):
        import torch
        L = N.lib()
        n_q = queries.shape[0]
        q = torch.zeros((n_q, ((self.dim + 1 + 3) // 4) * 4), device="cuda", dtype=torch.float32)
        q[:, : self.dim] = queries[:, : self.dim]
        q[:, self.dim] = 1.0
        N.check(L.fps_topk_pack_items_wide_live(N.ptr(table_rows), table_rows.shape[1], self.n_items, self.dim,
                                                N.ptr(live_ids), N.ptr(self.packed), N.current_stream()),
                "fps_topk_pack_items_wide_live")
        need = int(L.fps_topk_wide_scratch_bytes(n_q, k, self.n_items))
        if self.scratch is None or self.scratch.numel() < need:
            self.scratch = torch.empty(need, device="cuda", dtype=torch.uint8)
        out_s = torch.empty((n_q, k), device="cuda", dtype=torch.float32)
        out_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
        with nvtx_range("fps.topk.in_place"):
            N.check(L.fps_lemp_topk_wide(N.ptr(q), q.shape[1], n_q, N.ptr(self.packed), None, N.ptr(live_ids),
                                         self.n_items, self.dim + 1, k, 0, N.ptr(self.scratch), self.scratch.numel(),
                                         N.ptr(out_s), N.ptr(out_i), None, N.current_stream()), "fps_lemp_topk_wide (in place)")
        dead = out_i < 0
        return torch.where(dead, torch.full_like(out_s, -float("inf")), out_s), torch.where(dead, torch.full_like(out_i, -1), out_i)


def lemp_topk_table(queries, table_rows, dim: int, k: int, live_ids, *, scratch=None):
    """Top-``k`` of every query against a worker-resident table scored *in place*: ``table_rows`` is the
    shard's ``[rows, stride]`` tensor, ``live_ids[slot]`` the id of the item living in that slot or a
    negative number for a slot that is not a candidate (masked inside the kernel).  No length order,
    hence no LENGTH pruning: for catalogues small enough that a dump + sort + pack per model refresh
    would cost more than scoring every tile."""
    import torch
    q = _pad4(queries.float())
    n_q, n_items = q.shape[0], table_rows.shape[0]
    need = scratch_bytes(n_q, k, n_items, full=True)
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, device="cuda", dtype=torch.uint8)
    out_s = torch.empty((n_q, k), device="cuda", dtype=torch.float32)
    out_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
    N.check(N.lib().fps_lemp_topk(N.ptr(q), q.shape[1], n_q, N.ptr(table_rows), table_rows.shape[1], None,
                                  N.ptr(live_ids), n_items, dim, k, 2, N.ptr(scratch), scratch.numel(),
                                  N.ptr(out_s), N.ptr(out_i), None, N.current_stream()), "fps_lemp_topk (in place)")
    return out_s, out_i, scratch
