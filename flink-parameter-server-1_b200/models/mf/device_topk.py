"""Top-K recommendation on the device engine: the inference-only generator and the prequential
recommend-then-learn job.

Parity:
  * ``DeviceTopK``               - ``PSTopKGenerator.scala:47-107`` + ``PSTopKGeneratorWorker.scala:13-120``
  * ``DeviceLearnerGenerator``   - ``PSOnlineMatrixFactorizationAndTopKGenerator.scala:50-100`` +
    ``PSOnlineMatrixFactorizationAndTopKGeneratorWorker.scala:28-195``
  * merge                        - ``CollectTopKFromEachWorker.scala:23-74``

Roles are the reference's: **items live on the workers** (rank ``item % W`` owns an item, a
worker-resident table), **users live on the parameter server** (HBM table sharded over all GPUs).
A query batch is seen by every rank (X7 broadcast); each rank pulls the user rows P2P, scores them
against its local item tile with the tcgen05 kernel (LEMP bucket order + length bound), the
partial top-``worker_k`` lists are all-gathered (X8) and merged with the seen filter.
Learning re-uses the fused MF kernel with the tables swapped: the pulled row is the *user* row
(remote), the in-place update hits the *local item* row, the push is the user delta.
"""
from __future__ import annotations

import ctypes as C
import os
from collections import deque
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

from ...ops import _native as N
from ...ops import philox
from ...ops.table import ShardedTable, new_flow, read_flow
from ...ops.topk import (InPlaceScorer, PackedItems, lemp_topk, lemp_topk_filtered, lemp_topk_packed, lemp_topk_table, sort_items_by_length,
                         topk_merge)
from ...utils.observability import nvtx_range


class _LocalItems:
    """Length-sorted snapshot of this rank's live items (the GPU stand-in for the TreeSet of
    ``PSTopKGeneratorWorker.scala:19``; re-sorted when the model changed)."""

    def __init__(self):
        self.rows = self.lens = self.ids = None

    def refresh(self, ids, rows) -> None:
        self.packed = None
        if ids.numel() == 0:
            self.rows = self.lens = self.ids = None
            return
        self.rows, self.lens, self.ids = sort_items_by_length(rows.contiguous(), ids.int())
        if self.rows.shape[1] <= 128:
            # one-time split + swizzle so that every query batch streams the tiles with TMA
            self.packed = PackedItems(self.rows, self.lens, self.ids)

    def topk(self, q, k: int, use_pruning: bool, filters=None):
        if filters is not None:          # (strategy, bucket_size): the reference's loop with its candidate filters
            return lemp_topk_filtered(q, self.rows, self.lens, k, filters[0], filters[1], item_ids=self.ids)
        if getattr(self, "packed", None) is not None:
            return lemp_topk_packed(q, self.packed, k, use_pruning=use_pruning)
        return lemp_topk(q, self.rows, k, item_len=self.lens, item_ids=self.ids, use_pruning=use_pruning)

    @property
    def empty(self) -> bool:
        return self.rows is None


def _all_gather_parts(ctx, scores, ids):
    """X8: partial lists of every worker -> ``[n_q, W, k]`` on every rank."""
    import torch
    if not ctx.distributed:
        return scores[:, None, :], ids[:, None, :]
    import torch.distributed as dist
    W = ctx.world_size
    s_all = [torch.empty_like(scores) for _ in range(W)]
    i_all = [torch.empty_like(ids) for _ in range(W)]
    dist.all_gather(s_all, scores.contiguous())
    dist.all_gather(i_all, ids.contiguous())
    return torch.stack(s_all, dim=1), torch.stack(i_all, dim=1)


class _SeenMemory:
    """Per-user FIFO of consumed items (host side; ``CollectTopKFromEachWorker.scala:50-55``)."""

    def __init__(self, memory: int):
        self.memory = memory
        self.sets: Dict[int, set] = {}
        self.lists: Dict[int, deque] = {}

    def add(self, user: int, item: int) -> None:
        s = self.sets.setdefault(user, set())
        s.add(item)
        lst = self.lists.setdefault(user, deque())
        lst.append(item)
        if self.memory > -1 and len(lst) > self.memory:
            s.discard(lst.popleft())

    def matrix(self, users: List[int]):
        import torch
        width = max([len(self.sets.get(u, ())) for u in users] + [1])
        m = np.full((len(users), width), -1, dtype=np.int32)
        for r, u in enumerate(users):
            s = self.sets.get(u)
            if s:
                m[r, : len(s)] = list(s)
        return torch.from_numpy(m).cuda()


class DeviceTopK:
    def __init__(self, n_users: int, n_items: int, num_factors: int, *, K: int = 100, worker_k: int = 75,
                 user_memory: int = 0, use_pruning: bool = True, ctx=None, candidate_filters=None,
                 bucket_size: int = 100):
        """``candidate_filters``: a ``LEMPPruningStrategy`` to run the reference's bucket loop with its
        per-candidate filters on CUDA cores (``csrc/lemp_filters.cu``) instead of the tensor-core tile
        kernel with the bucket-level LENGTH bound - same results, the reference's pruning behaviour."""
        from ...parallel.device_context import get_context
        self.ctx = ctx or get_context()
        self.K, self.worker_k, self.use_pruning = K, min(worker_k, 128), use_pruning
        self.filters = None if candidate_filters is None else (candidate_filters, int(bucket_size))
        self.dim = num_factors
        # PS side: users; zero rows = unknown user (the reference's invalid param (-1, []))
        self.users = ShardedTable(n_users, num_factors, partition="hash", init_range=(0.0, 0.0), ctx=self.ctx)
        self.items = _LocalItems()
        self.seen = _SeenMemory(user_memory)
        self._item_ids: List[int] = []
        self._item_rows: List[np.ndarray] = []

    def load_model(self, model: Iterable) -> None:
        """``Left((user, (len, vec)))`` -> PS table; ``Right((item, (len, vec)))`` -> the owning
        worker ``item % W`` (``transform_with_double_model_load`` with key placement)."""
        import torch
        W, rank = self.ctx.world_size, self.ctx.rank
        us, uv = [], []
        for e in model:
            k, (_len, vec) = e.value
            if e.is_left:
                us.append(k)
                uv.append(np.asarray(vec, dtype=np.float32))
            elif k % W == rank:
                self._item_ids.append(k)
                self._item_rows.append(np.asarray(vec, dtype=np.float32))
        if us and rank == 0:            # one loader is enough: pushes are routed to the owners
            self.users.load(torch.tensor(us, dtype=torch.int64), torch.from_numpy(np.stack(uv)))
        torch.cuda.synchronize()
        self.ctx.barrier()
        if self._item_ids:
            self.items.refresh(torch.tensor(self._item_ids, dtype=torch.int32).cuda(),
                               torch.from_numpy(np.stack(self._item_rows)).cuda())

    def local_topk(self, q):
        import torch
        n_q = q.shape[0]
        if self.items.empty:
            return (torch.full((n_q, self.worker_k), -float("inf"), device="cuda"),
                    torch.full((n_q, self.worker_k), -1, device="cuda", dtype=torch.int32))
        return self.items.topk(q, min(self.worker_k, 128), self.use_pruning, self.filters)

    def query(self, user_ids: List[int], apply_seen: bool = True):
        """Every rank calls this with the same ``user_ids``.  Returns ``(scores, item_ids)``
        ``[n_q, K]`` (unknown users: all ``-1``)."""
        import torch
        ids = torch.tensor(user_ids, dtype=torch.int64)
        q = self.users.gather(ids)[:, : self.dim].contiguous()
        valid = torch.linalg.vector_norm(q, dim=1) > 0
        ls, li = self.local_topk(q)
        ps, pi = _all_gather_parts(self.ctx, ls, li)
        seen = self.seen.matrix(user_ids) if apply_seen else None
        k_out = min(self.K, ps.shape[1] * ps.shape[2])
        out_s, out_i = topk_merge(ps, pi, k_out, seen)
        out_i = torch.where(valid[:, None] & torch.isfinite(out_s), out_i, torch.full_like(out_i, -1))
        return out_s, out_i

    def recommend(self, ratings) -> list:
        """``ps_top_k_generator`` output: ``[(item, timestamp, [(score, item), ...]), ...]``."""
        users = [r.user for r in ratings]
        s, i = self.query(users)
        s, i = s.cpu().numpy(), i.cpu().numpy()
        out = []
        for row, r in enumerate(ratings):
            topk = [(float(a), int(b)) for a, b in zip(s[row], i[row]) if b >= 0]
            out.append((r.item, r.timestamp, topk))
            self.seen.add(r.user, r.item)
        return out


class PrequentialOutput:
    """The output of the prequential job kept as arrays (``scores`` / ``ids`` ``[n, K]``, ``-1`` = no
    item) and turned into the reference's records ``(user, item, timestamp, [(score, item), ...])``
    only when somebody indexes or iterates it - building 100-entry Python lists for every event was
    what bound the first device version to 30 K events/s."""

    def __init__(self, users, items, timestamps, scores, ids):
        self.users, self.items, self.timestamps, self.scores, self.ids = users, items, timestamps, scores, ids

    def __len__(self) -> int:
        return len(self.users)

    def _one(self, r: int):
        keep = self.ids[r] >= 0
        return (int(self.users[r]), int(self.items[r]), self.timestamps[r],
                list(zip(self.scores[r][keep].tolist(), self.ids[r][keep].tolist())))

    def __getitem__(self, key):
        if isinstance(key, slice):
            return [self._one(r) for r in range(*key.indices(len(self)))]
        return self._one(key if key >= 0 else len(self) + key)

    def __iter__(self):
        return (self._one(r) for r in range(len(self)))

    def ranks(self) -> np.ndarray:
        """1-based rank of the consumed item in its top-K list, 0 = not recommended (``nDCGSink``)."""
        hit = self.ids == self.items[:, None]
        return np.where(hit.any(axis=1), hit.argmax(axis=1) + 1, 0)

    def ndcg(self, start: int = 0) -> float:
        r = self.ranks()[start:]
        return float(np.where(r > 0, np.log(2.0) / np.log(1.0 + np.maximum(r, 1)), 0.0).mean()) if len(r) else 0.0

    def hit_rate(self, start: int = 0) -> float:
        r = self.ranks()[start:]
        return float((r > 0).mean()) if len(r) else 0.0


class DeviceLearnerGenerator:
    """Prequential recommend-then-learn (N8) as a fixed sequence of kernels per micro-batch with no
    host round trip in between: the whole batch is answered with the model as of the batch start
    (asynchronous semantics: reads may be stale by up to one batch), then learnt.

      recommend   pull of the batch's user rows (P2P gather) -> tcgen05 scoring of this worker's item
                  table *in place* (live-item mask; catalogues up to ``dense_catalog`` rows - beyond that
                  the length-sorted, packed snapshot with LEMP pruning, refreshed every
                  ``refresh_every`` batches) -> all-gather of the partial lists (X8) -> merge with
                  the per-user seen filter read from the device ring
      learn       only for events whose item this worker owns (``:128``): the item becomes live, its
                  negatives are drawn by the device sampler from this worker's live items and
                  rejected against the user's ring (``:143-156``), then one fused SGD launch with the
                  tables swapped - the pulled / pushed row is the *user's* (PS side), the row updated
                  in place is the local *item's* (``:150-163``)

    The user's recent-items memory is one device ring per user id on every rank (all ranks see all
    events), shared by the merge filter and the sampler; memories longer than ``max_ring`` are cut."""

    def __init__(self, n_users: int, n_items: int, num_factors: int = 10, range_min: float = -0.001,
                 range_max: float = 0.001, learning_rate: float = 0.01, negative_sample_rate: int = 0,
                 user_memory: int = 65535, K: int = 100, worker_k: int = 75, pull_limit: int = 500,
                 seed: int = 0, use_pruning: bool = True, ctx=None, dense_catalog: int = 1 << 18,
                 refresh_every: int = 8, max_ring: int = 512, candidate_filters=None, bucket_size: int = 100,
                 cuda_graph: bool = False):
        """``candidate_filters`` (a ``LEMPPruningStrategy``): score with the reference's bucket loop and the
        strategy's per-candidate filters (``ops.topk.lemp_topk_filtered``) on a length-sorted snapshot refreshed
        every batch, instead of the tensor-core kernels.

        ``cuda_graph=True`` (one GPU, item table scored in place): the ~25 launches of a step are captured once
        per batch size and replayed; the step tag of the negative sampler lives on the device for that reason.
        Off by default: measured on config 7 the step is bound by its kernels and by the Python loop around it,
        not by launches (batch 64: 0.553 -> 0.532 ms per step), and at large batches the copies into and out of
        the static buffers cost more than the launches saved."""
        import torch
        from ...parallel.device_context import get_context
        self.ctx = ctx or get_context()
        W = self.ctx.world_size
        self.n_users, self.n_items = n_users, n_items
        self.K, self.worker_k, self.dim = K, min(worker_k, 128), num_factors
        self.negative_sample_rate, self.use_pruning = negative_sample_rate, use_pruning
        self.users = ShardedTable(n_users, num_factors, partition="hash", init_range=(range_min, range_max),
                                  seed=seed, stream_tag=philox.STREAM_USER, ctx=self.ctx)
        # worker-resident items: every row holds init(id) from the start, but an item only becomes a
        # candidate (live) with its first rating
        self.items = ShardedTable(n_items, num_factors, partition="hash", init_range=(range_min, range_max),
                                  seed=seed, stream_tag=philox.STREAM_ITEM, lazy=False, ctx=self.ctx,
                                  local_only=True)
        self._shard = self.ctx.rank if self.ctx.distributed else 0
        if not self.ctx.distributed and self.items.n_shards != 1:
            raise ValueError("one local item shard per process")
        self.live_ids = torch.full((self.items.rows_per_shard,), -1, device="cuda", dtype=torch.int32)
        self.params = N.FpsMfParams()
        self.params.lr, self.params.lambda_ = learning_rate, 0.0
        self.params.pull_limit = int(pull_limit)
        self.params.local_shard = self.ctx.rank if self.ctx.distributed else -1
        self.flow = new_flow()
        # per-user ring on every rank + this worker's pool of live items
        self.mem = max(1, min(int(user_memory) if user_memory > 0 else max_ring, max_ring))
        self._seen = torch.full((n_users * self.mem,), -1, device="cuda", dtype=torch.int32)
        self._seen_count = torch.zeros(n_users, device="cuda", dtype=torch.int32)
        self._bitmap = torch.zeros((n_items + 31) // 32, device="cuda", dtype=torch.int32)
        self._pool = torch.zeros(n_items, device="cuda", dtype=torch.int32)
        self._pool_count = torch.zeros(1, device="cuda", dtype=torch.int32)
        st = N.FpsNegSampler()
        st.seen, st.seen_count = self._seen.data_ptr(), self._seen_count.data_ptr()
        st.item_bitmap, st.item_pool, st.pool_count = self._bitmap.data_ptr(), self._pool.data_ptr(), self._pool_count.data_ptr()
        st.n_items, st.seed = n_items, (seed * 0x9E3779B97F4A7C15 + 977 * (self.ctx.rank + 1)) & (2 ** 64 - 1)
        st.user_memory, st.rate, st.local_shard, st.n_rings = self.mem, int(negative_sample_rate), -2, n_users
        self._sampler = st
        self._tag = 0
        self._tag_dev = torch.zeros(1, device="cuda", dtype=torch.int64)
        self._graphs = {}             # batch size -> (graph, static inputs, static outputs)
        self._eager_steps = {}        # batch size -> eager steps so far (a graph is captured after two)
        self._want_graph = cuda_graph
        self.filters = None if candidate_filters is None else (candidate_filters, int(bucket_size))
        self.dense = self.filters is None and self.items.rows_per_shard <= dense_catalog and num_factors <= 128
        if self.filters is not None:
            refresh_every = 1          # the filters need exact lengths: no stale snapshot
        self.refresh_every = max(1, int(refresh_every))
        self.local = _LocalItems()
        self._since_refresh = 1 << 30
        self._scratch = None
        self._scorer = None
        self._skip = torch.tensor(N.SKIP_ID, device="cuda", dtype=torch.int32)
        self.batches = 0

    # -- recommend ---------------------------------------------------------------------------------
    def _local_topk(self, q):
        import torch
        n_q = q.shape[0]
        if self.dense:
            rows = self.items.rows[self._shard]
            # the TMA + tcgen05 pipeline over a per-step packed image pays from ~1K queries on (measured, rank 10,
            # 19K items, ms per step packed / in-kernel staging: 64 queries 0.64 / 0.35, 256: 0.68 / 0.50,
            # 2048: 0.93 / 1.06, 8192: 1.88 / 3.08); FPS_TOPK_INPLACE_PACKED=0 / 1 forces one of them
            force = os.environ.get("FPS_TOPK_INPLACE_PACKED")
            if self.dim < 128 and (force == "1" or (force != "0" and n_q >= 1024)):
                if self._scorer is None:
                    self._scorer = InPlaceScorer(rows.shape[0], self.dim)
                return self._scorer.topk(q, rows, self.worker_k, self.live_ids)
            ls, li, self._scratch = lemp_topk_table(q, rows, self.dim, self.worker_k, self.live_ids, scratch=self._scratch)
            return ls, li
        # large catalogue: length-sorted packed snapshot, refreshed every few batches
        if self._since_refresh >= self.refresh_every:
            live = torch.nonzero(self.live_ids >= 0).flatten()
            self.local.refresh(self.live_ids[live], self.items.rows[self._shard][live][:, : self.dim])
            self._since_refresh = 0
        self._since_refresh += 1
        if self.local.empty:
            return (torch.full((n_q, self.worker_k), -float("inf"), device="cuda"),
                    torch.full((n_q, self.worker_k), -1, device="cuda", dtype=torch.int32))
        return self.local.topk(q, self.worker_k, self.use_pruning, self.filters)

    def step_tensors(self, users, items, ratings):
        """``users`` / ``items`` int32, ``ratings`` fp32 device tensors - the same batch on every rank.
        Returns ``(scores, ids)`` ``[n, K]`` device tensors; nothing here synchronises with the host."""
        import torch
        n = int(users.numel())
        if not self._want_graph or self.ctx.world_size != 1 or not self.dense or n == 0:
            return self._step_impl(users, items, ratings)
        entry = self._graphs.get(n)
        if entry is None:
            done = self._eager_steps.get(n, 0)
            if done < 2:              # the first steps of a batch size run eagerly (scratch buffers, lazy CUDA state)
                self._eager_steps[n] = done + 1
                return self._step_impl(users, items, ratings)
            su, si, sr = users.clone(), items.clone(), ratings.clone()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out = self._step_impl(su, si, sr)
            self.batches -= 1         # the capture ran the host side of a step, not the step
            entry = self._graphs[n] = (g, (su, si, sr), out)
        g, (su, si, sr), out = entry
        su.copy_(users)
        si.copy_(items)
        sr.copy_(ratings)
        g.replay()
        self.batches += 1
        return out[0].clone(), out[1].clone()

    def _step_impl(self, users, items, ratings):
        import torch
        W, rank = self.ctx.world_size, self.ctx.rank
        n = users.numel()
        with nvtx_range("fps.learner_generator.step"):
            uid = users.long()
            q = self.users.gather(uid)[:, : self.dim].contiguous()
            ls, li = self._local_topk(q)
            ps, pi = _all_gather_parts(self.ctx, ls, li)
            seen = self._seen.view(self.n_users, self.mem)[uid]
            out_s, out_i = topk_merge(ps, pi, min(self.K, ps.shape[1] * ps.shape[2]), seen)
            # ---- learn (item owner only); the rings take every event
            owned = (items % W) == rank if W > 1 else None
            u_own = users if owned is None else torch.where(owned, users, self._skip)
            if owned is not None:
                u_other = torch.where(owned, torch.full_like(users, -1), users)
                N.check(N.lib().fps_seen_ring_append(N.ptr(self._seen), N.ptr(self._seen_count), self.mem, self.n_users,
                                                     N.ptr(u_other), N.ptr(items), n, N.current_stream()),
                        "fps_seen_ring_append")
            N.check(N.lib().fps_mark_live_items(self.items.view_ref(), self.params.local_shard if W > 1 else 0,
                                                N.ptr(u_own), N.ptr(items), n, N.ptr(self.live_ids),
                                                N.current_stream()), "fps_mark_live_items")
            m = n * (self.negative_sample_rate + 1)
            ou = torch.empty(m, device="cuda", dtype=torch.int32)
            oi = torch.empty(m, device="cuda", dtype=torch.int32)
            orr = torch.empty(m, device="cuda", dtype=torch.float32)
            self._tag_dev += 1          # on the stream: a replayed graph draws with a new tag every time
            N.check(N.lib().fps_mf_sample_negatives_dtag(self.users.view_ref(), C.byref(self._sampler), N.ptr(u_own),
                                                         N.ptr(items), N.ptr(ratings), n, N.ptr(self._tag_dev),
                                                         N.ptr(ou), N.ptr(oi), N.ptr(orr), N.current_stream()),
                    "fps_mf_sample_negatives")
            # tables swapped: the kernel's "item" operand = PS-side users (pulled / pushed remotely), its
            # "user" operand = worker-resident items (updated in place) - so the skip marker goes with the items
            skipped = ou == self._skip
            wi = torch.where(skipped, self._skip, oi)
            wu = torch.where(skipped, torch.zeros_like(ou), ou)
            N.check(N.lib().fps_mf_sgd_step(self.users.view_ref(), self.items.view_ref(), N.ptr(wi), N.ptr(wu),
                                            N.ptr(orr), m, C.byref(self.params), N.ptr(self.flow),
                                            N.current_stream()), "fps_mf_sgd_step (learner)")
        self.batches += 1
        return out_s, out_i

    def step(self, ratings) -> list:
        """``ratings``: list of ``Rating`` (same list on every rank).  Returns
        ``[(user, item, timestamp, [(score, item), ...]), ...]`` for the batch."""
        import torch
        u = torch.tensor([r.user for r in ratings], dtype=torch.int32).cuda()
        i = torch.tensor([r.item for r in ratings], dtype=torch.int32).cuda()
        v = torch.tensor([r.rating for r in ratings], dtype=torch.float32).cuda()
        s, ids = self.step_tensors(u, i, v)
        out = PrequentialOutput(np.array([r.user for r in ratings]), np.array([r.item for r in ratings]),
                                [r.timestamp for r in ratings], s.cpu().numpy(), ids.cpu().numpy())
        return out[:]

    def stats(self) -> dict:
        return read_flow(self.flow)


def ps_top_k_generator_device(src, model, num_factors: int = 10, user_memory: int = 0, K: int = 100,
                              worker_k: int = 75, n_users: Optional[int] = None, use_pruning: bool = True,
                              candidate_filters=None, bucket_size: int = 100):
    """Device twin of ``ps_top_k_generator``."""
    model = list(model)
    ratings = list(src)
    n_users = n_users or (max([e.value[0] for e in model if e.is_left] + [r.user for r in ratings]) + 1)
    n_items = max(e.value[0] for e in model if e.is_right) + 1
    gen = DeviceTopK(n_users, n_items, num_factors, K=K, worker_k=worker_k, user_memory=user_memory,
                     use_pruning=use_pruning, candidate_filters=candidate_filters, bucket_size=bucket_size)
    gen.load_model(model)
    return gen.recommend(ratings)


def ps_online_learner_and_generator_device(src, n_users: int, n_items: int, batch_size: int = 256, **kw):
    """Device twin of ``ps_online_learner_and_generator`` (mini-batched prequential loop).  Returns
    ``(PrequentialOutput, DeviceLearnerGenerator)``; the output behaves like the reference's list of
    ``(user, item, timestamp, topK)`` records and also offers vectorised ``ndcg()`` / ``hit_rate()``.
    ``src``: an iterable of ``Rating`` or, columnar, a tuple ``(users, items, ratings[, timestamps])`` of arrays."""
    import torch
    lg = DeviceLearnerGenerator(n_users, n_items, **kw)
    if isinstance(src, tuple) and len(src) in (3, 4) and not hasattr(src[0], "user"):
        # columnar input ``(users, items, ratings[, timestamps])``: no per-record Python objects (unpacking 69K
        # ``Rating`` objects costs 13 ms - half of the whole config-7 job at 2.6 M events/s)
        u = np.ascontiguousarray(np.asarray(src[0]), dtype=np.int32)
        i = np.ascontiguousarray(np.asarray(src[1]), dtype=np.int32)
        v = np.ascontiguousarray(np.asarray(src[2]), dtype=np.float32)
        n = len(u)
        ts = list(np.asarray(src[3]).tolist()) if len(src) == 4 else list(range(n))
    else:
        ratings = src if isinstance(src, list) else list(src)
        n = len(ratings)
        u = np.fromiter((r.user for r in ratings), dtype=np.int32, count=n)
        i = np.fromiter((r.item for r in ratings), dtype=np.int32, count=n)
        v = np.fromiter((r.rating for r in ratings), dtype=np.float32, count=n)
        ts = [r.timestamp for r in ratings]
    du, di, dv = (torch.from_numpy(x).cuda() for x in (u, i, v))
    parts = []
    for s in range(0, n, batch_size):
        parts.append(lg.step_tensors(du[s:s + batch_size], di[s:s + batch_size], dv[s:s + batch_size]))
    if parts:
        scores = torch.cat([p[0] for p in parts]).cpu().numpy()
        ids = torch.cat([p[1] for p in parts]).cpu().numpy()
    else:
        scores, ids = np.zeros((0, lg.K), np.float32), np.zeros((0, lg.K), np.int32)
    torch.cuda.synchronize()
    lg.ctx.barrier()
    return PrequentialOutput(u, i, ts, scores, ids), lg
