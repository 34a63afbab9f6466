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
from collections import deque
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

from ...ops import _native as N
from ...ops import philox
from ...ops.table import ShardedTable, new_flow, read_flow
from ...ops.topk import PackedItems, lemp_topk, lemp_topk_packed, sort_items_by_length, topk_merge


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

    def topk(self, q, k: int, use_pruning: bool):
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
                 user_memory: int = 0, use_pruning: bool = True, ctx=None):
        from ...parallel.device_context import get_context
        self.ctx = ctx or get_context()
        self.K, self.worker_k, self.use_pruning = K, min(worker_k, 128), use_pruning
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
        return self.items.topk(q, min(self.worker_k, 128), self.use_pruning)

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


class DeviceLearnerGenerator:
    """Prequential recommend-then-learn in mini-batches: the whole batch is answered with the
    model as of the batch start (asynchronous semantics: reads may be stale by up to one batch),
    then learnt."""

    def __init__(self, n_users: int, n_items: int, num_factors: int = 10, range_min: float = -0.001,
                 range_max: float = 0.001, learning_rate: float = 0.01, negative_sample_rate: int = 0,
                 user_memory: int = 65535, K: int = 100, worker_k: int = 75, pull_limit: int = 500,
                 seed: int = 0, use_pruning: bool = True, ctx=None):
        import torch
        from ...parallel.device_context import get_context
        self.ctx = ctx or get_context()
        W = self.ctx.world_size
        self.K, self.worker_k, self.dim = K, min(worker_k, 128), num_factors
        self.negative_sample_rate, self.use_pruning = negative_sample_rate, use_pruning
        self.users = ShardedTable(n_users, num_factors, partition="hash", init_range=(range_min, range_max),
                                  seed=seed, stream_tag=philox.STREAM_USER, ctx=self.ctx)
        # worker-resident items, lazily created by their first rating (only rated items are candidates)
        self.items = ShardedTable(n_items, num_factors, partition="hash", init_range=(range_min, range_max),
                                  seed=seed, stream_tag=philox.STREAM_ITEM, lazy=True, ctx=self.ctx,
                                  local_only=True)
        self.params = N.FpsMfParams()
        self.params.lr, self.params.lambda_ = learning_rate, 0.0
        self.params.pull_limit = int(pull_limit)
        self.params.local_shard = self.ctx.rank if self.ctx.distributed else -1
        self.flow = new_flow()
        self.local = _LocalItems()
        self.seen = _SeenMemory(user_memory)
        self._np_rng = np.random.default_rng(seed * 977 + self.ctx.rank)
        self._dirty = False

    def _refresh_items(self) -> None:
        if self._dirty or self.local.empty:
            ids, rows = self.items.dump()
            self.local.refresh(ids, rows[:, : self.dim])
            self._dirty = False

    def step(self, ratings) -> list:
        """``ratings``: list of ``Rating`` (same list on every rank).  Returns
        ``[(user, item, timestamp, [(score, item), ...]), ...]`` for the batch."""
        import torch
        W, rank = self.ctx.world_size, self.ctx.rank
        users = [r.user for r in ratings]
        # ---- recommend with the current model
        self._refresh_items()
        uid = torch.tensor(users, dtype=torch.int64)
        q = self.users.gather(uid)[:, : self.dim].contiguous()
        n_q = len(users)
        if self.local.empty:
            ls = torch.full((n_q, self.worker_k), -float("inf"), device="cuda")
            li = torch.full((n_q, self.worker_k), -1, device="cuda", dtype=torch.int32)
        else:
            ls, li = self.local.topk(q, self.worker_k, self.use_pruning)
        ps, pi = _all_gather_parts(self.ctx, ls, li)
        out_s, out_i = topk_merge(ps, pi, min(self.K, ps.shape[1] * ps.shape[2]), self.seen.matrix(users))
        s_np, i_np = out_s.cpu().numpy(), out_i.cpu().numpy()
        out = []
        for row, r in enumerate(ratings):
            out.append((r.user, r.item, r.timestamp,
                        [(float(a), int(b)) for a, b in zip(s_np[row], i_np[row]) if b >= 0]))
            self.seen.add(r.user, r.item)
        # ---- learn: only the owner of the rated item (item % W == rank)
        mine = [r for r in ratings if r.item % W == rank]
        if mine:
            u = torch.tensor([r.user for r in mine], dtype=torch.int32)
            it = torch.tensor([r.item for r in mine], dtype=torch.int32)
            rt = torch.tensor([r.rating for r in mine], dtype=torch.float32)
            if self.negative_sample_rate > 0 and not self.local.empty:
                # negatives are drawn from this worker's live items (rating 0) and retried against the
                # user's seen set like the reference (":145-149", which retries up to 32 times; here 8)
                # (drawn in one vectorised call: 8 candidates per wanted negative, first unseen one wins;
                # the per-draw retry loop was the slowest part of a step)
                pool = self.local.ids.cpu().numpy()
                neg_u, neg_i = [], []
                rate = self.negative_sample_rate
                cand = pool[self._np_rng.integers(0, len(pool), size=(len(mine), rate, 8))]
                for row, r in enumerate(mine):
                    seen = self.seen.sets.get(r.user, ())
                    for j in range(min(rate, max(len(pool) - len(seen), 0))):
                        for c in cand[row, j]:
                            if int(c) not in seen:
                                neg_u.append(r.user)
                                neg_i.append(int(c))
                                break
                if neg_u:
                    u = torch.cat([torch.tensor(neg_u, dtype=torch.int32), u])
                    it = torch.cat([torch.tensor(neg_i, dtype=torch.int32), it])
                    rt = torch.cat([torch.zeros(len(neg_u)), rt])
            u, it, rt = u.cuda(), it.cuda(), rt.cuda()
            # tables swapped: "items" operand = PS-side users (pulled / pushed remotely),
            # "users" operand = worker-resident items (updated in place)
            N.check(N.lib().fps_mf_sgd_step(self.users.view_ref(), self.items.view_ref(), N.ptr(it), N.ptr(u),
                                            N.ptr(rt), u.numel(), C.byref(self.params), N.ptr(self.flow),
                                            N.current_stream()), "fps_mf_sgd_step (learner)")
            self._dirty = True
        torch.cuda.synchronize()
        self.ctx.barrier()
        return out

    def stats(self) -> dict:
        return read_flow(self.flow)


def ps_top_k_generator_device(src, model, num_factors: int = 10, user_memory: int = 0, K: int = 100,
                              worker_k: int = 75, n_users: Optional[int] = None, use_pruning: bool = True):
    """Device twin of ``ps_top_k_generator``."""
    model = list(model)
    ratings = list(src)
    n_users = n_users or (max([e.value[0] for e in model if e.is_left] + [r.user for r in ratings]) + 1)
    n_items = max(e.value[0] for e in model if e.is_right) + 1
    gen = DeviceTopK(n_users, n_items, num_factors, K=K, worker_k=worker_k, user_memory=user_memory,
                     use_pruning=use_pruning)
    gen.load_model(model)
    return gen.recommend(ratings)


def ps_online_learner_and_generator_device(src, n_users: int, n_items: int, batch_size: int = 256, **kw):
    """Device twin of ``ps_online_learner_and_generator`` (mini-batched prequential loop)."""
    lg = DeviceLearnerGenerator(n_users, n_items, **kw)
    out, buf = [], []
    for r in src:
        buf.append(r)
        if len(buf) >= batch_size:
            out.extend(lg.step(buf))
            buf = []
    if buf:
        out.extend(lg.step(buf))
    return out, lg
