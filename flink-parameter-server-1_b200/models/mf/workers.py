"""Matrix-factorisation worker logics for the host engine (fp64).

Parity:
  * ``BaseMFWorkerLogic``                               - ``workers/BaseMFWorkerLogic.scala:8-188``
  * ``PSOfflineMatrixFactorizationWorker``              - ``workers/PSOfflineMatrixFactorizationWorker.scala:27-150``
  * ``PSOnlineMatrixFactorizationWorker``               - ``workers/PSOnlineMatrixFactorizationWorker.scala:22-90``
  * ``PSTopKGeneratorWorker``                           - ``workers/PSTopKGeneratorWorker.scala:13-120``
  * ``PSOnlineMatrixFactorizationAndTopKGeneratorWorker`` - ``workers/...AndTopKGeneratorWorker.scala:28-195``

Host-engine differences, all deliberate:
  * no second thread in the offline worker - the epoch loop runs inside the EOF handler and the
    pull limiter queues the excess pulls, which gives the same message order without the
    unsynchronised sharing of SURVEY Q13;
  * every RNG is seeded (negative sampling) or counter-based (initialisers);
  * ``shuffle=True`` really shuffles the blocks each epoch (the reference discards the shuffled
    copy, Q2); ``shuffle=False`` reproduces the reference order.
"""
from __future__ import annotations

import random
from collections import deque
from typing import Deque, Dict, List, Optional, Tuple

import numpy as np

from ...api import ParameterServerClient, RuntimeContext, WorkerLogic
from ...entities import EOF
from ...ops import philox
from .factors import (FactorInitializerDescriptor, FactorUpdater,
                      RangedRandomFactorInitializerDescriptor, SGDUpdater)
from .pruning import (LEMPPruningStrategy, LI, bucket_filter, focus_coordinate, focus_set)
from .rating import Rating, RichRating, TopKQueue
from .vector import attach_length, dot_product, vector_sum


class BaseMFWorkerLogic(WorkerLogic):
    """A worker logic with a worker-resident model half."""

    def __init__(self):
        self.model: Dict[int, object] = {}

    def update_model(self, param_id: int, param) -> None:
        raise NotImplementedError


class _SeenMemory:
    """Per-user FIFO of the most recent ``capacity`` items + membership set."""

    def __init__(self, capacity: int):
        self.capacity = capacity
        self.sets: Dict[int, set] = {}
        self.queues: Dict[int, Deque[int]] = {}

    def touch(self, user: int, item: int) -> set:
        s = self.sets.setdefault(user, set())
        q = self.queues.setdefault(user, deque())
        if len(q) >= self.capacity:
            s.discard(q.popleft())
        s.add(item)
        q.append(item)
        return s


class PSOfflineMatrixFactorizationWorker(WorkerLogic):
    """Multi-epoch MF.  User vectors live on the worker, item vectors on the PS.
    Input: ``Right(Rating)`` records followed by one ``Left(EOF)``."""

    def __init__(self, num_factors: int, range_min: float, range_max: float, learning_rate: float,
                 negative_sample_rate: int, user_memory: int, iterations: int, *,
                 seed: int = 0, shuffle: bool = True, lambda_: float = 0.0):
        self.factor_init_desc = RangedRandomFactorInitializerDescriptor(
            num_factors, range_min, range_max, seed, philox.STREAM_USER)
        self.factor_update = SGDUpdater(learning_rate, lambda_)
        self.negative_sample_rate = negative_sample_rate
        self.user_memory = user_memory
        self.iterations = iterations
        self.shuffle = shuffle
        self.seed = seed
        self.rbs: List[List[Rating]] = []
        self.user_vectors: Dict[int, np.ndarray] = {}
        self.rating_buffer: Dict[int, Deque[Tuple[int, float]]] = {}
        self.seen = _SeenMemory(user_memory)
        self.all_item_ids_set: set = set()
        self.all_item_ids: List[int] = []
        self.started = False
        self._rng = random.Random(seed)
        self._init = None
        self.updates = 0

    def open(self, runtime_context: RuntimeContext | None = None) -> None:
        idx = runtime_context.index_of_this_subtask if runtime_context else 0
        self._rng = random.Random((self.seed << 8) ^ idx)
        self._init = self.factor_init_desc.open()

    def on_recv(self, value, ps: ParameterServerClient) -> None:
        if value.is_right:
            rating: Rating = value.value
            if self.started:
                raise RuntimeError("Should not have started the epochs while waiting for further elements.")
            if rating.item not in self.all_item_ids_set:
                self.all_item_ids_set.add(rating.item)
                self.all_item_ids.append(rating.item)
            seen = self.seen.touch(rating.user, rating.item)
            rs: List[Rating] = []
            for _ in range(min(len(self.all_item_ids_set) - len(seen), self.negative_sample_rate)):
                neg = self._rng.choice(self.all_item_ids)
                while neg in seen:
                    neg = self._rng.choice(self.all_item_ids)
                rs.append(Rating.from_tuple(rating.user, neg, 0.0))
            rs.append(rating)
            self.rbs.append(rs)
        else:
            self.started = True
            for _ in range(self.iterations):
                blocks = list(self.rbs)
                if self.shuffle:
                    self._rng.shuffle(blocks)
                for rs in blocks:
                    for r in rs:
                        self.rating_buffer.setdefault(r.item, deque()).append((r.user, r.rating))
                        ps.pull(r.item)

    def on_pull_recv(self, item: int, item_vec, ps: ParameterServerClient) -> None:
        user, rating = self.rating_buffer[item].popleft()
        if self._init is None:
            self._init = self.factor_init_desc.open()
        user_vec = self.user_vectors.get(user)
        if user_vec is None:
            user_vec = self._init.next_factor(user)
        d_user, d_item = self.factor_update.delta(rating, user_vec, item_vec)
        new_user = vector_sum(user_vec, d_user)
        self.user_vectors[user] = new_user
        self.updates += 1
        ps.output((user, new_user))
        ps.push(item, d_item)


class PSOnlineMatrixFactorizationWorker(WorkerLogic):
    """Single-pass streaming MF with negative sampling from the items seen so far."""

    def __init__(self, num_factors: int, range_min: float, range_max: float, learning_rate: float,
                 user_memory: int, negative_sample_rate: int, *, seed: int = 0, lambda_: float = 0.0):
        self.factor_init_desc = RangedRandomFactorInitializerDescriptor(
            num_factors, range_min, range_max, seed, philox.STREAM_USER)
        self.factor_update = SGDUpdater(learning_rate, lambda_)
        self.user_memory = user_memory
        self.negative_sample_rate = negative_sample_rate
        self.seed = seed
        self.user_vectors: Dict[int, np.ndarray] = {}
        self.rating_buffer: Dict[int, Deque[Rating]] = {}
        self.item_ids: List[int] = []
        self.seen = _SeenMemory(user_memory)
        self._rng = random.Random(seed)
        self._init = None

    def open(self, runtime_context: RuntimeContext | None = None) -> None:
        idx = runtime_context.index_of_this_subtask if runtime_context else 0
        self._rng = random.Random((self.seed << 8) ^ idx)
        self._init = self.factor_init_desc.open()

    def on_recv(self, data: Rating, ps: ParameterServerClient) -> None:
        seen = self.seen.touch(data.user, data.item)
        # the current item only counts against the candidate pool if it is already in it
        pool = len(self.item_ids) - sum(1 for i in seen if i in self.rating_buffer)
        for _ in range(min(pool, self.negative_sample_rate)):
            neg = self._rng.choice(self.item_ids)
            while neg in seen:
                neg = self._rng.choice(self.item_ids)
            self.rating_buffer[neg].append(Rating(data.user, neg, 0.0, data.timestamp))
            ps.pull(neg)
        q = self.rating_buffer.get(data.item)
        if q is None:
            self.item_ids.append(data.item)
            q = self.rating_buffer[data.item] = deque()
        q.append(data)
        ps.pull(data.item)

    def on_pull_recv(self, param_id: int, param_value, ps: ParameterServerClient) -> None:
        rating = self.rating_buffer[param_id].popleft()
        if self._init is None:
            self._init = self.factor_init_desc.open()
        user = self.user_vectors.get(rating.user)
        if user is None:
            user = self._init.next_factor(rating.user)
        d_user, d_item = self.factor_update.delta(rating.rating, user, param_value)
        new_user = vector_sum(user, d_user)
        self.user_vectors[rating.user] = new_user
        ps.output((rating.user, new_user))
        ps.push(param_id, d_item)


def lemp_top_k(user_and_len, items_desc: List[Tuple[float, int]], model: Dict[int, tuple],
               worker_k: int, bucket_size: int, pruning: LEMPPruningStrategy,
               stats: Optional[dict] = None) -> TopKQueue:
    """Bucketed LEMP retrieval over items sorted by length descending.

    ``items_desc`` is ``[(length, id), ...]``; ``model[id] = (length, vector)``.
    Parity: ``PSTopKGeneratorWorker.scala:49-111`` (same loop in the learner+generator worker)."""
    q_len, q = user_and_len
    topk = TopKQueue()
    focus = focus_coordinate(q) if len(q) else 0
    fset = focus_set(q, getattr(pruning, "num_focus", 0))
    user = (q_len, q)
    for b in range(0, len(items_desc), bucket_size):
        bucket = items_desc[b:b + bucket_size]
        if len(topk) >= worker_k and bucket[0][0] * q_len <= topk.head[0]:
            break
        theta = 0.0 if len(topk) < worker_k else topk.head[0]
        keep = bucket_filter(pruning, user, theta, bucket[0][0], bucket[-1][0], focus, fset)
        for plen, pid in bucket:
            item = (pid, model[pid])
            if len(topk) >= worker_k and not keep(item):
                if stats is not None:
                    stats["pruned"] = stats.get("pruned", 0) + 1
                continue
            if stats is not None:
                stats["scored"] = stats.get("scored", 0) + 1
            topk.offer(dot_product(q, item[1][1]), pid, worker_k)
    return topk


class _LengthSortedItems:
    """Items kept sorted by length descending; re-sorted lazily (the GPU path keeps tiles and
    re-sorts periodically instead of maintaining a tree)."""

    def __init__(self):
        self.model: Dict[int, tuple] = {}
        self._sorted: Optional[List[Tuple[float, int]]] = None

    def put(self, item_id: int, len_and_vec) -> None:
        self.model[item_id] = len_and_vec
        self._sorted = None

    def desc(self) -> List[Tuple[float, int]]:
        if self._sorted is None:
            self._sorted = sorted(((lv[0], i) for i, lv in self.model.items()),
                                  key=lambda t: (-t[0], -t[1]))
        return self._sorted


class PSTopKGeneratorWorker(BaseMFWorkerLogic):
    """Inference-only top-K: items on the workers, users on the PS."""

    def __init__(self, worker_k: int, bucket_size: int, worker_parallelism: int,
                 pruning: LEMPPruningStrategy):
        super().__init__()
        self.worker_k, self.bucket_size, self.W, self.pruning = worker_k, bucket_size, worker_parallelism, pruning
        self.items = _LengthSortedItems()
        self.model = self.items.model
        self.rating_buffer: Dict[int, Deque[RichRating]] = {}
        self.stats: dict = {}

    @staticmethod
    def invalid_user(length: float) -> bool:
        return length == -1

    def on_recv(self, data: RichRating, ps) -> None:
        self.rating_buffer.setdefault(data.user, deque()).append(data)
        ps.pull(data.user)

    def on_pull_recv(self, param_id: int, user_and_len, ps) -> None:
        rate = self.rating_buffer[param_id].popleft()
        if self.invalid_user(user_and_len[0]):
            ps.output((rate, TopKQueue()))
            return
        user = (user_and_len[0], np.asarray(user_and_len[1], dtype=np.float64))
        ps.output((rate, lemp_top_k(user, self.items.desc(), self.items.model, self.worker_k,
                                    self.bucket_size, self.pruning, self.stats)))

    def update_model(self, param_id: int, param) -> None:
        self.items.put(param_id, (param[0], np.asarray(param[1], dtype=np.float64)))


class PSOnlineMatrixFactorizationAndTopKGeneratorWorker(BaseMFWorkerLogic):
    """Prequential recommend-then-learn: answer the query with the current model, then - only on
    the worker that owns the rated item - learn from it (negatives are local items)."""

    def __init__(self, negative_sample_rate: int, user_memory: int, worker_k: int, bucket_size: int,
                 pruning_algorithm: LEMPPruningStrategy, worker_parallelism: int,
                 factor_init_desc: FactorInitializerDescriptor, factor_update: FactorUpdater,
                 *, seed: int = 0):
        super().__init__()
        self.negative_sample_rate, self.user_memory = negative_sample_rate, user_memory
        self.worker_k, self.bucket_size, self.pruning = worker_k, bucket_size, pruning_algorithm
        self.W = worker_parallelism
        self.factor_init_desc, self.factor_update = factor_init_desc, factor_update
        self.items = _LengthSortedItems()
        self.model = self.items.model
        self.item_ids_buffer: List[int] = []
        self.rating_buffer: Dict[int, Deque[RichRating]] = {}
        self.seen_list: Dict[int, Deque[int]] = {}
        self.seen_set: Dict[int, set] = {}
        self.worker_id = -1
        self.seed = seed
        self._rng = random.Random(seed)
        self._init = None

    def open(self, runtime_context: RuntimeContext | None = None) -> None:
        if runtime_context is not None:
            self.worker_id = runtime_context.index_of_this_subtask   # explicit rank
            self._rng = random.Random((self.seed << 8) ^ self.worker_id)
        self._init = self.factor_init_desc.open()

    def owns(self, item: int) -> bool:
        return item % self.W == self.worker_id          # Python % is non-negative

    def on_recv(self, data: RichRating, ps) -> None:
        if self.worker_id == -1:
            self.worker_id = data.target_worker
        self.rating_buffer.setdefault(data.user, deque()).append(data)
        ps.pull(data.user)

    def update_model(self, item_id: int, param) -> None:
        if item_id not in self.items.model:
            self.item_ids_buffer.append(item_id)
        self.items.put(item_id, (param[0], np.asarray(param[1], dtype=np.float64)))

    def initialize(self, item_id: int):
        if self._init is None:
            self._init = self.factor_init_desc.open()
        lv = attach_length(self._init.next_factor(item_id))
        self.update_model(item_id, lv)
        return lv

    def on_pull_recv(self, param_id: int, user_and_len, ps) -> None:
        rate = self.rating_buffer[param_id].popleft()
        q_len, q = user_and_len[0], np.asarray(user_and_len[1], dtype=np.float64)
        topk = lemp_top_k((q_len, q), self.items.desc(), self.items.model, self.worker_k,
                          self.bucket_size, self.pruning)
        ps.output((rate, topk))
        if not self.owns(rate.item):
            return
        s = self.seen_set.setdefault(rate.user, set())
        if rate.item not in s:
            s.add(rate.item)
            lst = self.seen_list.setdefault(rate.user, deque())
            lst.append(rate.item)
            if len(lst) > self.user_memory:
                s.discard(lst.popleft())
        u_delta = np.zeros_like(q)
        for _ in range(min(len(self.items.model) - len(s), self.negative_sample_rate)):
            neg = self._rng.choice(self.item_ids_buffer)
            counter = 32
            while counter > 0 and neg in s:
                neg = self._rng.choice(self.item_ids_buffer)
                counter -= 1
            if counter > 0:
                _, neg_vec = self.items.model[neg]
                uu, idelta = self.factor_update.delta(0.0, q, neg_vec)
                u_delta = vector_sum(u_delta, uu)
                self.items.put(neg, attach_length(vector_sum(neg_vec, idelta)))
        lv = self.items.model.get(rate.item)
        if lv is None:
            lv = self.initialize(rate.item)
        user_delta, item_delta = self.factor_update.delta(rate.rating, q, lv[1])
        self.items.put(rate.item, attach_length(vector_sum(lv[1], item_delta)))
        ps.push(param_id, (float("nan"), vector_sum(u_delta, user_delta)))
