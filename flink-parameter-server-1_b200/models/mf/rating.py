"""``Rating`` / ``RichRating`` (``M/matrix/factorization/utils/Rating.scala:9-31``) and the small
helpers of ``utils/Utils.scala:7-68`` (top-K min-heap, id generator, hash partitioner)."""
from __future__ import annotations

import heapq
import itertools
import threading
from dataclasses import dataclass
from typing import List, Tuple


@dataclass(frozen=True)
class Rating:
    user: int
    item: int
    rating: float
    timestamp: int = 0

    @staticmethod
    def from_tuple(user: int, item: int, rating: float) -> "Rating":
        return Rating(user, item, rating, 0)

    def enrich(self, worker_id: int, rating_id: float) -> "RichRating":
        return RichRating(self.user, self.item, self.rating, worker_id, rating_id, self.timestamp)


@dataclass(frozen=True)
class RichRating:
    user: int
    item: int
    rating: float
    target_worker: int
    rating_id: float
    timestamp: int = 0

    def reduce(self) -> Rating:
        return Rating(self.user, self.item, self.rating, self.timestamp)


class TopKQueue:
    """Bounded min-heap of ``(score, item)``: ``head`` is the *smallest* retained score
    (the reference's reverse-ordered PriorityQueue, ``Utils.scala:28-30``)."""

    def __init__(self, items=()):
        self._h: List[Tuple[float, int]] = list(items)
        heapq.heapify(self._h)

    def __len__(self):
        return len(self._h)

    @property
    def head(self) -> Tuple[float, int]:
        return self._h[0]

    def push(self, score: float, item: int) -> None:
        heapq.heappush(self._h, (score, item))

    def replace_head(self, score: float, item: int) -> None:
        heapq.heapreplace(self._h, (score, item))

    def offer(self, score: float, item: int, k: int) -> None:
        if len(self._h) < k:
            heapq.heappush(self._h, (score, item))
        elif self._h[0][0] < score:
            heapq.heapreplace(self._h, (score, item))

    def to_list(self) -> List[Tuple[float, int]]:
        return list(self._h)

    def sorted_desc(self) -> List[Tuple[float, int]]:
        return sorted(self._h, key=lambda t: (-t[0], t[1]))

    def __iter__(self):
        return iter(self._h)


class IDGenerator:
    """Process-wide monotonically increasing id (``Utils.scala:39-46``)."""

    _counter = itertools.count()
    _lock = threading.Lock()

    @classmethod
    def next(cls) -> int:
        with cls._lock:
            return next(cls._counter)


def hash_partition(param_id: int, ps_parallelism: int) -> int:
    """``Utils.Partitioner`` (``:51-65``): non-negative ``hash(id) % n``."""
    return abs(int(param_id)) % ps_parallelism
