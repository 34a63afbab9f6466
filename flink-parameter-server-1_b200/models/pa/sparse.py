"""Sparse vectors for the passive-aggressive workloads.

``SparseVector`` replaces Breeze's ``SparseVector[Double]`` (index/value arrays + logical length,
``M/passive/aggressive/PassiveAggressiveParameterServer.scala:29``).  ``LegacySparseVector`` /
``EOFSign`` restate the hand-rolled map-backed vector of the offline jobs
(``M/passive/aggressive/entities/SparseVector.scala:6-42``).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Generic, Iterable, Iterator, Tuple, TypeVar

import numpy as np

E = TypeVar("E")


class SparseVector:
    __slots__ = ("indices", "values", "length")

    def __init__(self, indices, values, length: int):
        idx = np.asarray(indices, dtype=np.int64)
        val = np.asarray(values, dtype=np.float64)
        order = np.argsort(idx, kind="stable")
        self.indices, self.values, self.length = idx[order], val[order], int(length)

    @staticmethod
    def from_pairs(pairs: Iterable[Tuple[int, float]], length: int) -> "SparseVector":
        pairs = list(pairs)
        return SparseVector([p[0] for p in pairs], [p[1] for p in pairs], length)

    @property
    def active_size(self) -> int:
        return len(self.indices)

    def index_at(self, off: int) -> int:
        return int(self.indices[off])

    def value_at(self, off: int) -> float:
        return float(self.values[off])

    def active_iterator(self) -> Iterator[Tuple[int, float]]:
        return zip(self.indices.tolist(), self.values.tolist())

    def dot(self, other) -> float:
        if isinstance(other, SparseVector):
            common, ia, ib = np.intersect1d(self.indices, other.indices, return_indices=True)
            return float(np.dot(self.values[ia], other.values[ib]))
        return float(np.dot(self.values, np.asarray(other)[self.indices]))

    def norm_sqr(self) -> float:
        return float(np.dot(self.values, self.values))

    def scale(self, a: float) -> "SparseVector":
        return SparseVector(self.indices, self.values * a, self.length)

    # hashable so a vector can be its own prediction id (VecId = the vector itself)
    def _key(self):
        return (self.length, self.indices.tobytes(), self.values.tobytes())

    def __hash__(self):
        return hash(self._key())

    def __eq__(self, o):
        return isinstance(o, SparseVector) and self._key() == o._key()

    def __repr__(self):
        return f"SparseVector(nnz={self.active_size}, length={self.length})"


class LegacySparseVector(Generic[E]):
    """Immutable ``Map[Long, E]`` vector with value equality."""

    def __init__(self, data: Dict[int, E]):
        self._data = dict(data)

    @staticmethod
    def build_from_keys_values(keys: Iterable[int], values: Iterable[E]) -> "LegacySparseVector":
        return LegacySparseVector(dict(zip(keys, values)))

    @staticmethod
    def build_from_pairs(pairs: Iterable[Tuple[int, E]]) -> "LegacySparseVector":
        return LegacySparseVector(dict(pairs))

    def get_values(self) -> Dict[int, E]:
        return self._data

    def keys(self):
        return self._data.keys()

    def __len__(self):
        return len(self._data)

    def __eq__(self, o):
        return isinstance(o, LegacySparseVector) and self._data == o._data

    def __hash__(self):
        return hash(frozenset(self._data.items()))


class EOFSign(LegacySparseVector):
    """Sentinel vector: ``EOFSign(workerId, minusSourceId)``."""

    def __init__(self, worker_id: int, minus_source_id: int):
        super().__init__({})
        self.worker_id, self.minus_source_id = worker_id, minus_source_id

    def __eq__(self, o):
        return isinstance(o, EOFSign) and (self.worker_id, self.minus_source_id) == (o.worker_id, o.minus_source_id)

    def __hash__(self):
        return hash(("EOFSign", self.worker_id, self.minus_source_id))
