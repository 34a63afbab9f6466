"""A small partitioned-stream abstraction standing in for the slice of Flink's DataStream API
the reference uses to *feed* a parameter-server job:
``rebalance``, ``partitionCustom``, ``broadcast``-by-copy, ``map`` / ``flatMap`` / ``filter``,
``union`` and ``setParallelism``.  Streams are lazy: partitions are iterators, so an unbounded
generator works for the online workloads.

Parity: call sites ``M/FlinkParameterServer.scala:215-336``;
``M/matrix/factorization/PSOnlineMatrixFactorization.scala:62-64`` (partition by ``user % n``);
``M/matrix/factorization/PSTopKGenerator.scala:78-89`` (copy to all workers).
"""
from __future__ import annotations

import itertools
from collections import deque
from typing import Any, Callable, Iterable, Iterator, List, Optional, Sequence


class _Router:
    """Lazily routes one upstream iterator into ``n`` downstream FIFO partitions."""

    def __init__(self, upstream: Iterator, n: int, route: Callable[[Any], Iterable[int]]):
        self._up = upstream
        self._route = route
        self._queues = [deque() for _ in range(n)]
        self._done = False

    def _pump(self) -> bool:
        if self._done:
            return False
        try:
            rec = next(self._up)
        except StopIteration:
            self._done = True
            return False
        for t in self._route(rec):
            self._queues[t].append(rec)
        return True

    def partition(self, i: int) -> Iterator:
        q = self._queues[i]
        while True:
            if q:
                yield q.popleft()
            elif not self._pump():
                if q:
                    continue
                return


class DataStream:
    """``partitions[i]`` is the iterator feeding subtask ``i``."""

    def __init__(self, partitions: Sequence[Iterable]):
        self._partitions: List[Iterator] = [iter(p) for p in partitions]

    # -- construction ---------------------------------------------------------------------
    @staticmethod
    def from_collection(records: Iterable, parallelism: int = 1) -> "DataStream":
        if parallelism == 1:
            return DataStream([records])
        return DataStream([records]).rebalance(parallelism)

    @staticmethod
    def from_partitions(partitions: Sequence[Iterable]) -> "DataStream":
        return DataStream(partitions)

    @property
    def parallelism(self) -> int:
        return len(self._partitions)

    def partitions(self) -> List[Iterator]:
        return self._partitions

    # -- element-wise ---------------------------------------------------------------------
    def map(self, f: Callable) -> "DataStream":
        return DataStream([map(f, p) for p in self._partitions])

    def flat_map(self, f: Callable[[Any], Iterable]) -> "DataStream":
        return DataStream([itertools.chain.from_iterable(map(f, p)) for p in self._partitions])

    def filter(self, f: Callable[[Any], bool]) -> "DataStream":
        return DataStream([filter(f, p) for p in self._partitions])

    # -- repartitioning -------------------------------------------------------------------
    def _merged(self) -> Iterator:
        """Round-robin interleave of the upstream partitions (a fair stand-in for the
        arbitrary interleaving of a network shuffle)."""
        active = list(self._partitions)
        while active:
            nxt = []
            for it in active:
                try:
                    yield next(it)
                    nxt.append(it)
                except StopIteration:
                    pass
            active = nxt

    def rebalance(self, parallelism: int) -> "DataStream":
        counter = itertools.count()
        r = _Router(self._merged(), parallelism, lambda rec: (next(counter) % parallelism,))
        return DataStream([r.partition(i) for i in range(parallelism)])

    def partition_custom(self, key_selector: Callable[[Any], int], parallelism: int,
                         partitioner: Optional[Callable[[int, int], int]] = None) -> "DataStream":
        """``partitioner(key, n)`` defaults to non-negative ``key % n`` (the
        reference forgets ``abs``)."""
        part = partitioner or (lambda k, n: k % n)

        def route(rec):
            t = part(key_selector(rec), parallelism)
            if not 0 <= t < parallelism:
                raise RuntimeError(f"partitioner returned {t} for {parallelism} partitions")
            return (t,)

        r = _Router(self._merged(), parallelism, route)
        return DataStream([r.partition(i) for i in range(parallelism)])

    def broadcast(self, parallelism: int) -> "DataStream":
        r = _Router(self._merged(), parallelism, lambda rec: range(parallelism))
        return DataStream([r.partition(i) for i in range(parallelism)])

    def forward(self, parallelism: int) -> "DataStream":
        if parallelism == self.parallelism:
            return self
        return self.rebalance(parallelism)

    set_parallelism = forward

    def union(self, other: "DataStream") -> "DataStream":
        if other.parallelism != self.parallelism:
            other = other.rebalance(self.parallelism)

        def interleave(a, b):
            ia, ib = iter(a), iter(b)
            live = [ia, ib]
            while live:
                for it in list(live):
                    try:
                        yield next(it)
                    except StopIteration:
                        live.remove(it)

        return DataStream([interleave(a, b) for a, b in zip(self._partitions, other._partitions)])

    def concat(self, other: "DataStream") -> "DataStream":
        """Partition-wise ``self`` then ``other`` (first stream fully consumed first)."""
        if other.parallelism != self.parallelism:
            other = other.rebalance(self.parallelism)
        return DataStream([itertools.chain(a, b) for a, b in zip(self._partitions, other._partitions)])

    def collect(self) -> list:
        return list(self._merged())


def as_stream(data, parallelism: int) -> DataStream:
    """Anything iterable becomes a stream of the wanted parallelism (``forward`` when it already
    has it, ``rebalance`` otherwise - what ``trainingData.map(x=>x).forward.setParallelism(W)``
    amounts to, ``M/FlinkParameterServer.scala:331-334``)."""
    if isinstance(data, DataStream):
        return data.forward(parallelism)
    return DataStream.from_collection(data, parallelism)
