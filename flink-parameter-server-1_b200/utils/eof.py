"""End-of-input barrier for finite streams.

Parity: ``M/utils/FlinkEOF.scala:9-124``.  Every upstream subtask, when it finishes, sends one
``EOF(source, target)`` to *each* downstream subtask; data records are routed by the user's
partitioner; a downstream subtask calls ``on_eof`` once it has counted an EOF from every
upstream subtask.  Because each (source, target) channel is FIFO, no data record can follow the
last EOF.  An empty upstream partition still sends its EOFs (the reference throws).
"""
from __future__ import annotations

import itertools
from collections import deque
from typing import Any, Callable, Iterable, Iterator, List

from ..entities import EOF
from ..parallel.streams import DataStream, as_stream


class EOFHandler:
    """Implement ``flat_map(value) -> iterable`` and ``on_eof() -> iterable``."""

    def open(self, subtask_index: int, parallelism: int) -> None:
        pass

    def flat_map(self, value) -> Iterable:
        raise NotImplementedError

    def on_eof(self) -> Iterable:
        return ()

    def close(self) -> None:
        pass


class _EOFRouter:
    """Pumps the upstream partitions fairly and files records / EOF marks into the downstream
    FIFO queues."""

    def __init__(self, upstream: List[Iterator], n_down: int, route: Callable[[Any], int]):
        self.up = upstream
        self.alive = [True] * len(upstream)
        self.queues = [deque() for _ in range(n_down)]
        self.route = route
        self.n_down = n_down
        self._rr = 0

    def pump(self) -> bool:
        n = len(self.up)
        for k in range(n):
            i = (self._rr + k) % n
            if not self.alive[i]:
                continue
            self._rr = (i + 1) % n
            try:
                rec = next(self.up[i])
            except StopIteration:
                self.alive[i] = False
                for t in range(self.n_down):
                    self.queues[t].append(EOF(i, t))
                return True
            t = self.route(rec)
            if not 0 <= t < self.n_down:
                raise RuntimeError(f"partitioner returned {t} for {self.n_down} partitions")
            self.queues[t].append(rec)
            return True
        return False


def flat_map_with_eof(stream, handler_factory: Callable[[], EOFHandler], downstream_parallelism: int,
                      partitioner: Callable[[Any, int], int], key_selector: Callable[[Any], Any]) -> DataStream:
    """``FlinkEOF.flatMapWithEOF(in, flatMapFunction, downstreamParallelism, partitioner, keyFn)``.
    ``handler_factory`` builds one handler per downstream subtask."""
    s = stream if isinstance(stream, DataStream) else as_stream(stream, 1)
    in_par = s.parallelism
    router = _EOFRouter(s.partitions(), downstream_parallelism,
                        lambda rec: partitioner(key_selector(rec), downstream_parallelism))

    def downstream(t: int):
        h = handler_factory()
        h.open(t, downstream_parallelism)
        eof_count = 0
        q = router.queues[t]
        try:
            while True:
                if not q:
                    if not router.pump():
                        if not q:
                            return
                    continue
                rec = q.popleft()
                if isinstance(rec, EOF):
                    eof_count += 1
                    if eof_count >= in_par:
                        yield from h.on_eof()
                else:
                    if eof_count >= in_par:
                        raise RuntimeError("data record after end-of-input")
                    yield from h.flat_map(rec)
        finally:
            h.close()

    return DataStream([downstream(t) for t in range(downstream_parallelism)])


class WrapWithEOF(EOFHandler):
    """The handler every reference call site uses: pass data through as ``Right(x)`` and emit
    ``Left(EOF)`` at the barrier (``M/matrix/factorization/PSOfflineMatrixFactorization.scala:60-75``)."""

    def flat_map(self, value):
        from ..entities import Right
        return (Right(value),)

    def on_eof(self):
        from ..entities import Left
        return (Left(EOF()),)
