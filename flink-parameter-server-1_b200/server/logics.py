"""Parameter-server side logics (host engine).

Parity:
  * ``SimplePSLogic``           - ``M/server/SimplePSLogic.scala:7-26``
  * ``SimplePSLogicWithClose``  - ``M/server/SimplePSLogicWithClose.scala:7-32``
  * ``RangePSLogicWithClose``   - ``M/server/RangePSLogicWithClose.scala:7-63``
  * ``LockPSLogicA`` / ``B``    - ``M/server/LockPSLogicA.scala:13-46``, ``LockPSLogicB.scala:15-50``

``param_init`` / ``param_update`` may be passed either as the functions themselves or as
zero-argument *factories* (the reference takes them by-name and re-materialises them per
task so they can capture an RNG); pass ``by_name=True`` for the factory form.
"""
from __future__ import annotations

import math
from collections import deque
from typing import Any, Callable, Deque, Dict, List, Optional

from ..api import ParameterServer, ParameterServerLogic, RuntimeContext


def _materialise(fn, by_name: bool):
    return fn() if by_name else fn


class SimplePSLogic(ParameterServerLogic):
    """Hash-map store.  Pull initialises lazily; a push to an absent id stores the delta
    as the value; *every push emits* ``(id, new_value)``."""

    def __init__(self, param_init: Callable[[int], Any], param_update: Callable[[Any, Any], Any],
                 by_name: bool = False):
        self._init_src, self._update_src, self._by_name = param_init, param_update, by_name
        self.params: Dict[int, Any] = {}
        self.init = _materialise(param_init, by_name)
        self.update = _materialise(param_update, by_name)

    def open(self, parameters=None, runtime_context: RuntimeContext | None = None) -> None:
        # per-task re-materialisation of by-name arguments
        self.init = _materialise(self._init_src, self._by_name)
        self.update = _materialise(self._update_src, self._by_name)

    def on_pull_recv(self, param_id: int, worker_partition_index: int, ps: ParameterServer) -> None:
        if param_id not in self.params:
            self.params[param_id] = self.init(param_id)
        ps.answer_pull(param_id, self.params[param_id], worker_partition_index)

    def _apply(self, param_id: int, delta_update):
        if param_id in self.params:
            c = self.update(self.params[param_id], delta_update)
        else:
            c = delta_update
        self.params[param_id] = c
        return c

    def on_push_recv(self, param_id: int, delta_update, ps: ParameterServer) -> None:
        ps.output((param_id, self._apply(param_id, delta_update)))


class SimplePSLogicWithClose(SimplePSLogic):
    """Same store; pushes are silent, ``close`` dumps every live ``(id, value)``."""

    def on_push_recv(self, param_id: int, delta_update, ps: ParameterServer) -> None:
        self._apply(param_id, delta_update)

    def close(self, ps: ParameterServer) -> None:
        for k, v in self.params.items():
            ps.output((k, v))


def range_partition_div(feature_count: int, ps_parallelism: int) -> int:
    """``ceil(featureCount / psParallelism)`` - the one constant shared by the range
    partitioner (``PassiveAggressiveParameterServer.scala:372-375``) and the shard-local slot
    arithmetic (``RangePSLogicWithClose.scala:53-61``).  The device table uses the same."""
    return int(math.ceil(feature_count / float(ps_parallelism)))


class RangePSLogicWithClose(ParameterServerLogic):
    """Dense per-shard array for a contiguous id range, dumped on close (never-touched slots
    are skipped)."""

    def __init__(self, feature_count: int, param_init, param_update, by_name: bool = False):
        self.feature_count = int(feature_count)
        self._init_src, self._update_src, self._by_name = param_init, param_update, by_name
        self.start_index = 0
        self.params: List[Optional[Any]] = []
        self._live: List[bool] = []
        self.init = _materialise(param_init, by_name)
        self.update = _materialise(param_update, by_name)

    def open(self, parameters=None, runtime_context: RuntimeContext | None = None) -> None:
        ctx = runtime_context or RuntimeContext(0, 1)
        n, idx = ctx.number_of_parallel_subtasks, ctx.index_of_this_subtask
        div = range_partition_div(self.feature_count, n)
        mod = self.feature_count - (n - 1) * div
        size = mod if (mod != 0 and idx + 1 == n) else div
        size = max(size, 0)
        self.params = [None] * size
        self._live = [False] * size
        self.start_index = idx * div
        self.init = _materialise(self._init_src, self._by_name)
        self.update = _materialise(self._update_src, self._by_name)

    def _slot(self, param_id: int) -> int:
        slot = param_id - self.start_index
        if slot < 0 or slot >= len(self.params):
            raise IndexError(
                f"parameter id {param_id} is outside this shard's range "
                f"[{self.start_index}, {self.start_index + len(self.params)})")
        return slot

    def on_pull_recv(self, param_id: int, worker_partition_index: int, ps: ParameterServer) -> None:
        s = self._slot(param_id)
        if not self._live[s]:
            self.params[s] = self.init(param_id)
            self._live[s] = True
        ps.answer_pull(param_id, self.params[s], worker_partition_index)

    def on_push_recv(self, param_id: int, delta_update, ps: ParameterServer) -> None:
        s = self._slot(param_id)
        self.params[s] = self.update(self.params[s], delta_update) if self._live[s] else delta_update
        self._live[s] = True

    def close(self, ps: ParameterServer) -> None:
        for i, live in enumerate(self._live):
            if live:
                ps.output((self.start_index + i, self.params[i]))


class _LockEntry:
    __slots__ = ("locked", "value", "queue")

    def __init__(self, value):
        self.locked = False
        self.value = value
        self.queue: Deque[int] = deque()


class LockPSLogicA(ParameterServerLogic):
    """Per-key lock giving read-modify-write consistency: a pull takes the lock, the matching
    push releases it or hands the *new* value to the next waiter.  Duplicate waiters kept."""

    dedupe_waiters = False

    def __init__(self, param_init, param_update):
        self.init, self.update = param_init, param_update
        self.params: Dict[int, _LockEntry] = {}

    def on_pull_recv(self, param_id: int, worker_partition_index: int, ps: ParameterServer) -> None:
        e = self.params.get(param_id)
        if e is None:
            e = self.params[param_id] = _LockEntry(self.init(param_id))
        if not e.locked:
            ps.answer_pull(param_id, e.value, worker_partition_index)
            e.locked = True
        elif not (self.dedupe_waiters and worker_partition_index in e.queue):
            e.queue.append(worker_partition_index)

    def on_push_recv(self, param_id: int, delta_update, ps: ParameterServer) -> None:
        e = self.params.get(param_id)
        if e is None:
            raise RuntimeError("Not existed model was not able to update by any delta.")
        e.value = self.update(e.value, delta_update)
        if e.queue:
            ps.answer_pull(param_id, e.value, e.queue.popleft())
            e.locked = True
        else:
            e.locked = False
        ps.output((param_id, e.value))


class LockPSLogicB(LockPSLogicA):
    """As A, but a worker already waiting on a key is not queued twice."""

    dedupe_waiters = True
