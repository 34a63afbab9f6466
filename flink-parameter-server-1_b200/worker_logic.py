"""Pull limiters and the future-style worker API.

Parity:
  * ``add_pull_limiter``           - ``M/WorkerLogic.scala:176-225`` (non-blocking; FIFO queue of
    excess pulls, one re-issued per answer *after* the wrapped ``on_pull_recv`` ran)
  * ``add_blocking_pull_limiter``  - ``M/WorkerLogic.scala:82-156`` (pull blocks on a condition
    variable; the worker must pull from a second thread)
  * ``WorkerLogicWithFuture``      - ``M/WorkerLogic.scala:239-356``.  The reference registers the
    future but never sends the pull; this one does send it and chains callbacks
    without recursion.

The device engine implements the same bound with a credit counter in device memory
(``csrc/flow.cuh``, per-CTA allowances in ``csrc/mf_tiles.cu``); ``pull_limit`` means the same thing in both engines:
*the number of unanswered pulls per worker instance never exceeds the limit*.
"""
from __future__ import annotations

import threading
from collections import deque
from typing import Any, Callable, Deque, Dict, List, Optional, Tuple

from .api import ParameterServerClient, RuntimeContext, WorkerLogic


class _ForwardingLogic(WorkerLogic):
    """Shared plumbing: forwards life-cycle hooks (and ``update_model`` for MF logics,
    cf. ``BaseMFWorkerLogic.scala:8-188``) to the wrapped logic."""

    def __init__(self, inner: WorkerLogic):
        self.inner = inner

    def open(self, runtime_context: RuntimeContext | None = None) -> None:
        self.inner.open(runtime_context)

    def close(self) -> None:
        self.inner.close()

    def update_model(self, param_id, param) -> None:  # BaseMFWorkerLogic passthrough
        self.inner.update_model(param_id, param)

    @property
    def needs_realtime(self) -> bool:
        return bool(getattr(self.inner, "needs_realtime", False))


class _LimitedClient(ParameterServerClient):
    def __init__(self, owner: "PullLimiter"):
        self._owner = owner
        self.ps: ParameterServerClient = None  # type: ignore[assignment]

    def pull(self, param_id: int) -> None:
        o = self._owner
        if o.pull_counter < o.pull_limit:
            o.pull_counter += 1
            o.max_in_flight = max(o.max_in_flight, o.pull_counter)
            self.ps.pull(param_id)
        else:
            o.stalled_pulls += 1
            o.pull_queue.append(param_id)

    def push(self, param_id: int, delta_update) -> None:
        self.ps.push(param_id, delta_update)

    def output(self, out) -> None:
        self.ps.output(out)


class PullLimiter(_ForwardingLogic):
    """Non-blocking pull limiter."""

    def __init__(self, inner: WorkerLogic, pull_limit: int):
        super().__init__(inner)
        if pull_limit <= 0:
            raise ValueError("pull_limit must be positive")
        self.pull_limit = int(pull_limit)
        self.pull_counter = 0
        self.pull_queue: Deque[int] = deque()
        # observability (no reference counterpart): high-water mark and number of queued pulls
        self.max_in_flight = 0
        self.stalled_pulls = 0
        self._client = _LimitedClient(self)

    def on_recv(self, data, ps: ParameterServerClient) -> None:
        self._client.ps = ps
        self.inner.on_recv(data, self._client)

    def on_pull_recv(self, param_id: int, param_value, ps: ParameterServerClient) -> None:
        self._client.ps = ps
        self.inner.on_pull_recv(param_id, param_value, self._client)
        self.pull_counter -= 1
        if self.pull_queue:
            self._client.pull(self.pull_queue.popleft())


def add_pull_limiter(worker_logic: WorkerLogic, pull_limit: int) -> PullLimiter:
    return PullLimiter(worker_logic, pull_limit)


class BlockingPullLimiter(_ForwardingLogic):
    """Blocking variant: ``pull`` waits while ``pull_counter >= pull_limit``."""

    needs_realtime = True

    def __init__(self, inner: WorkerLogic, pull_limit: int):
        super().__init__(inner)
        self.pull_limit = int(pull_limit)
        self.pull_counter = 0
        self.max_in_flight = 0
        self._lock = threading.RLock()
        self._can_pull = threading.Condition(self._lock)
        outer = self

        class _Client(ParameterServerClient):
            ps: ParameterServerClient = None  # type: ignore[assignment]

            def pull(self, param_id: int) -> None:
                with outer._lock:
                    while outer.pull_counter >= outer.pull_limit:
                        outer._can_pull.wait()
                    outer.pull_counter += 1
                    outer.max_in_flight = max(outer.max_in_flight, outer.pull_counter)
                    self.ps.pull(param_id)

            def push(self, param_id: int, delta_update) -> None:
                with outer._lock:
                    self.ps.push(param_id, delta_update)

            def output(self, out) -> None:
                with outer._lock:
                    self.ps.output(out)

        self._client = _Client()

    def _set_ps(self, ps):
        with self._lock:
            self._client.ps = ps

    def on_recv(self, data, ps) -> None:
        self._set_ps(ps)
        self.inner.on_recv(data, self._client)

    def on_pull_recv(self, param_id, param_value, ps) -> None:
        self._set_ps(ps)
        self.inner.on_pull_recv(param_id, param_value, self._client)
        with self._lock:
            self.pull_counter -= 1
            self._can_pull.notify()


def add_blocking_pull_limiter(worker_logic: WorkerLogic, pull_limit: int) -> BlockingPullLimiter:
    return BlockingPullLimiter(worker_logic, pull_limit)


# --- future-style API -------------------------------------------------------------------------

class PullAnswerFuture:
    """A minimal future completed by the matching pull answer."""

    def __init__(self, param_id: int):
        self.param_id = param_id
        self._callbacks: List[Callable[[Tuple[int, Any]], Any]] = []
        self._answer: Optional[Tuple[int, Any]] = None

    def pull_arrived(self, param_id: int, param) -> None:
        self._answer = (param_id, param)
        callbacks, self._callbacks = self._callbacks, []
        for cb in callbacks:
            cb(self._answer)

    def on_complete(self, f: Callable[[Tuple[int, Any]], Any]) -> None:
        if self._answer is None:
            self._callbacks.append(f)
        else:
            f(self._answer)

    @property
    def is_completed(self) -> bool:
        return self._answer is not None

    @property
    def value(self) -> Optional[Tuple[int, Any]]:
        return self._answer

    def result(self) -> Tuple[int, Any]:
        if self._answer is None:
            raise RuntimeError("pull answer has not arrived; register on_complete instead")
        return self._answer


class PSClientWithFuture:
    def __init__(self, waiters: Dict[int, Deque[PullAnswerFuture]]):
        self._waiters = waiters
        self.ps: ParameterServerClient = None  # type: ignore[assignment]

    def pull(self, param_id: int) -> PullAnswerFuture:
        fut = PullAnswerFuture(param_id)
        self._waiters.setdefault(param_id, deque()).append(fut)
        self.ps.pull(param_id)  # the reference forgets this line
        return fut

    def push(self, param_id: int, delta_update) -> None:
        self.ps.push(param_id, delta_update)

    def output(self, out) -> None:
        self.ps.output(out)


class WorkerLogicWithFuture(WorkerLogic):
    """Subclass and implement ``on_data_recv(data, ps)`` where ``ps.pull`` returns a future.
    The n-th answer for an id completes the n-th future for that id (per-key FIFO,
    SURVEY §2.9 "Ordering and matching")."""

    def __init__(self):
        self._pull_waiter: Dict[int, Deque[PullAnswerFuture]] = {}
        self._future_client = PSClientWithFuture(self._pull_waiter)

    def on_data_recv(self, data, ps: PSClientWithFuture) -> None:
        raise NotImplementedError

    def on_recv(self, data, ps: ParameterServerClient) -> None:
        self._future_client.ps = ps
        self.on_data_recv(data, self._future_client)

    def on_pull_recv(self, param_id: int, param_value, ps: ParameterServerClient) -> None:
        self._future_client.ps = ps
        self._pull_waiter[param_id].popleft().pull_arrived(param_id, param_value)
