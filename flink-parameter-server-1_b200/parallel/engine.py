"""E-host: the generic host engine - a single-process event loop that runs arbitrary Python
worker / server logics over arbitrary parameter types with the reference's message semantics.

It replaces (not wraps) the Flink job graph of ``M/FlinkParameterServer.scala:195-336``:

  * ``W`` worker subtasks, ``S`` server subtasks, each with its own deep copy of the user logic
    (Flink ships a serialised copy to every subtask);
  * one FIFO channel per (worker, shard) and per (shard, worker) pair - X1 and X2 in SURVEY §2.7;
    per-channel order is the *only* ordering guarantee, exactly as in the reference (§2.9);
  * ``param_partitioner(msg) % S`` picks the shard (``:270-274``); a pull answer's key must equal
    a valid worker index or the job fails (``:309-317``);
  * outputs ``Left(worker_out)`` and ``Right(ps_out)`` interleaved in emission order (``:325-328``);
  * termination after the feedback edge has been idle for ``iteration_wait_time`` ms
    (``:49-52``).  For a finite input and logics without background threads the loop is
    *quiescent* the moment every queue is empty, so waiting adds nothing and is skipped
    (``realtime=False``); with timers / blocking limiters / threaded logics the wait is real.

The event-selection policy is pluggable so tests can drive the same job through different
interleavings (``round_robin``, ``random`` with a seed, ``input_first`` = maximum staleness,
``feedback_first`` = serial).  This engine is the correctness oracle for the device engine
(``parallel/device_engine.py``) and the home of the pluggable wire components.
"""
from __future__ import annotations

import copy
import random
import threading
import time
from collections import deque
from typing import Any, Callable, Deque, Dict, List, Optional

from ..api import (ParameterServer, ParameterServerClient, ParameterServerLogic, PSReceiver,
                   PSSender, RuntimeContext, WorkerLogic, WorkerReceiver, WorkerSender)
from ..entities import Left, Right
from .streams import DataStream, as_stream


class MessagingPSClient(ParameterServerClient):
    """Binds the client interface to the engine's collector
    (``M/FlinkParameterServer.scala:846-873``)."""

    def __init__(self, sender: WorkerSender, partition_id: int, collect_msg, collect_out):
        self.sender = sender
        self.partition_id = partition_id
        self._collect_msg = collect_msg
        self._collect_out = collect_out

    def pull(self, param_id: int) -> None:
        self.sender.on_pull(param_id, self._collect_msg, self.partition_id)

    def push(self, param_id: int, delta_update) -> None:
        self.sender.on_push(param_id, delta_update, self._collect_msg, self.partition_id)

    def output(self, out) -> None:
        self._collect_out(out)


class MessagingPS(ParameterServer):
    """``M/FlinkParameterServer.scala:821-841``"""

    def __init__(self, sender: PSSender, collect_msg, collect_out):
        self.sender = sender
        self._collect_msg = collect_msg
        self._collect_out = collect_out

    def answer_pull(self, param_id: int, value, worker_partition_index: int) -> None:
        self.sender.on_pull_answer(param_id, value, worker_partition_index, self._collect_msg)

    def output(self, out) -> None:
        self._collect_out(out)


class EngineStats:
    def __init__(self):
        self.worker_to_ps_msgs = 0
        self.ps_to_worker_msgs = 0
        self.input_records = 0
        self.worker_outputs = 0
        self.ps_outputs = 0
        self.events = 0
        self.idle_waits = 0

    def as_dict(self) -> Dict[str, int]:
        return dict(self.__dict__)


def _clone(obj, n: int) -> list:
    """One private copy per subtask.  Objects that cannot be deep-copied (e.g. they hold a
    thread or a lock) may provide ``clone()``."""
    out = []
    for _ in range(n):
        if hasattr(obj, "clone"):
            out.append(obj.clone())
        else:
            out.append(copy.deepcopy(obj))
    return out


class HostEngine:
    def __init__(self, training_data, worker_logic: WorkerLogic, ps_logic: ParameterServerLogic,
                 param_partitioner: Callable[[Any], int], w_in_partition: Callable[[Any], int],
                 worker_parallelism: int, ps_parallelism: int,
                 worker_receiver: WorkerReceiver, worker_sender: WorkerSender,
                 ps_receiver: PSReceiver, ps_sender: PSSender, iteration_wait_time: int,
                 *, schedule: str = "round_robin", seed: int = 0, realtime: Optional[bool] = None,
                 flush_on_close: bool = True, max_events: Optional[int] = None,
                 share_logic_instances: bool = False):
        W, S = int(worker_parallelism), int(ps_parallelism)
        if W <= 0 or S <= 0:
            raise ValueError("parallelism must be positive")
        self.W, self.S = W, S
        self.iteration_wait_time = int(iteration_wait_time)
        self.schedule = schedule
        self.rng = random.Random(seed)
        self.flush_on_close = flush_on_close
        self.max_events = max_events
        self.stats = EngineStats()
        self.param_partitioner = param_partitioner
        self.w_in_partition = w_in_partition

        if share_logic_instances and W == 1 and S == 1:
            self.worker_logics, self.ps_logics = [worker_logic], [ps_logic]
        else:
            self.worker_logics = _clone(worker_logic, W)
            self.ps_logics = _clone(ps_logic, S)
        self.worker_senders = _clone(worker_sender, W)
        self.worker_receivers = _clone(worker_receiver, W)
        self.ps_senders = _clone(ps_sender, S)
        self.ps_receivers = _clone(ps_receiver, S)

        if realtime is None:
            realtime = any(getattr(o, "needs_realtime", False) for o in
                           self.worker_logics + self.ps_logics + self.worker_senders + self.ps_senders)
        self.realtime = bool(realtime)

        self.inputs = as_stream(training_data, W).partitions()
        self._input_done = [False] * W
        self._input_peek: List[Deque] = [deque() for _ in range(W)]
        # channels
        self.w2ps: List[List[Deque]] = [[deque() for _ in range(S)] for _ in range(W)]
        self.ps2w: List[List[Deque]] = [[deque() for _ in range(W)] for _ in range(S)]
        self.outputs: List[Any] = []
        self._cv = threading.Condition()
        self._out_lock = threading.Lock()

        self.clients: List[MessagingPSClient] = []
        for w in range(W):
            self.clients.append(MessagingPSClient(
                self.worker_senders[w], w, self._make_w_collect(w), self._make_out_collect(True)))
        self.servers: List[MessagingPS] = []
        for s in range(S):
            self.servers.append(MessagingPS(
                self.ps_senders[s], self._make_ps_collect(s), self._make_out_collect(False)))

    # -- collectors ------------------------------------------------------------------------
    def _make_w_collect(self, w: int):
        def collect(msg):
            key = self.param_partitioner(msg)
            shard = key % self.S  # the reference's final ``key % numPartitions``
            if shard < 0:
                raise RuntimeError("param_partitioner must return a non-negative key")
            with self._cv:
                self.w2ps[w][shard].append(msg)
                self.stats.worker_to_ps_msgs += 1
                self._cv.notify()
        return collect

    def _make_ps_collect(self, s: int):
        def collect(msg):
            key = self.w_in_partition(msg)
            if not 0 <= key < self.W:
                raise RuntimeError("Pull answer key should be the partition ID itself!")
            with self._cv:
                self.ps2w[s][key].append(msg)
                self.stats.ps_to_worker_msgs += 1
                self._cv.notify()
        return collect

    def _make_out_collect(self, is_worker: bool):
        def collect(out):
            with self._out_lock:
                if is_worker:
                    self.outputs.append(Left(out))
                    self.stats.worker_outputs += 1
                else:
                    self.outputs.append(Right(out))
                    self.stats.ps_outputs += 1
        return collect

    # -- event sources ---------------------------------------------------------------------
    def _fill_input(self, w: int) -> bool:
        if self._input_peek[w]:
            return True
        if self._input_done[w]:
            return False
        try:
            self._input_peek[w].append(next(self.inputs[w]))
            return True
        except StopIteration:
            self._input_done[w] = True
            return False

    def _ready(self):
        """All non-empty event sources as ``(kind, a, b)`` in a fixed canonical order."""
        ready_in, ready_msg = [], []
        for w in range(self.W):
            if self._fill_input(w):
                ready_in.append(("in", w, 0))
        for w in range(self.W):
            for s in range(self.S):
                if self.w2ps[w][s]:
                    ready_msg.append(("w2ps", w, s))
        for s in range(self.S):
            for w in range(self.W):
                if self.ps2w[s][w]:
                    ready_msg.append(("ps2w", s, w))
        return ready_in, ready_msg

    def _process(self, ev) -> None:
        kind, a, b = ev
        self.stats.events += 1
        if kind == "in":
            data = self._input_peek[a].popleft()
            self.stats.input_records += 1
            self.worker_logics[a].on_recv(data, self.clients[a])
        elif kind == "w2ps":
            with self._cv:
                msg = self.w2ps[a][b].popleft()
            logic, ps = self.ps_logics[b], self.servers[b]
            self.ps_receivers[b].on_worker_msg(
                msg,
                lambda pid, widx: logic.on_pull_recv(pid, widx, ps),
                lambda pid, delta: logic.on_push_recv(pid, delta, ps))
        else:
            with self._cv:
                msg = self.ps2w[a][b].popleft()
            logic, client = self.worker_logics[b], self.clients[b]
            self.worker_receivers[b].on_pull_answer_recv(
                msg, lambda ans: logic.on_pull_recv(ans.param_id, ans.param, client))

    def _pick(self, ready_in, ready_msg):
        sched = self.schedule
        if sched == "random":
            allr = ready_in + ready_msg
            return [allr[self.rng.randrange(len(allr))]]
        if sched == "input_first":
            return ready_in if ready_in else ready_msg
        if sched == "feedback_first":
            return ready_msg if ready_msg else ready_in[:1]
        # round_robin: one event from every non-empty source per sweep
        return ready_in + ready_msg

    # -- main loop -------------------------------------------------------------------------
    def _pump_until_quiescent(self) -> None:
        while True:
            ready_in, ready_msg = self._ready()
            if not ready_in and not ready_msg:
                if not self.realtime:
                    return
                # idle: wait for a background thread (timer / worker thread) to emit
                self.stats.idle_waits += 1
                deadline = None if self.iteration_wait_time == 0 else \
                    time.monotonic() + self.iteration_wait_time / 1000.0
                with self._cv:
                    while not self._any_msg():
                        remaining = None if deadline is None else deadline - time.monotonic()
                        if remaining is not None and remaining <= 0:
                            return
                        self._cv.wait(timeout=remaining if remaining is not None else 0.25)
                        if deadline is None and self._stop_requested:
                            return
                continue
            for ev in self._pick(ready_in, ready_msg):
                # a sweep's snapshot can go stale in random / threaded runs: re-check
                if ev[0] == "in" and not self._input_peek[ev[1]]:
                    continue
                if ev[0] == "w2ps" and not self.w2ps[ev[1]][ev[2]]:
                    continue
                if ev[0] == "ps2w" and not self.ps2w[ev[1]][ev[2]]:
                    continue
                self._process(ev)
                if self.max_events is not None and self.stats.events >= self.max_events:
                    raise RuntimeError(f"engine exceeded max_events={self.max_events}")

    _stop_requested = False

    def request_stop(self) -> None:
        """Ends an ``iteration_wait_time == 0`` ("run forever") job from another thread."""
        self._stop_requested = True
        with self._cv:
            self._cv.notify_all()

    def _any_msg(self) -> bool:
        return any(q for row in self.w2ps for q in row) or any(q for row in self.ps2w for q in row)

    def run(self) -> List[Any]:
        for s, logic in enumerate(self.ps_logics):
            logic.open({}, RuntimeContext(s, self.S, "ps"))
        for w, logic in enumerate(self.worker_logics):
            logic.open(RuntimeContext(w, self.W, "worker"))
        try:
            self._pump_until_quiescent()
            if self.flush_on_close:
                flushed = False
                for w, snd in enumerate(self.worker_senders):
                    if hasattr(snd, "flush") and snd.has_pending():
                        snd.flush(self.clients[w]._collect_msg)
                        flushed = True
                for s, snd in enumerate(self.ps_senders):
                    if hasattr(snd, "flush") and snd.has_pending():
                        snd.flush(self.servers[s]._collect_msg)
                        flushed = True
                while flushed:
                    was_rt, self.realtime = self.realtime, False
                    self._pump_until_quiescent()
                    self.realtime = was_rt
                    flushed = False
                    for s, snd in enumerate(self.ps_senders):
                        if hasattr(snd, "flush") and snd.has_pending():
                            snd.flush(self.servers[s]._collect_msg)
                            flushed = True
                    for w, snd in enumerate(self.worker_senders):
                        if hasattr(snd, "flush") and snd.has_pending():
                            snd.flush(self.clients[w]._collect_msg)
                            flushed = True
        finally:
            for snd in self.worker_senders + self.ps_senders:
                if hasattr(snd, "shutdown"):
                    snd.shutdown()
        for logic in self.worker_logics:
            logic.close()
        for s, logic in enumerate(self.ps_logics):
            logic.close(self.servers[s])
        return self.outputs
