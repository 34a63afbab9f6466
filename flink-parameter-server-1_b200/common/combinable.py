"""Message batching triggers.

Parity: ``M/common/Combinable.scala:5-27``, ``CountLogic.scala:5-29``, ``TimerLogic.scala:6-51``,
``CombinationLogic.scala:6-35``.

Kept on purpose: the *sticky* send flag - once a combinable fired, ``should_send``
stays true until the combination actually flushes - and ``CountLogic`` zeroing its counter even
when the combined condition refused to flush.  ``T/SenderReceiverTest.scala:193-204`` (the AND
case: one pull plus ``count_limit - 1`` pushes delivered) depends on both.

Fixed on purpose: the reference timer thread never stops and mutates the batch
buffer without synchronisation.  Here every ``CombinationLogic`` owns one lock shared with its
combinables, and timers are daemon threads that ``shutdown()`` joins.
"""
from __future__ import annotations

import threading
from typing import Callable, Generic, List, Optional, Sequence, TypeVar

W = TypeVar("W")


class Combinable(Generic[W]):
    def __init__(self):
        self._send = False
        self._lock: threading.RLock = threading.RLock()  # replaced by the owning CombinationLogic

    def send_condition(self) -> bool:
        raise NotImplementedError

    def logic(self, adder, callback: Callable[[Callable], None], collect: Callable[[List[W]], None]) -> None:
        raise NotImplementedError

    # state logic
    def should_send(self) -> bool:
        return self._send

    def send(self, callback: Callable[[Callable], None], collect: Callable[[List[W]], None]) -> None:
        self._send = True
        callback(collect)

    def reset(self) -> None:
        self._send = False

    def shutdown(self) -> None:
        pass

    needs_realtime = False


class CountLogic(Combinable[W]):
    def __init__(self, max_count: int):
        super().__init__()
        self.max = int(max_count)
        self.count = 0

    def clone(self) -> "CountLogic":
        return CountLogic(self.max)

    def send_condition(self) -> bool:
        return self.count >= self.max

    def logic(self, adder, callback, collect) -> None:
        self.count += 1
        if self.send_condition():
            self.send(callback, collect)
            self.count = 0


class TimerLogic(Combinable[W]):
    """Fires every ``interval`` seconds if at least one message arrived since the last firing."""

    needs_realtime = True

    def __init__(self, interval_seconds: float):
        super().__init__()
        self.interval = float(interval_seconds)
        self.contains_data = False
        self._thread: Optional[threading.Thread] = None
        self._stop = threading.Event()

    def clone(self) -> "TimerLogic":
        return TimerLogic(self.interval)

    def send_condition(self) -> bool:
        return self.contains_data

    def _run(self, callback, collect) -> None:
        while not self._stop.wait(self.interval):
            with self._lock:
                if self.send_condition():
                    self.send(callback, collect)
                    self.contains_data = False

    def logic(self, adder, callback, collect) -> None:
        self.contains_data = True
        if self._thread is None:
            self._stop.clear()
            self._thread = threading.Thread(target=self._run, args=(callback, collect), daemon=True)
            self._thread.start()

    def shutdown(self) -> None:
        self._stop.set()
        t, self._thread = self._thread, None
        if t is not None and t is not threading.current_thread():
            t.join(timeout=self.interval + 1.0)


class CombinationLogic(Generic[W]):
    """Shared pending-message buffer poked by every combinable; flushes the whole buffer as one
    list when ``condition(combinables)`` holds."""

    def __init__(self, condition: Callable[[Sequence[Combinable[W]]], bool],
                 combinables: Sequence[Combinable[W]]):
        self.condition = condition
        self.combinables = list(combinables)
        self.data: List[W] = []
        self._lock = threading.RLock()
        for c in self.combinables:
            c._lock = self._lock

    def clone(self):
        """A fresh, empty copy for another subtask (locks and timer threads are not copyable)."""
        return type(self)(self.condition, [c.clone() for c in self.combinables])

    @property
    def needs_realtime(self) -> bool:
        return any(c.needs_realtime for c in self.combinables)

    def check_and_send(self, collect: Callable[[List[W]], None]) -> None:
        with self._lock:
            if self.condition(self.combinables):
                batch, self.data = self.data, []
                collect(batch)
                for c in self.combinables:
                    c.reset()

    def logic(self, func: Callable[[List[W]], None], collect: Callable[[List[W]], None]) -> None:
        with self._lock:
            func(self.data)
            for c in self.combinables:
                c.logic(func, self.check_and_send, collect)

    def has_pending(self) -> bool:
        with self._lock:
            return bool(self.data)

    def flush(self, collect: Callable[[List[W]], None]) -> None:
        """Emit whatever is buffered regardless of the condition (used at job shutdown so a
        finite job never strands messages; no reference counterpart - the reference relies on
        the timer eventually firing)."""
        with self._lock:
            if self.data:
                batch, self.data = self.data, []
                collect(batch)
            for c in self.combinables:
                c.reset()

    def shutdown(self) -> None:
        for c in self.combinables:
            c.shutdown()


def any_should_send(combinables) -> bool:
    return any(c.should_send() for c in combinables)


def all_should_send(combinables) -> bool:
    return all(c.should_send() for c in combinables)
