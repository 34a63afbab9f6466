"""PS-side logics and wire components.

Parity: ``M/server/*.scala``, ``M/server/sender/*.scala``, ``M/server/receiver/*.scala``.
"""
from __future__ import annotations

from ..api import PSReceiver, PSSender
from ..common.combinable import CombinationLogic, CountLogic, TimerLogic
from ..entities import PSToWorker, PullAnswer
from .logics import (LockPSLogicA, LockPSLogicB, RangePSLogicWithClose, SimplePSLogic,
                     SimplePSLogicWithClose, range_partition_div)


class SimplePSReceiver(PSReceiver):
    """``M/server/receiver/SimplePSReceiver.scala:6-22``"""

    def on_worker_msg(self, msg, on_pull_recv, on_push_recv):
        inner = msg.msg
        if inner.is_left:
            on_pull_recv(inner.value.param_id, msg.worker_partition_index)
        else:
            on_push_recv(inner.value.param_id, inner.value.delta)


class MultiplePSReceiver(PSReceiver):
    """``M/server/receiver/MultiplePSReceiver.scala:6-26`` - wire type is a list."""

    _one = SimplePSReceiver()

    def on_worker_msg(self, msg, on_pull_recv, on_push_recv):
        for m in msg:
            self._one.on_worker_msg(m, on_pull_recv, on_push_recv)


class SimplePSSender(PSSender):
    """``M/server/sender/SimplePSSender.scala:6-12``"""

    def on_pull_answer(self, param_id, value, worker_partition_index, collect):
        collect(PSToWorker(worker_partition_index, PullAnswer(param_id, value)))


class CombinationPSSender(CombinationLogic, PSSender):
    """``M/server/sender/CombinationPSSender.scala:9-24`` - wire type ``list[PSToWorker]``."""

    def on_pull_answer(self, param_id, value, worker_partition_index, collect):
        self.logic(lambda buf: buf.append(PSToWorker(worker_partition_index, PullAnswer(param_id, value))),
                   collect)


def CountPSSender(max_count: int) -> CountLogic:  # noqa: N802
    return CountLogic(max_count)


def TimerPSSender(interval_seconds: float) -> TimerLogic:  # noqa: N802
    return TimerLogic(interval_seconds)


__all__ = ["SimplePSLogic", "SimplePSLogicWithClose", "RangePSLogicWithClose", "LockPSLogicA",
           "LockPSLogicB", "range_partition_div", "SimplePSReceiver", "MultiplePSReceiver",
           "SimplePSSender", "CombinationPSSender", "CountPSSender", "TimerPSSender"]
