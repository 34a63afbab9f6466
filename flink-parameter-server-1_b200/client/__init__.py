"""Worker-side wire components.

Parity: ``M/client/sender/SimpleWorkerSender.scala:6-15``, ``CombinationWorkerSender.scala:9-36``,
``CountClientSender.scala:6-8``, ``TimerClientSender.scala:8-10``,
``M/client/receiver/SimpleWorkerReceiver.scala:6-13``, ``MultipleWorkerReceiver.scala:6-13``.
"""
from __future__ import annotations

from ..api import WorkerReceiver, WorkerSender
from ..common.combinable import CombinationLogic, CountLogic, TimerLogic
from ..entities import Left, Pull, Push, Right, WorkerToPS


class SimpleWorkerSender(WorkerSender):
    """One ``WorkerToPS`` per pull / push."""

    def on_pull(self, param_id, collect, partition_id):
        collect(WorkerToPS(partition_id, Left(Pull(param_id))))

    def on_push(self, param_id, delta_update, collect, partition_id):
        collect(WorkerToPS(partition_id, Right(Push(param_id, delta_update))))


class SimpleWorkerReceiver(WorkerReceiver):
    def on_pull_answer_recv(self, msg, pull_handler):
        pull_handler(msg.msg)


class MultipleWorkerReceiver(WorkerReceiver):
    """Wire type is a list of ``PSToWorker``."""

    def on_pull_answer_recv(self, msg, pull_handler):
        for m in msg:
            pull_handler(m.msg)


class CombinationWorkerSender(CombinationLogic, WorkerSender):
    """Batching sender: wire type is ``list[WorkerToPS]``."""

    def on_pull(self, param_id, collect, partition_id):
        self.logic(lambda buf: buf.append(WorkerToPS(partition_id, Left(Pull(param_id)))), collect)

    def on_push(self, param_id, delta_update, collect, partition_id):
        self.logic(lambda buf: buf.append(WorkerToPS(partition_id, Right(Push(param_id, delta_update)))),
                   collect)


def CountClientSender(max_count: int) -> CountLogic:  # noqa: N802 - mirrors the reference's case-class name
    return CountLogic(max_count)


def TimerClientSender(interval_seconds: float) -> TimerLogic:  # noqa: N802
    return TimerLogic(interval_seconds)


__all__ = ["SimpleWorkerSender", "SimpleWorkerReceiver", "MultipleWorkerReceiver",
           "CombinationWorkerSender", "CountClientSender", "TimerClientSender"]
