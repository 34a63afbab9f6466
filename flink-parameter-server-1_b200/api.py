"""Public callback interfaces of the parameter server.

Parity map (reference file:line):
  * ``WorkerLogic``            - ``M/WorkerLogic.scala:23-58``
  * ``ParameterServerClient``  - ``M/ParameterServerClient.scala:12-20``
  * ``ParameterServerLogic``   - ``M/FlinkParameterServer.scala:891-926``
  * ``ParameterServer``        - ``M/FlinkParameterServer.scala:928-932``
  * ``PSReceiver`` / ``PSSender`` / ``WorkerReceiver`` / ``WorkerSender``
                               - ``M/FlinkParameterServer.scala:934-987``
  * ``RuntimeContext``         - the slice of Flink's RuntimeContext the reference reads
                                 (subtask index + parallelism, ``RangePSLogicWithClose.scala:51-62``)
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Generic, TypeVar

T = TypeVar("T")
P = TypeVar("P")
WOut = TypeVar("WOut")
PSOut = TypeVar("PSOut")


@dataclass
class RuntimeContext:
    index_of_this_subtask: int
    number_of_parallel_subtasks: int
    task_name: str = ""
    # free-form, e.g. {"device": torch.device, "rank": r} on the device engine
    extras: Dict[str, Any] = field(default_factory=dict)

    # Flink-style accessors so logics read the same either way
    def get_index_of_this_subtask(self) -> int:
        return self.index_of_this_subtask

    def get_number_of_parallel_subtasks(self) -> int:
        return self.number_of_parallel_subtasks


class ParameterServerClient(Generic[P, WOut]):
    """What a worker may do.  All three calls are fire-and-forget; the answer to a pull
    arrives later through ``WorkerLogic.on_pull_recv``."""

    def pull(self, param_id: int) -> None:
        raise NotImplementedError

    def push(self, param_id: int, delta_update: P) -> None:
        raise NotImplementedError

    def output(self, out: WOut) -> None:
        raise NotImplementedError


class WorkerLogic(Generic[T, P, WOut]):
    def open(self, runtime_context: RuntimeContext | None = None) -> None:
        """Called once per worker subtask before any record."""

    def on_recv(self, data: T, ps: ParameterServerClient) -> None:
        raise NotImplementedError

    def on_pull_recv(self, param_id: int, param_value: P, ps: ParameterServerClient) -> None:
        raise NotImplementedError

    def close(self) -> None:
        """Called once when the job terminates."""


class ParameterServer(Generic[P, PSOut]):
    """What server-side logic may do."""

    def answer_pull(self, param_id: int, value: P, worker_partition_index: int) -> None:
        raise NotImplementedError

    def output(self, out: PSOut) -> None:
        raise NotImplementedError


class ParameterServerLogic(Generic[P, PSOut]):
    def open(self, parameters: dict | None = None, runtime_context: RuntimeContext | None = None) -> None:
        pass

    def on_pull_recv(self, param_id: int, worker_partition_index: int, ps: ParameterServer) -> None:
        raise NotImplementedError

    def on_push_recv(self, param_id: int, delta_update: P, ps: ParameterServer) -> None:
        raise NotImplementedError

    def close(self, ps: ParameterServer) -> None:
        pass


# --- pluggable wire codecs -------------------------------------------------------------------

class PSReceiver:
    def on_worker_msg(self, msg, on_pull_recv: Callable[[int, int], None],
                      on_push_recv: Callable[[int, Any], None]) -> None:
        raise NotImplementedError


class PSSender:
    def on_pull_answer(self, param_id: int, value, worker_partition_index: int,
                       collect: Callable[[Any], None]) -> None:
        raise NotImplementedError


class WorkerReceiver:
    def on_pull_answer_recv(self, msg, pull_handler: Callable[[Any], None]) -> None:
        raise NotImplementedError


class WorkerSender:
    def on_pull(self, param_id: int, collect: Callable[[Any], None], partition_id: int) -> None:
        raise NotImplementedError

    def on_push(self, param_id: int, delta_update, collect: Callable[[Any], None],
                partition_id: int) -> None:
        raise NotImplementedError
