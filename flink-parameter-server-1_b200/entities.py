"""Wire messages and the ``Either`` output type.

Parity: ``M/entities/Messages.scala:3-8`` (``WorkerToPS``, ``PSToWorker``, ``Pull``, ``Push``,
``PullAnswer``) and Scala's ``Either`` as used for the job output
(``M/FlinkParameterServer.scala:325-328``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Generic, TypeVar, Union

P = TypeVar("P")
A = TypeVar("A")
B = TypeVar("B")


@dataclass(frozen=True)
class Left(Generic[A]):
    value: A

    is_left = True
    is_right = False


@dataclass(frozen=True)
class Right(Generic[B]):
    value: B

    is_left = False
    is_right = True


Either = Union[Left[A], Right[B]]


@dataclass(frozen=True)
class Pull:
    param_id: int


@dataclass(frozen=True)
class Push(Generic[P]):
    param_id: int
    delta: Any


@dataclass(frozen=True)
class PullAnswer(Generic[P]):
    param_id: int
    param: Any


@dataclass(frozen=True)
class WorkerToPS(Generic[P]):
    """A worker request; ``msg`` is ``Left(Pull)`` or ``Right(Push)``."""

    worker_partition_index: int
    msg: Any

    @property
    def param_id(self) -> int:
        return self.msg.value.param_id


@dataclass(frozen=True)
class PSToWorker(Generic[P]):
    worker_partition_index: int
    msg: PullAnswer


@dataclass(frozen=True)
class EOF:
    """End-of-input marker (``M/matrix/factorization/utils/Utils.scala:67``,
    ``M/utils/FlinkEOF.scala``)."""

    source: int = -1
    target: int = -1


def lefts(stream):
    return [e.value for e in stream if e.is_left]


def rights(stream):
    return [e.value for e in stream if e.is_right]


def last_write_wins(pairs) -> dict:
    """Fold ``(id, value)`` records the way every reference sink does
    (``T/matrix/factorization/PSOfflineMatrixFactorizationTest.scala:75-82``)."""
    out = {}
    for k, v in pairs:
        out[k] = v
    return out
