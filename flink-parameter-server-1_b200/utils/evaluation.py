"""Evaluation sinks and metrics (the reference keeps these in its test tree).

Parity: ``T/matrix/factorization/sink/nDCGSink.scala:13-272`` (nDCG + hit-rate, optional
per-period breakdown, text / CSV output), ``T/test/utils/PassiveAggressiveBinaryModelEvaluation.scala:14-41``,
``T/test/utils/PassiveAggressiveMultiModelEvaluation.scala:13-26``, RMSE of
``T/matrix/factorization/PSOfflineMatrixFactorizationTest.scala:36-47``, and the
``SuccessException`` result-extraction trick of ``T/test/utils/FlinkTestUtils.scala:6-30``.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

LN2 = math.log(2.0)


def ndcg_of_rank(rank: int) -> float:
    """``ln2 / ln(1 + rank)`` for a 1-based rank; 0 if the item is not in the list."""
    return LN2 / math.log(1.0 + rank) if rank > 0 else 0.0


def rank_in_topk(item: int, topk: Sequence[Tuple[float, int]]) -> int:
    for pos, (_score, i) in enumerate(topk):
        if i == item:
            return pos + 1
    return 0


class NDCGSink:
    """Feed ``(user, item, timestamp, topK)`` records (the output of the recommend-then-learn
    job); ``period_length > 0`` additionally buckets by ``timestamp // period_length``."""

    def __init__(self, period_length: int = 0):
        self.period_length = period_length
        self.sum_ndcg = 0.0
        self.hits = 0
        self.count = 0
        self.periods: Dict[int, List[float]] = {}

    def invoke(self, record) -> None:
        _user, item, ts, topk = record
        rank = rank_in_topk(item, topk)
        nd = ndcg_of_rank(rank)
        self.sum_ndcg += nd
        self.hits += 1 if rank > 0 else 0
        self.count += 1
        if self.period_length > 0:
            p = self.periods.setdefault(int(ts) // self.period_length, [0.0, 0, 0])
            p[0] += nd
            p[1] += 1 if rank > 0 else 0
            p[2] += 1

    def consume(self, records: Iterable) -> "NDCGSink":
        for r in records:
            self.invoke(r)
        return self

    @property
    def ndcg(self) -> float:
        return self.sum_ndcg / self.count if self.count else 0.0

    @property
    def hit_rate(self) -> float:
        return self.hits / self.count if self.count else 0.0

    def per_period(self) -> List[Tuple[int, float, float, int]]:
        return [(p, v[0] / v[2], v[1] / v[2], v[2]) for p, v in sorted(self.periods.items())]

    def write_text(self, path: str) -> None:
        with open(path, "w") as f:
            f.write(f"nDCG: {self.ndcg}\nhits: {self.hits}\ncount: {self.count}\n")
            for p, nd, hr, n in self.per_period():
                f.write(f"period {p}: nDCG {nd} hit-rate {hr} count {n}\n")

    def write_csv(self, path: str) -> None:
        with open(path, "w") as f:
            f.write("period,ndcg,hit_rate,count\n")
            for p, nd, hr, n in self.per_period():
                f.write(f"{p},{nd},{hr},{n}\n")
            f.write(f"all,{self.ndcg},{self.hit_rate},{self.count}\n")


def rmse(ratings, users: Dict[int, np.ndarray], items: Dict[int, np.ndarray]) -> float:
    se, n = 0.0, 0
    for r in ratings:
        if r.user in users and r.item in items:
            se += (r.rating - float(np.dot(users[r.user], items[r.item]))) ** 2
            n += 1
    return math.sqrt(se / n) if n else float("nan")


def binary_accuracy(model: Dict[int, float], test: Iterable, method) -> float:
    """% of ``(vector, bool label)`` pairs the model classifies correctly."""
    tp = tn = fp = fn = 0
    for vec, label in test:
        pred = method.predict(vec, model)
        if pred and label:
            tp += 1
        elif not pred and not label:
            tn += 1
        elif pred:
            fp += 1
        else:
            fn += 1
    tot = tp + tn + fp + fn
    return 100.0 * (tp + tn) / tot if tot else 0.0


def multiclass_accuracy(model: Dict[int, np.ndarray], test: Iterable, method) -> float:
    hit = tot = 0
    for vec, label in test:
        hit += int(method.predict(vec, model) == label)
        tot += 1
    return 100.0 * hit / tot if tot else 0.0


class SuccessException(Exception):
    def __init__(self, content):
        super().__init__("success")
        self.content = content


class NoSuccessExceptionReceived(AssertionError):
    pass


def execute_with_success_check(job: Callable[[], object], check: Callable[[object], None]) -> None:
    """Run ``job``; it must raise ``SuccessException(result)`` (possibly wrapped as ``__cause__``)."""
    try:
        job()
    except Exception as e:  # noqa: BLE001
        root = e
        while not isinstance(root, SuccessException) and (root.__cause__ or root.__context__):
            root = root.__cause__ or root.__context__
        if isinstance(root, SuccessException):
            check(root.content)
            return
        raise
    raise NoSuccessExceptionReceived("job finished without a SuccessException")
