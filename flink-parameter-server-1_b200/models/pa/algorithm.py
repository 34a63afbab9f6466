"""Passive-aggressive update rules (Crammer et al. 2006), host fp64.

Parity:
  * ``PassiveAggressiveAlgorithm``            - ``algorithm/PassiveAggressiveAlgorithm.scala:8-16``
  * binary PA / PA-I / PA-II                  - ``algorithm/PassiveAggressiveBinaryAlgorithm.scala:11-112``
  * ``PassiveAggressiveMulticlassAlgorithm``  - ``algorithm/PassiveAggressiveMulticlassAlgorithm.scala:10-24``
  * one-vs-all PA / PA-I / PA-II              - ``algorithm/PassiveAggressiveOneVersusAll.scala:14-123``
  * cost-based PB / ML                        - ``algorithm/PassiveAggressiveCostBased.scala:13-140``
    (paper-correct two-column update; the reference's never-cleared VectorBuilder is not
    reproduced)
  * initialisers                              - ``PassiveAggressiveParameterInitializer.scala:7-11``,
                                                ``RandomModelInitializer.scala:3-5``
  * legacy filters PAF / PAF-I / PAF-II       - ``algorithm/binary/PassiveAggressiveFilter.scala:7-49``
    (with the Breeze-path formulas: real division in PAF-II) and the abstract
    multiclass stub ``algorithm/multi/PassiveAggressiveFilter.scala:7-56``.

A *model* handed to ``delta`` / ``predict`` is restricted to the example's active features:
binary -> ``{feature: w}``; multiclass -> ``{feature: np.ndarray[label_count]}``.
"""
from __future__ import annotations

import math
from typing import Callable, Dict, Iterable, List, Tuple

import numpy as np

from .sparse import LegacySparseVector, SparseVector


class PassiveAggressiveAlgorithm:
    def delta(self, data_point: SparseVector, model, label) -> Iterable[Tuple[int, object]]:
        raise NotImplementedError

    def predict(self, data_point: SparseVector, model):
        raise NotImplementedError


def _margin(x: SparseVector, model: Dict[int, float]) -> float:
    return float(sum(v * model.get(i, 0.0) for i, v in x.active_iterator()))


class PassiveAggressiveBinaryAlgorithm(PassiveAggressiveAlgorithm):
    """Param = one float per feature, label = bool."""

    VARIANT = 0

    def __init__(self, aggressiveness: float = 0.0):
        self.aggressiveness = float(aggressiveness)

    @staticmethod
    def build_pa():
        return PassiveAggressiveBinaryAlgorithm(0.0)

    @staticmethod
    def build_pa_i(aggressiveness: float):
        return PassiveAggressiveBinaryAlgorithmI(aggressiveness)

    @staticmethod
    def build_pa_ii(aggressiveness: float):
        return PassiveAggressiveBinaryAlgorithmII(aggressiveness)

    def _quotient(self, x: SparseVector, loss: float, denom_const: float) -> float:
        n2 = x.norm_sqr() + denom_const
        return loss / n2 if n2 > 0 else 0.0

    def tau(self, x: SparseVector, loss: float) -> float:
        return self._quotient(x, loss, 0.0)

    def delta(self, x: SparseVector, model, label: bool):
        y = 1 if label else -1
        loss = max(0.0, 1.0 - y * _margin(x, model))
        mult = self.tau(x, loss) * y
        if mult == 0.0:
            return []
        return [(i, v * mult) for i, v in x.active_iterator()]

    def predict(self, x: SparseVector, model) -> bool:
        return _margin(x, model) > 0


class PassiveAggressiveBinaryAlgorithmI(PassiveAggressiveBinaryAlgorithm):
    VARIANT = 1

    def tau(self, x, loss):
        return min(self.aggressiveness, self._quotient(x, loss, 0.0))


class PassiveAggressiveBinaryAlgorithmII(PassiveAggressiveBinaryAlgorithm):
    VARIANT = 2

    def tau(self, x, loss):
        return self._quotient(x, loss, 1.0 / (2.0 * self.aggressiveness))


class PassiveAggressiveMulticlassAlgorithm(PassiveAggressiveAlgorithm):
    """Param = dense ``label_count`` vector per feature, label = class index."""

    label_count: int = 0

    def decision(self, x: SparseVector, model: Dict[int, np.ndarray]) -> np.ndarray:
        d = None
        for i, v in x.active_iterator():
            row = model.get(i)
            if row is not None:
                d = v * np.asarray(row, dtype=np.float64) if d is None else d + v * np.asarray(row, dtype=np.float64)
        return d if d is not None else np.zeros(self.label_count)

    def delta_mtx(self, x: SparseVector, model, label: int) -> List[Tuple[int, np.ndarray]]:
        raise NotImplementedError

    def delta(self, x, model, label):
        return self.delta_mtx(x, model, label)

    def predict(self, x: SparseVector, model) -> int:
        return int(np.argmax(self.decision(x, model)))


class PassiveAggressiveOneVersusAll(PassiveAggressiveMulticlassAlgorithm):
    VARIANT = 0

    def __init__(self, label_count: int, aggressiveness: float = 0.0):
        self.label_count, self.aggressiveness = int(label_count), float(aggressiveness)

    @staticmethod
    def build_pa(label_count: int):
        return PassiveAggressiveOneVersusAll(label_count)

    @staticmethod
    def build_pa_i(label_count: int, aggressiveness: float):
        return PassiveAggressiveOneVersusAllI(label_count, aggressiveness)

    @staticmethod
    def build_pa_ii(label_count: int, aggressiveness: float):
        return PassiveAggressiveOneVersusAllII(label_count, aggressiveness)

    def tau(self, norm_sqr: float, loss: np.ndarray) -> np.ndarray:
        return loss / norm_sqr

    def delta_mtx(self, x, model, label):
        y = -np.ones(self.label_count)
        y[label] = 1.0
        loss = np.maximum(0.0, 1.0 - self.decision(x, model) * y)
        mult = self.tau(x.norm_sqr(), loss) * y
        if not mult.any():
            return []
        return [(i, v * mult) for i, v in x.active_iterator()]


class PassiveAggressiveOneVersusAllI(PassiveAggressiveOneVersusAll):
    VARIANT = 1

    def tau(self, norm_sqr, loss):
        return np.minimum(self.aggressiveness, loss / norm_sqr)


class PassiveAggressiveOneVersusAllII(PassiveAggressiveOneVersusAll):
    VARIANT = 2

    def tau(self, norm_sqr, loss):
        return loss / (norm_sqr + 1.0 / (2.0 * self.aggressiveness))


class PassiveAggressiveCostBased(PassiveAggressiveMulticlassAlgorithm):
    """Cost-sensitive multiclass PA.  ``mode`` 'PB' (prediction-based) or 'ML' (max-loss)."""

    def __init__(self, cost: Callable[[int, int], float], label_count: int, mode: str = "PB"):
        self.cost, self.label_count, self.mode = cost, int(label_count), mode

    @staticmethod
    def build_pb(cost, label_count):
        return PassiveAggressiveCostBased(cost, label_count, "PB")

    @staticmethod
    def build_ml(cost, label_count):
        return PassiveAggressiveCostBased(cost, label_count, "ML")

    def quotient(self, d: np.ndarray, label: int) -> int:
        if self.mode == "PB":
            return int(np.argmax(d))
        c = np.array([math.sqrt(self.cost(label, i)) for i in range(self.label_count)])
        return int(np.argmax(d - d[label] + c))

    def delta_mtx(self, x, model, label):
        d = self.decision(x, model)
        q = self.quotient(d, label)
        if q == label:
            return []
        loss = d[q] - d[label] + math.sqrt(self.cost(label, q))
        t = loss / (2.0 * x.norm_sqr())
        out = []
        for i, v in x.active_iterator():
            row = np.zeros(self.label_count)
            row[label] = t * v
            row[q] = -t * v
            out.append((i, row))
        return out


def zero_one_cost(y: int, q: int) -> float:
    return 0.0 if y == q else 1.0


# --- initialisers ------------------------------------------------------------------------------

def binary_initializer(_feature_id: int) -> float:
    return 0.0


def multiclass_initializer(label_count: int) -> Callable[[int], np.ndarray]:
    return lambda _feature_id: np.zeros(label_count)


class RandomModelInitializer:
    @staticmethod
    def init() -> float:
        return 0.0


# --- legacy map-vector filters ---------------------------------------------------------------

class PassiveAggressiveFilter:
    """Binary PA over ``LegacySparseVector``; label in {+1, -1}."""

    def __init__(self, C: float = 1.0):
        self.C = float(C)

    def tau(self, norm_sqr: float, loss: float) -> float:
        return loss / norm_sqr

    def delta(self, data: LegacySparseVector, model: Dict[int, float], label: int) -> Dict[int, float]:
        vals = data.get_values()
        margin = sum(v * model.get(k, 0.0) for k, v in vals.items())
        loss = max(0.0, 1.0 - label * margin)
        n2 = sum(v * v for v in vals.values())
        t = self.tau(n2, loss) if n2 > 0 else 0.0
        return {k: t * label * v for k, v in vals.items()}

    def predict(self, data: LegacySparseVector, model: Dict[int, float]) -> int:
        return 1 if sum(v * model.get(k, 0.0) for k, v in data.get_values().items()) > 0 else -1


class PassiveAggressiveFilterI(PassiveAggressiveFilter):
    def tau(self, n2, loss):
        return min(self.C, loss / n2)


class PassiveAggressiveFilterII(PassiveAggressiveFilter):
    def tau(self, n2, loss):
        return loss / (n2 + 1.0 / (2.0 * self.C))


class PassiveAggressiveMultiFilter:
    """Abstract upstream (implementations commented out); the working multiclass rules are
    ``PassiveAggressiveOneVersusAll`` / ``PassiveAggressiveCostBased``."""

    def delta(self, data, model, label):
        raise NotImplementedError

    def predict(self, data, model):
        raise NotImplementedError
