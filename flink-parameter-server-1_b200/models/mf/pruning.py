"""LEMP pruning (Teflioudi et al.) for maximum-inner-product top-K.

Parity: ``pruning/LEMPPruningFunctions.scala:10-90`` (``lengthPruning``, ``coordPruning``,
``incrPruning``) and ``pruning/LEMPPruningStrategy.scala:6-76`` (ADT + ``fromString``).

Every predicate answers "may this item still beat theta?" - ``True`` keeps the candidate.
They are *safe* filters: an item whose score exceeds theta is never dropped, so top-K results
are exact whichever strategy is chosen.  Deviations (documented in SURVEY Q9/Q10):
``length_pruning`` here takes the bound on the *length* (``||p|| >= theta/||q||``), the reference
call site compares the squared length with the unsquared bound; the INCR focus set may use all
``k`` coordinates (the reference never picks the last one).
"""
from __future__ import annotations

import math
import re
from dataclasses import dataclass
from typing import Callable, Sequence, Tuple

import numpy as np

Item = Tuple[int, Tuple[float, np.ndarray]]   # (id, (length, vector))


def length_pruning(min_length: float) -> Callable[[Item], bool]:
    return lambda p: p[1][0] >= min_length


def coord_pruning(f: int, user: Tuple[float, np.ndarray], theta_b_q: float) -> Callable[[Item], bool]:
    """Feasible interval [L_f, U_f] of the normalised focus coordinate."""
    qlen, q = user
    q_bar_f = q[f] / qlen
    t = min(max(theta_b_q, -1.0), 1.0)
    a = q_bar_f * t
    b = math.sqrt(max(0.0, (1 - t * t) * (1 - q_bar_f * q_bar_f)))
    lf_p, uf_p = a - b, a + b
    l_f = lf_p if (q_bar_f >= 0 or lf_p > t / q_bar_f) else -1.0
    u_f = uf_p if (q_bar_f <= 0 or uf_p < t / q_bar_f) else 1.0

    def keep(p: Item) -> bool:
        plen, pv = p[1]
        if plen == 0.0:
            return theta_b_q <= 0.0
        p_bar_f = pv[f] / plen
        return l_f <= p_bar_f <= u_f
    return keep


def incr_pruning(F: Sequence[int], user: Tuple[float, np.ndarray], theta: float) -> Callable[[Item], bool]:
    qlen, q = user
    F = list(F)
    qF = q[F] if F else np.zeros(0)
    q_mF_sqr = qlen * qlen - float(np.dot(qF, qF))

    def keep(p: Item) -> bool:
        plen, pv = p[1]
        pF = pv[F] if F else np.zeros(0)
        u_bound = theta - float(np.dot(qF, pF))
        return u_bound < 0.0 or q_mF_sqr * (plen * plen - float(np.dot(pF, pF))) >= u_bound * u_bound
    return keep


class LEMPPruningStrategy:
    num_focus = 0


@dataclass(frozen=True)
class LENGTH(LEMPPruningStrategy):
    pass


@dataclass(frozen=True)
class COORD(LEMPPruningStrategy):
    pass


@dataclass(frozen=True)
class INCR(LEMPPruningStrategy):
    num_focus_coordinates: int

    @property
    def num_focus(self):
        return self.num_focus_coordinates


@dataclass(frozen=True)
class LC(LEMPPruningStrategy):
    algorithm_switch_threshold: float


@dataclass(frozen=True)
class LI(LEMPPruningStrategy):
    num_focus_coordinates: int
    algorithm_switch_threshold: float

    @property
    def num_focus(self):
        return self.num_focus_coordinates


_RX = [
    (re.compile(r"length"), lambda m: LENGTH()),
    (re.compile(r"coord"), lambda m: COORD()),
    (re.compile(r"incr:(\d+)"), lambda m: INCR(int(m.group(1)))),
    (re.compile(r"lc:([0-9.]+)"), lambda m: LC(float(m.group(1)))),
    (re.compile(r"li:(\d+):([0-9.]+)"), lambda m: LI(int(m.group(1)), float(m.group(2)))),
]


def from_string(s: str) -> LEMPPruningStrategy:
    for rx, mk in _RX:
        m = rx.fullmatch(s)
        if m:
            return mk(m)
    raise ValueError(f"Invalid LEMP Pruning strategy string {s}")


def focus_coordinate(q: np.ndarray) -> int:
    return int(np.argmax(q * q))


def focus_set(q: np.ndarray, n: int) -> list:
    if n <= 0:
        return []
    return list(np.argsort(-(q * q), kind="stable")[:n])


def bucket_filter(strategy: LEMPPruningStrategy, user: Tuple[float, np.ndarray], theta: float,
                  bucket_max_len: float, bucket_min_len: float, focus: int, fset: Sequence[int]):
    """Pick the per-bucket predicate exactly like ``PSTopKGeneratorWorker.scala:82-96``."""
    qlen = user[0]
    denom = bucket_max_len * qlen
    theta_b_q = theta / denom if denom > 0 else 0.0
    min_len = theta / qlen if qlen > 0 else 0.0
    if isinstance(strategy, LENGTH):
        return length_pruning(min_len)
    if isinstance(strategy, COORD):
        return coord_pruning(focus, user, theta_b_q)
    if isinstance(strategy, INCR):
        return incr_pruning(fset, user, theta)
    if isinstance(strategy, LC):
        if bucket_max_len > bucket_min_len * strategy.algorithm_switch_threshold:
            return length_pruning(min_len)
        return coord_pruning(focus, user, theta_b_q)
    if isinstance(strategy, LI):
        if bucket_max_len > bucket_min_len * strategy.algorithm_switch_threshold:
            return length_pruning(min_len)
        return incr_pruning(fset, user, theta)
    raise TypeError(strategy)
