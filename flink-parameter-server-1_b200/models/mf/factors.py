"""Factor initialisers and the SGD updater (host, fp64 maths on fp32-representable inits).

Parity:
  * ``FactorInitializer`` + descriptor/``open()`` pattern - ``factors/FactorInitializer.scala:8-32``
  * ``RandomFactorInitializer`` (U[0,1))                  - ``factors/RandomFactorInitializer.scala:5-17``
  * ``RangedRandomFactorInitializer``                     - ``factors/RangedRandomFactorInitializer.scala:5-18``
  * ``PseudoRandomFactorInitializer`` (seeded by id)      - ``PseudoRandomFactorInitializer.scala:7-20``
  * ``FactorUpdater`` / ``SGDUpdater``                    - ``factors/FactorUpdater.scala:9-24``, ``SGDUpdater.scala:3-13``

Deliberate change: every random initialiser is counter-based and keyed by
``(seed, stream, id)`` (``ops/philox.py``) so host and device - and any two GPUs - agree.
``lambda_`` (L2 regularisation) is exposed with the reference's behaviour (none) as default (Q4).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import numpy as np

from ...ops import philox


class FactorInitializer:
    def next_factor(self, id_: int) -> np.ndarray:
        raise NotImplementedError

    def batch(self, ids) -> np.ndarray:
        return np.stack([self.next_factor(int(i)) for i in ids])


class FactorInitializerDescriptor:
    """Serialisable recipe; ``open()`` builds the live initialiser on the task."""

    def open(self) -> FactorInitializer:
        raise NotImplementedError


@dataclass(frozen=True)
class RangedRandomFactorInitializerDescriptor(FactorInitializerDescriptor):
    num_factors: int
    range_min: float
    range_max: float
    seed: int = 0
    stream: int = philox.STREAM_ITEM

    def open(self) -> "RangedRandomFactorInitializer":
        return RangedRandomFactorInitializer(self)


class RangedRandomFactorInitializer(FactorInitializer):
    def __init__(self, d: RangedRandomFactorInitializerDescriptor):
        self.d = d

    def next_factor(self, id_: int) -> np.ndarray:
        return self.batch([id_])[0]

    def batch(self, ids) -> np.ndarray:
        d = self.d
        return philox.ranged_uniform(ids, d.num_factors, d.range_min, d.range_max, d.seed,
                                     d.stream).astype(np.float64)


def RandomFactorInitializerDescriptor(num_factors: int, seed: int = 0,  # noqa: N802
                                      stream: int = philox.STREAM_ITEM):
    """U[0,1) - the initialiser of the reference's offline-MF acceptance test."""
    return RangedRandomFactorInitializerDescriptor(num_factors, 0.0, 1.0, seed, stream)


@dataclass(frozen=True)
class PseudoRandomFactorInitializerDescriptor(FactorInitializerDescriptor):
    """Deterministic per id, independent of any global seed (``new Random(id)`` upstream)."""
    num_factors: int

    def open(self):
        return RangedRandomFactorInitializer(
            RangedRandomFactorInitializerDescriptor(self.num_factors, 0.0, 1.0, seed=0x5EED,
                                                    stream=philox.STREAM_ITEM))


class FactorUpdater:
    def delta(self, rating: float, user: np.ndarray, item: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
        raise NotImplementedError


class SGDUpdater(FactorUpdater):
    """``e = r - u.v``; ``du = lr*(e*v - lambda*u)``; ``dv = lr*(e*u - lambda*v)``."""

    def __init__(self, learning_rate: float, lambda_: float = 0.0):
        self.learning_rate = float(learning_rate)
        self.lambda_ = float(lambda_)

    def delta(self, rating, user, item):
        user = np.asarray(user, dtype=np.float64)
        item = np.asarray(item, dtype=np.float64)
        e = rating - float(np.dot(user, item))
        lr, lam = self.learning_rate, self.lambda_
        if lam == 0.0:
            return lr * e * item, lr * e * user
        return lr * (e * item - lam * user), lr * (e * user - lam * item)
