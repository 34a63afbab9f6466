"""Dense-vector helpers for matrix factorisation (host, fp64).

Parity: ``M/matrix/factorization/utils/Vector.scala:3-100`` - ``vectorLengthSqr``, ``dotProduct``,
``vectorSum`` (raises on NaN, ``:78-80``), ``attachLength``.
"""
from __future__ import annotations

import math
from typing import Tuple

import numpy as np

Vector = np.ndarray
LengthAndVector = Tuple[float, np.ndarray]


class FactorIsNotANumberException(ArithmeticError):
    pass


def vector_length_sqr(u) -> float:
    u = np.asarray(u, dtype=np.float64)
    return float(np.dot(u, u))


def dot_product(u, v) -> float:
    return float(np.dot(np.asarray(u, dtype=np.float64), np.asarray(v, dtype=np.float64)))


def vector_sum(u, v) -> np.ndarray:
    r = np.asarray(u, dtype=np.float64) + np.asarray(v, dtype=np.float64)
    if np.isnan(r).any():
        raise FactorIsNotANumberException("factor became NaN")
    return r


def attach_length(u) -> LengthAndVector:
    u = np.asarray(u, dtype=np.float64)
    return (math.sqrt(vector_length_sqr(u)), u)
