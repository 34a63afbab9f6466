"""Checkpoint / resume through the output stream.

The reference has no checkpoint subsystem; its de-facto mechanism is the job output
(``Left((user, vec))`` / ``Right((id, P))`` records) written as text, one coordinate per
line ``id;value`` (``T/matrix/factorization/PSOnlineMatrixFactorizationImplicitTest.scala:77-91``), and
fed back through ``transformWithModelLoad`` / ``transformWithDoubleModelLoad``.  This module writes
and reads that text format, plus a compact ``.npz`` form for large device dumps, and turns either
back into the model stream both engines accept.
"""
from __future__ import annotations

from typing import Dict, Iterable, Iterator, List, Tuple

import numpy as np

from ..entities import Left, Right, last_write_wins


def fold(stream: Iterable) -> Tuple[Dict[int, np.ndarray], Dict[int, np.ndarray]]:
    """Last-write-wins per id for both halves of an output stream."""
    left_values, right_values = [], []
    for e in stream:                      # one pass: `stream` may be a generator
        (left_values if e.is_left else right_values).append(e.value)
    return last_write_wins(left_values), last_write_wins(right_values)


def write_id_value(path: str, vectors: Dict[int, Iterable[float]]) -> None:
    """The reference drivers' format: one ``id;value`` line per coordinate."""
    with open(path, "w") as f:
        for k, v in vectors.items():
            for x in np.atleast_1d(np.asarray(v, dtype=np.float64)):
                f.write(f"{k};{repr(float(x))}\n")


def read_id_value(path: str) -> Dict[int, np.ndarray]:
    acc: Dict[int, List[float]] = {}
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line:
                continue
            k, x = line.split(";")
            acc.setdefault(int(k), []).append(float(x))
    return {k: np.asarray(v, dtype=np.float64) for k, v in acc.items()}


def save_stream(stream: Iterable, left_path: str, right_path: str) -> Tuple[int, int]:
    """Write the folded output stream as two ``id;value`` files (worker half, server half)."""
    lefts, rights = fold(list(stream))
    write_id_value(left_path, lefts)
    write_id_value(right_path, rights)
    return len(lefts), len(rights)


def load_stream(left_path: str, right_path: str) -> list:
    """Model stream for ``transform_with_double_model_load`` / ``DeviceMF.load``:
    ``Left((id, vec))`` records then ``Right((id, vec))`` records."""
    out = [Left((k, v)) for k, v in read_id_value(left_path).items()]
    out += [Right((k, v)) for k, v in read_id_value(right_path).items()]
    return out


def save_npz(path: str, stream: Iterable) -> None:
    """Binary form: ids and dense row matrices of both halves."""
    lefts, rights = fold(list(stream))

    def pack(d):
        ids = np.fromiter(d.keys(), dtype=np.int64, count=len(d))
        rows = np.stack([np.atleast_1d(np.asarray(v, dtype=np.float32)) for v in d.values()]) if d else np.zeros((0, 0), np.float32)
        return ids, rows
    li, lr = pack(lefts)
    ri, rr = pack(rights)
    np.savez_compressed(path, left_ids=li, left_rows=lr, right_ids=ri, right_rows=rr)


def load_npz(path: str) -> list:
    z = np.load(path)
    out = [Left((int(k), v)) for k, v in zip(z["left_ids"], z["left_rows"])]
    out += [Right((int(k), v)) for k, v in zip(z["right_ids"], z["right_rows"])]
    return out
