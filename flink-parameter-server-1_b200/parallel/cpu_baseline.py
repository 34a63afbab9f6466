"""BL2 front-end: the multi-threaded C++ CPU parameter server (``csrc/cpu/ps_cpu.cpp``).

This is the *timed CPU baseline* that stands in for the reference's Flink mini-cluster run
(unavailable on this box) and a second oracle for the MF maths.  Always
report its numbers as "CPU restatement, not Flink".
"""
from __future__ import annotations

import ctypes as C
import pathlib

import numpy as np

_LIB = pathlib.Path(__file__).resolve().parent.parent / "ops" / "_lib" / "libfps_cpu.so"
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not _LIB.exists():
            import importlib.util
            import sys
            root = pathlib.Path(__file__).resolve().parents[2]
            spec = importlib.util.spec_from_file_location("build_native", root / "build_native.py")
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.build()
        L = C.CDLL(str(_LIB))
        L.fps_cpu_mf_run.restype = C.c_double
        L.fps_cpu_mf_run.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_double, C.c_int,
                                     C.POINTER(C.c_double)]
        L.fps_cpu_pa_binary_run.restype = C.c_double
        L.fps_cpu_pa_binary_run.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_int,
                                            C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.fps_cpu_pa_ova_run.restype = C.c_double
        L.fps_cpu_pa_ova_run.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_double, C.c_int,
                                         C.POINTER(C.c_double), C.POINTER(C.c_int64)]
        L.fps_cpu_fill_uniform.restype = None
        L.fps_cpu_fill_uniform.argtypes = [C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_uint64]
        _lib = L
    return _lib


def run_mf(users, items, ratings, user_table, item_table, lr: float, pull_limit: int = 1600):
    """``users`` / ``items`` / ``ratings``: ``[W, n]`` arrays (worker-local user slots, global item
    ids); ``user_table`` ``[W, U, dim]`` and ``item_table`` ``[S, I, dim]`` fp64, updated in place.
    Returns ``(seconds, sum of squared errors)``."""
    users = np.ascontiguousarray(users, dtype=np.int32)
    items = np.ascontiguousarray(items, dtype=np.int32)
    ratings = np.ascontiguousarray(ratings, dtype=np.float32)
    W, n = users.shape
    S, I, dim = item_table.shape
    assert user_table.shape[0] == W and user_table.shape[2] == dim
    assert user_table.flags.c_contiguous and item_table.flags.c_contiguous
    sse = C.c_double(0.0)
    secs = lib().fps_cpu_mf_run(W, S, dim, n, users.ctypes.data, items.ctypes.data, ratings.ctypes.data,
                                user_table.ctypes.data, user_table.shape[1], item_table.ctypes.data, I,
                                float(lr), int(pull_limit), C.byref(sse))
    return secs, sse.value


def run_pa_binary(indptr, indices, values, labels, weights, variant: str = "pa", C_: float = 1.0,
                  pull_limit: int = 10_000):
    """Binary passive-aggressive through the C++ PS loop (``PassiveAggressiveParameterServer.scala:283-338``
    semantics: one pull and one push per active feature, limiter counted in features, range-partitioned
    weights).  ``indptr`` ``[W, n + 1]`` (offsets local to each worker's block), ``indices`` / ``values``
    ``[W, nnz]``, ``labels`` ``[W, n]`` (+1 / -1), ``weights`` ``[S, div]`` fp64 updated in place
    (feature ``f`` lives at ``weights[f // div, f % div]``).  Returns ``(seconds, hinge loss sum, mistakes)``."""
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    values = np.ascontiguousarray(values, dtype=np.float32)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    W, n1 = indptr.shape
    S, div = weights.shape
    assert indices.shape == values.shape and indices.shape[0] == W and labels.shape == (W, n1 - 1)
    assert weights.flags.c_contiguous and weights.dtype == np.float64
    loss, wrong = C.c_double(0.0), C.c_int64(0)
    secs = lib().fps_cpu_pa_binary_run(W, S, n1 - 1, indices.shape[1], indptr.ctypes.data, indices.ctypes.data,
                                       values.ctypes.data, labels.ctypes.data, weights.ctypes.data, div,
                                       {"pa": 0, "pa_i": 1, "pa_ii": 2}[variant], float(C_), int(pull_limit),
                                       C.byref(loss), C.byref(wrong))
    return secs, loss.value, wrong.value


def run_pa_ova(indptr, indices, values, labels, weights, variant: str = "pa", C_: float = 1.0,
               pull_limit: int = 10_000):
    """One-vs-all multiclass passive-aggressive through the C++ PS loop
    (``PassiveAggressiveOneVersusAll.scala:52-69``): parameters are rows of ``L`` class weights, ``labels``
    ``[W, n]`` class indices, ``weights`` ``[S, div, L]`` fp64 updated in place.  Same array layout as
    ``run_pa_binary``.  Returns ``(seconds, summed hinge loss over classes, argmax mistakes)``."""
    indptr = np.ascontiguousarray(indptr, dtype=np.int64)
    indices = np.ascontiguousarray(indices, dtype=np.int32)
    values = np.ascontiguousarray(values, dtype=np.float32)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    W, n1 = indptr.shape
    S, div, L = weights.shape
    assert indices.shape == values.shape and indices.shape[0] == W and labels.shape == (W, n1 - 1)
    assert weights.flags.c_contiguous and weights.dtype == np.float64
    loss, wrong = C.c_double(0.0), C.c_int64(0)
    secs = lib().fps_cpu_pa_ova_run(W, S, L, n1 - 1, indices.shape[1], indptr.ctypes.data, indices.ctypes.data,
                                    values.ctypes.data, labels.ctypes.data, weights.ctypes.data, div,
                                    {"pa": 0, "pa_i": 1, "pa_ii": 2}[variant], float(C_), int(pull_limit),
                                    C.byref(loss), C.byref(wrong))
    return secs, loss.value, wrong.value


def bench_pa_ova(n_features: int = 1_000_000, n_classes: int = 100, nnz: int = 100, examples: int = 1 << 13,
                 workers: int = 4, shards: int = 4, variant: str = "pa_i", pull_limit: int = 10_000,
                 seed: int = 0) -> dict:
    """BASELINE config 4 on the CPU restatement (same shapes as ``benchmarks/run_configs.py`` config 4)."""
    rng = np.random.default_rng(seed)
    n = examples // workers
    div = (n_features + shards - 1) // shards
    idx = np.sort(rng.integers(0, n_features, size=(workers, n, nnz), dtype=np.int32), axis=2).reshape(workers, -1)
    val = rng.standard_normal((workers, n * nnz)).astype(np.float32)
    lab = rng.integers(0, n_classes, size=(workers, n)).astype(np.int32)
    indptr = np.tile(np.arange(0, (n + 1) * nnz, nnz, dtype=np.int64), (workers, 1))
    w = np.zeros((shards, div, n_classes), dtype=np.float64)
    run_pa_ova(indptr[:, : n // 8 + 1], idx[:, : (n // 8) * nnz], val[:, : (n // 8) * nnz], lab[:, : n // 8], w,
               variant, 0.1, pull_limit)
    secs, loss, wrong = run_pa_ova(indptr, idx, val, lab, w, variant, 0.1, pull_limit)
    return {"impl": "cpu", "note": "C++ CPU restatement of the PS loop (not Flink), fp64", "workers": workers,
            "shards": shards, "examples": n * workers, "seconds": secs, "examples_per_sec": n * workers / secs,
            "row_pulls_per_sec": n * workers * nnz / secs, "row_bytes": n_classes * 8}


def bench_pa_binary(n_features: int = 47_236, nnz: int = 75, examples: int = 1 << 17, workers: int = 4,
                    shards: int = 4, variant: str = "pa_i", pull_limit: int = 10_000, seed: int = 0) -> dict:
    """BASELINE config 3 on the CPU restatement (same shapes as ``benchmarks/run_configs.py`` config 3)."""
    rng = np.random.default_rng(seed)
    n = examples // workers
    div = (n_features + shards - 1) // shards
    idx = np.sort(rng.integers(0, n_features, size=(workers, n, nnz), dtype=np.int32), axis=2).reshape(workers, -1)
    val = rng.standard_normal((workers, n * nnz)).astype(np.float32)
    lab = (rng.integers(0, 2, size=(workers, n)) * 2 - 1).astype(np.int32)
    indptr = np.tile(np.arange(0, (n + 1) * nnz, nnz, dtype=np.int64), (workers, 1))
    w = np.zeros((shards, div), dtype=np.float64)
    run_pa_binary(indptr[:, : n // 8 + 1], idx[:, : (n // 8) * nnz], val[:, : (n // 8) * nnz], lab[:, : n // 8], w,
                  variant, 1.0, pull_limit)                                         # warm-up on a slice
    secs, loss, wrong = run_pa_binary(indptr, idx, val, lab, w, variant, 1.0, pull_limit)
    return {"impl": "cpu", "note": "C++ CPU restatement of the PS loop (not Flink), fp64", "workers": workers,
            "shards": shards, "examples": n * workers, "seconds": secs, "examples_per_sec": n * workers / secs,
            "feature_pulls_per_sec": n * workers * nnz / secs, "mean_hinge_loss": loss / (n * workers)}


def bench_mf(n_users: int, n_items: int, dim: int, batch: int, steps: int, warmup: int, workers: int,
             shards: int, lr: float = 0.01, pull_limit: int = 1600, seed: int = 0) -> dict:
    rng = np.random.default_rng(seed)
    upw = (n_users + workers - 1) // workers
    ips = (n_items + shards - 1) // shards
    ut = np.empty((workers, upw, dim), dtype=np.float64)
    it = np.empty((shards, ips, dim), dtype=np.float64)
    lib().fps_cpu_fill_uniform(ut.ctypes.data, ut.size, -0.01, 0.01, 1)
    lib().fps_cpu_fill_uniform(it.ctypes.data, it.size, -0.01, 0.01, 2)
    n = batch // workers

    def batch_arrays(k):
        u = rng.integers(0, upw, size=(workers, n), dtype=np.int32)
        i = (rng.random((workers, n)) ** 3 * n_items).astype(np.int32)
        r = rng.random((workers, n), dtype=np.float32)
        return u, i, r

    for k in range(warmup):
        run_mf(*batch_arrays(k), ut, it, lr, pull_limit)
    total = 0.0
    for k in range(steps):
        secs, _ = run_mf(*batch_arrays(k), ut, it, lr, pull_limit)
        total += secs
    return {"ratings_per_sec": n * workers * steps / total, "ms_per_step": total / steps * 1e3,
            "workers": workers, "shards": shards}
