"""BL2: the C++ CPU parameter server computes the same SGD as a serial numpy loop when run with
one worker, and stays sane (finite, loss decreasing) with several threads."""
import numpy as np

from fps_b200.parallel import cpu_baseline


def serial(u, i, r, U, V, lr, S):
    U, V = U.copy(), V.copy()
    sse = 0.0
    for a, b, y in zip(u, i, r):
        row = V[b % S, b // S]
        e = float(y) - U[0, a] @ row
        sse += e * e
        d = lr * e * U[0, a]
        U[0, a] += lr * e * row
        row += d
    return U, V, sse


def test_single_worker_pull_limit_1_equals_serial_sgd():
    rng = np.random.default_rng(0)
    dim, nu, ni, n, S = 8, 50, 40, 600, 3
    U = rng.uniform(0, 0.3, size=(1, nu, dim))
    V = rng.uniform(0, 0.3, size=(S, (ni + S - 1) // S, dim))
    u = rng.integers(0, nu, size=(1, n), dtype=np.int32)
    i = rng.integers(0, ni, size=(1, n), dtype=np.int32)
    r = rng.random((1, n), dtype=np.float32)
    U1, V1 = U.copy(), V.copy()
    _secs, sse = cpu_baseline.run_mf(u, i, r, U1, V1, 0.05, pull_limit=1)
    U2, V2, sse2 = serial(u[0], i[0], r[0], U, V, 0.05, S)
    np.testing.assert_allclose(U1, U2, rtol=1e-12, atol=1e-14)
    np.testing.assert_allclose(V1, V2, rtol=1e-12, atol=1e-14)
    assert abs(sse - sse2) < 1e-9


def test_multi_threaded_run_learns():
    rng = np.random.default_rng(1)
    W, S, dim, nu, ni, n = 3, 2, 6, 30, 20, 4000
    U = rng.uniform(0, 0.3, size=(W, nu, dim))
    V = rng.uniform(0, 0.3, size=(S, (ni + S - 1) // S, dim))
    u = rng.integers(0, nu, size=(W, n), dtype=np.int32)
    i = rng.integers(0, ni, size=(W, n), dtype=np.int32)
    r = np.full((W, n), 0.7, dtype=np.float32)
    _s, sse1 = cpu_baseline.run_mf(u, i, r, U, V, 0.05, pull_limit=64)
    _s, sse2 = cpu_baseline.run_mf(u, i, r, U, V, 0.05, pull_limit=64)
    assert np.isfinite(U).all() and np.isfinite(V).all() and sse2 < sse1
    res = cpu_baseline.bench_mf(2000, 1000, 16, 8000, steps=2, warmup=1, workers=2, shards=2)
    assert res["ratings_per_sec"] > 0
