"""BL2: the C++ CPU parameter server computes the same SGD as a serial numpy loop when run with
one worker, and stays sane (finite, loss decreasing) with several threads."""
import numpy as np
import pytest

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


@pytest.mark.parametrize("variant", ["pa", "pa_i", "pa_ii"])
def test_cpu_pa_binary_equals_serial_rule_at_pull_limit_one(variant):
    """BL2 for passive-aggressive: with one worker and pull limit 1 the threaded PS loop is the serial
    algorithm (an example's pushes are on the channel before the next example's first pull)."""
    rng = np.random.default_rng(1)
    F, n, nnz, Cc = 50, 200, 6, 0.5
    idx = np.stack([np.sort(rng.choice(F, nnz, replace=False)) for _ in range(n)]).astype(np.int32)
    val = rng.standard_normal((n, nnz)).astype(np.float32)
    lab = np.where((val * rng.standard_normal(F)[idx]).sum(1) > 0, 1, -1).astype(np.int32)
    indptr = np.arange(0, (n + 1) * nnz, nnz, dtype=np.int64)[None]
    w = np.zeros((2, 25))                                     # two range-partitioned shards
    _secs, loss, wrong = cpu_baseline.run_pa_binary(indptr, idx.reshape(1, -1), val.reshape(1, -1), lab[None], w,
                                                    variant, Cc, 1)
    ref, want_loss, want_wrong = np.zeros(F), 0.0, 0
    for e in range(n):
        x = val[e].astype(np.float64)
        m, y = float(x @ ref[idx[e]]), float(lab[e])
        l, n2 = max(0.0, 1.0 - y * m), float(x @ x)
        want_loss += l
        want_wrong += (m > 0) != (y > 0)
        tau = {"pa": l / n2, "pa_i": min(Cc, l / n2), "pa_ii": l / (n2 + 1 / (2 * Cc))}[variant]
        ref[idx[e]] += tau * y * x
    assert np.allclose(w.reshape(-1), ref) and abs(loss - want_loss) < 1e-9 and wrong == want_wrong


def test_cpu_pa_binary_multi_worker_learns_and_respects_limit():
    rng = np.random.default_rng(2)
    F, W, n, nnz = 400, 3, 600, 10
    truth = rng.standard_normal(F)
    idx = np.sort(rng.integers(0, F, size=(W, n, nnz), dtype=np.int32), axis=2)
    val = rng.standard_normal((W, n, nnz)).astype(np.float32)
    lab = np.where((val * truth[idx]).sum(2) > 0, 1, -1).astype(np.int32)
    indptr = np.tile(np.arange(0, (n + 1) * nnz, nnz, dtype=np.int64), (W, 1))
    w = np.zeros((4, 100))
    _s, _loss, wrong_first = cpu_baseline.run_pa_binary(indptr, idx.reshape(W, -1), val.reshape(W, -1), lab, w, "pa_i", 1.0, 50)
    _s, _loss, wrong_again = cpu_baseline.run_pa_binary(indptr, idx.reshape(W, -1), val.reshape(W, -1), lab, w, "pa_i", 1.0, 50)
    assert wrong_again < 0.5 * wrong_first                       # second pass over the same data: far fewer mistakes


def test_cpu_ova_restatement_equals_the_serial_rule_at_limit_one():
    """One worker, pull limit 1: the threaded loop is serial, so it must reproduce the host algorithm."""
    from fps_b200.models.pa import SparseVector
    from fps_b200.models.pa.algorithm import PassiveAggressiveOneVersusAll
    from fps_b200.parallel import cpu_baseline
    rng = np.random.default_rng(4)
    F, L, n, nnz = 60, 5, 40, 6
    idx = np.sort(np.stack([rng.choice(F, nnz, replace=False) for _ in range(n)]), axis=1).astype(np.int32)
    val = rng.standard_normal((n, nnz)).astype(np.float32)
    lab = rng.integers(0, L, n).astype(np.int32)
    S = 3
    div = (F + S - 1) // S
    w = np.zeros((S, div, L))
    indptr = np.arange(0, (n + 1) * nnz, nnz, dtype=np.int64)[None, :]
    secs, loss, wrong = cpu_baseline.run_pa_ova(indptr, idx.reshape(1, -1), val.reshape(1, -1), lab[None, :], w,
                                                "pa_i", 0.5, pull_limit=1)
    algo = PassiveAggressiveOneVersusAll.build_pa_i(L, 0.5)
    model = {}
    mistakes = 0
    for e in range(n):
        x = SparseVector(idx[e], val[e].astype(np.float64), F)
        sub = {int(k): model.get(int(k), np.zeros(L)) for k in idx[e]}
        mistakes += int(algo.predict(x, sub) != lab[e])
        for k, d in algo.delta(x, sub, int(lab[e])):
            model[k] = model.get(k, np.zeros(L)) + d
    flat = w.reshape(-1, L)
    for k, v in model.items():
        np.testing.assert_allclose(flat[k], v, rtol=1e-9, atol=1e-9)
    assert wrong == mistakes and secs > 0
