"""GPU tests of the passive-aggressive kernels against the fp64 host rules."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from fps_b200 import Left, Right, last_write_wins, lefts, rights  # noqa: E402
from fps_b200.models.pa import (PassiveAggressiveBinaryAlgorithm, PassiveAggressiveCostBased,  # noqa: E402
                                PassiveAggressiveOneVersusAll, SparseVector, zero_one_cost)
from fps_b200.utils.evaluation import binary_accuracy  # noqa: E402


def make_binary(n=80, features=5000, nnz=120, seed=50):
    rnd = np.random.default_rng(seed)
    w_true = rnd.normal(size=features)
    data = []
    for _ in range(n):
        idx = rnd.choice(features, nnz, replace=False)
        val = rnd.normal(size=nnz).astype(np.float32).astype(np.float64)
        x = SparseVector(idx, val, features)
        data.append((x, bool(x.dot(w_true) > 0)))
    return data


def make_multi(n=60, features=300, nnz=25, classes=7, seed=3):
    rnd = np.random.default_rng(seed)
    W = rnd.normal(size=(features, classes))
    data = []
    for _ in range(n):
        idx = rnd.choice(features, nnz, replace=False)
        val = rnd.normal(size=nnz).astype(np.float32).astype(np.float64)
        x = SparseVector(idx, val, features)
        data.append((x, int(np.argmax(W[x.indices].T @ x.values))))
    return data


@pytest.mark.parametrize("variant,range_part,shards", [("pa", True, 3), ("pa_i", False, 2), ("pa_ii", True, 1)])
def test_binary_kernel_equals_serial_host_rule(variant, range_part, shards):
    from fps_b200.models.pa.device import DevicePA
    data = make_binary(n=40, features=3000, nnz=50)
    algo = {"pa": PassiveAggressiveBinaryAlgorithm.build_pa(),
            "pa_i": PassiveAggressiveBinaryAlgorithm.build_pa_i(0.05),
            "pa_ii": PassiveAggressiveBinaryAlgorithm.build_pa_ii(0.5)}[variant]
    C = {"pa": 1.0, "pa_i": 0.05, "pa_ii": 0.5}[variant]
    pa = DevicePA(3000, algorithm="binary", variant=variant, aggressiveness=C,
                  range_partitioning=range_part, n_shards=shards)
    w = {}
    for x, y in data:                         # one example per launch == the serial schedule
        want_pred = algo.predict(x, w)
        got = pa.step_vectors([x], [y]).item()
        assert (got > 0) == want_pred or abs(sum(v * w.get(i, 0.0) for i, v in x.active_iterator())) < 1e-4
        for i, d in algo.delta(x, w, y):
            w[i] = w.get(i, 0.0) + d
    dumped = dict(r.value for r in pa.dump())
    touched = {i for x, _ in data for i in x.indices.tolist()}
    assert set(dumped) == touched                # untouched features are not dumped
    for i in touched:
        assert dumped[i] == pytest.approx(w.get(i, 0.0), rel=2e-3, abs=2e-5)
    s = pa.stats()
    assert s["records"] == 40 and s["pulls"] == 40 * 50 and s["bad_ids"] == 0


def test_binary_minibatch_acceptance_and_predict_stream():
    """Reference acceptance (accuracy >= 80 % on the first 20 training vectors) with
    asynchronous mini-batches, then unlabelled records through the same entry point."""
    from fps_b200.models.pa.device import transform_binary_device
    data = make_binary()
    stream = [Left(d) for d in data] * 3 + [Right((100 + k, x)) for k, (x, _y) in enumerate(data[:20])]
    out, pa = transform_binary_device(stream, "pa", feature_count=5000, range_partitioning=True,
                                      batch_size=8, pull_limit=10000, n_shards=3)
    model = last_write_wins(rights(out))
    assert binary_accuracy(model, data[:20], PassiveAggressiveBinaryAlgorithm.build_pa()) >= 80.0
    preds = lefts(out)
    assert len(preds) == 20
    agree = sum(1 for (vec, p), (_x, y) in zip(preds, data[:20]) if p == y)
    assert agree >= 16
    s = pa.stats()
    assert 0 < s["max_in_flight"] <= 10000


@pytest.mark.parametrize("algorithm", ["ova", "pb", "ml"])
def test_multiclass_kernel_equals_serial_host_rule(algorithm):
    from fps_b200.models.pa.device import DevicePA
    L, F = 7, 300
    data = make_multi(classes=L, features=F)
    algo = {"ova": PassiveAggressiveOneVersusAll.build_pa_i(L, 0.3),
            "pb": PassiveAggressiveCostBased.build_pb(zero_one_cost, L),
            "ml": PassiveAggressiveCostBased.build_ml(zero_one_cost, L)}[algorithm]
    pa = DevicePA(F, label_count=L, algorithm=algorithm, variant="pa_i", aggressiveness=0.3,
                  range_partitioning=False, n_shards=2)
    W = {}
    for x, y in data:
        pa.step_vectors([x], [y])
        for i, d in algo.delta(x, {k: W[k] for k in x.indices.tolist() if k in W}, y):
            W[i] = W.get(i, np.zeros(L)) + d
    dumped = dict(r.value for r in pa.dump())
    for i, row in W.items():
        np.testing.assert_allclose(dumped[i], row, rtol=5e-3, atol=5e-5)
    # prediction path uses the same decision vector
    x0 = data[0][0]
    pred = pa.step_vectors([x0], None).item()
    assert pred == algo.predict(x0, {k: W[k] for k in x0.indices.tolist() if k in W})


def test_cost_matrix_and_wide_rows():
    """100 classes (BASELINE config 4 shape) with an explicit cost matrix; compares one update."""
    from fps_b200.models.pa.device import DevicePA
    L, F = 100, 2000
    rnd = np.random.default_rng(1)
    cost = rnd.uniform(0.5, 2.0, size=(L, L))
    np.fill_diagonal(cost, 0.0)
    algo = PassiveAggressiveCostBased.build_ml(lambda y, q: float(cost[y, q]), L)
    pa = DevicePA(F, label_count=L, algorithm="ml", cost=cost, n_shards=4)
    model0 = [(i, rnd.normal(size=L)) for i in range(0, F, 3)]
    pa.load(model0)
    W = {i: v.astype(np.float32).astype(np.float64) for i, v in model0}
    x = SparseVector(rnd.choice(F, 40, replace=False), rnd.normal(size=40).astype(np.float32), F)
    pa.step_vectors([x], [5])
    for i, d in algo.delta(x, {k: W[k] for k in x.indices.tolist() if k in W}, 5):
        W[i] = W.get(i, np.zeros(L)) + d
    dumped = dict(r.value for r in pa.dump())
    for i in x.indices.tolist():
        np.testing.assert_allclose(dumped[i], W.get(i, np.zeros(L)), rtol=2e-3, atol=2e-4)


@pytest.mark.parametrize("deferred", [True, False])
@pytest.mark.parametrize("scalar", [True, False])
def test_replicated_table_push_paths_on_one_gpu(deferred, scalar):
    """The replicated-table push paths, exercised on ONE GPU by pointing the view at two buffers that
    play two replicas: deferred (local replica + delta array, then fps_table_flush_delta) and
    immediate without multicast (one red.add per replica).  The multimem.red path needs two GPUs and
    is covered by scripts/multi_gpu_check.py."""
    import ctypes as C
    from fps_b200.ops import _native as N
    from fps_b200.ops.table import ShardedTable
    dim = 1 if scalar else 8
    t = ShardedTable(1000, dim, n_shards=1, init_range=(0.0, 0.0), scalar=scalar)
    other = torch.zeros_like(t.rows[0])
    delta = torch.zeros_like(t.rows[0])
    v = t._view
    v.n_replicas, v.replica_rank = 2, 0
    v.replica_rows[0], v.replica_rows[1] = v.rows[0], other.data_ptr()
    if deferred:
        v.delta_rows = delta.data_ptr()
    ids = torch.tensor([3, 3, 3, 10, 999, 10])
    vals = torch.arange(1, 7, dtype=torch.float32)[:, None].repeat(1, dim)
    t.scatter(ids, vals, "add")
    torch.cuda.synchronize()
    want = torch.zeros(1000, dim)
    want.index_add_(0, ids, vals)
    assert torch.equal(t.rows[0][:, :dim].cpu(), want)
    if deferred:
        assert torch.equal(delta[:, :dim].cpu(), want) and float(other.abs().sum()) == 0.0
        N.check(N.lib().fps_table_flush_delta(C.byref(v), N.current_stream()), "flush")
        torch.cuda.synchronize()
        assert float(delta.abs().sum()) == 0.0
    assert torch.equal(other[:, :dim].cpu(), want)
    t.load(torch.tensor([5]), torch.full((1, dim), 2.5))          # a model load reaches every replica
    torch.cuda.synchronize()
    assert float(other[5, 0]) == 2.5 and float(t.rows[0][5, 0]) == 2.5
