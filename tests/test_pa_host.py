"""Host passive-aggressive: restates ``T/passive/aggressive/PassiveAggressiveParameterServerTest.scala``
(accuracy >= 80 % on the first 20 training vectors, range partitioning, W=PS=3) with a smaller
nnz so it runs in seconds, ``T/passive/aggressive/entities/SparseVectorTest.scala``, and covers
what upstream never tested: PA-I / PA-II, one-vs-all, cost-based, the offline jobs, evaluators."""
import math
import random

import numpy as np
import pytest

from fps_b200 import Left, Right, last_write_wins, lefts, rights
from fps_b200.models.pa import (LegacySparseVector, PassiveAggressiveBinaryAlgorithm,
                                PassiveAggressiveCostBased, PassiveAggressiveFilter,
                                PassiveAggressiveFilterII, PassiveAggressiveOneVersusAll, SparseVector,
                                pa_binary_classification_offline, pa_multi_classification_offline,
                                transform_binary, transform_multiclass,
                                transform_multiclass_with_long_id, zero_one_cost)
from fps_b200.utils.evaluation import (NDCGSink, NoSuccessExceptionReceived, SuccessException,
                                       binary_accuracy, execute_with_success_check,
                                       multiclass_accuracy, ndcg_of_rank)


def make_binary(n=80, features=5000, nnz=120, seed=50):
    rnd = np.random.default_rng(seed)
    w_true = rnd.normal(size=features)
    data = []
    for _ in range(n):
        idx = rnd.choice(features, nnz, replace=False)
        val = rnd.normal(size=nnz)
        x = SparseVector(idx, val, features)
        data.append((x, bool(x.dot(w_true) > 0)))
    return data


def make_multi(n=150, features=400, nnz=30, classes=5, seed=3):
    rnd = np.random.default_rng(seed)
    W = rnd.normal(size=(features, classes))
    data = []
    for _ in range(n):
        idx = rnd.choice(features, nnz, replace=False)
        val = rnd.normal(size=nnz)
        x = SparseVector(idx, val, features)
        data.append((x, int(np.argmax(W[x.indices].T @ x.values))))
    return data


def test_sparse_vector_builders_equal():
    a = LegacySparseVector({1: 1.0, 5: 2.0})
    b = LegacySparseVector.build_from_keys_values([1, 5], [1.0, 2.0])
    c = LegacySparseVector.build_from_pairs([(5, 2.0), (1, 1.0)])
    assert a == b == c and hash(a) == hash(c)
    s = SparseVector([5, 1], [2.0, 1.0], 10)
    assert s.indices.tolist() == [1, 5] and s.dot(np.arange(10.0)) == 11.0 and s.norm_sqr() == 5.0
    assert s == SparseVector.from_pairs([(1, 1.0), (5, 2.0)], 10)


def test_binary_tau_variants_hand_computed():
    x = SparseVector([0, 2], [1.0, 2.0], 3)          # ||x||^2 = 5
    model = {0: 0.0, 2: 0.0}                          # margin 0 -> loss 1
    assert dict(PassiveAggressiveBinaryAlgorithm.build_pa().delta(x, model, True)) == pytest.approx({0: 0.2, 2: 0.4})
    assert dict(PassiveAggressiveBinaryAlgorithm.build_pa_i(0.1).delta(x, model, False)) == pytest.approx({0: -0.1, 2: -0.2})
    d2 = dict(PassiveAggressiveBinaryAlgorithm.build_pa_ii(0.5).delta(x, model, True))   # 1/(5+1)
    assert d2 == pytest.approx({0: 1 / 6, 2: 2 / 6})
    # no loss -> no update
    assert PassiveAggressiveBinaryAlgorithm.build_pa().delta(x, {0: 5.0, 2: 5.0}, True) == []
    # legacy filter with real division in PAF-II (reference: integer division -> 0)
    lv = LegacySparseVector({0: 1.0, 2: 2.0})
    assert PassiveAggressiveFilter().delta(lv, {}, 1) == pytest.approx({0: 0.2, 2: 0.4})
    assert PassiveAggressiveFilterII(1.0).delta(lv, {}, 1) == pytest.approx({0: 1 / 5.5, 2: 2 / 5.5})


def test_ova_and_cost_based_hand_computed():
    x = SparseVector([1], [2.0], 4)                  # ||x||^2 = 4
    ova = PassiveAggressiveOneVersusAll.build_pa(3)
    (i, row), = ova.delta(x, {1: np.zeros(3)}, 2)
    assert i == 1 and np.allclose(row, 2.0 * np.array([-0.25, -0.25, 0.25]))
    pb = PassiveAggressiveCostBased.build_pb(zero_one_cost, 3)
    model = {1: np.array([1.0, 0.0, 0.0])}           # d = [2,0,0]; argmax 0 != label 2
    (i, row), = pb.delta(x, model, 2)
    t = (2.0 - 0.0 + 1.0) / (2 * 4.0)
    assert np.allclose(row, [-t * 2.0, 0.0, t * 2.0])
    # second call gives the same delta (no state leaks between calls)
    (_, row2), = pb.delta(x, model, 2)
    assert np.allclose(row, row2)
    ml = PassiveAggressiveCostBased.build_ml(zero_one_cost, 3)
    assert ml.delta(x, {1: np.array([0.0, 0.0, 5.0])}, 2) == []   # already separated by > sqrt(cost)


@pytest.mark.parametrize("range_partitioning", [True, False])
def test_binary_pa_job_reference_acceptance(range_partitioning):
    data = make_binary()
    stream = [Left((x, y)) for x, y in data] * 3
    out = transform_binary()(stream, 3, 3, PassiveAggressiveBinaryAlgorithm.build_pa(), 10000, 5000,
                             range_partitioning, 100)
    model = last_write_wins(rights(out))
    acc = binary_accuracy(model, data[:20], PassiveAggressiveBinaryAlgorithm.build_pa())
    assert acc >= 80.0
    assert all(isinstance(v, float) for v in model.values())


def test_binary_pa_prediction_records_and_warm_start():
    data = make_binary(n=60)
    out = transform_binary()([Left(d) for d in data] * 3, 2, 2,
                             PassiveAggressiveBinaryAlgorithm.build_pa_i(1.0), 500, 5000, True, 50)
    model = last_write_wins(rights(out))
    # resume: load the dumped model, only predict
    queries = [Right((1000 + k, x)) for k, (x, _y) in enumerate(data[:10])]
    out2 = transform_binary(model=list(model.items()))(
        queries, 2, 2, PassiveAggressiveBinaryAlgorithm.build_pa_i(1.0), 500, 5000, True, 50)
    preds = lefts(out2)
    assert len(preds) == 10
    want = {x: PassiveAggressiveBinaryAlgorithm.build_pa().predict(x, model) for x, _ in data[:10]}
    assert all(want[vec] == p for vec, p in preds)
    # the reloaded model comes back unchanged in the dump
    model2 = last_write_wins(rights(out2))
    assert model2.keys() == model.keys() and all(abs(model2[k] - model[k]) < 1e-12 for k in model)


@pytest.mark.parametrize("algo", ["ova", "ova_ii", "pb", "ml"])
def test_multiclass_pa_job_learns(algo):
    data = make_multi()
    L = 5
    method = {"ova": PassiveAggressiveOneVersusAll.build_pa(L),
              "ova_ii": PassiveAggressiveOneVersusAll.build_pa_ii(L, 1.0),
              "pb": PassiveAggressiveCostBased.build_pb(zero_one_cost, L),
              "ml": PassiveAggressiveCostBased.build_ml(zero_one_cost, L)}[algo]
    stream = [Left(d) for d in data] * 4 + [Right((7, data[0][0]))]
    out = transform_multiclass_with_long_id()(stream, 3, 2, method, 2000, L, 400, False, 50)
    model = last_write_wins(rights(out))
    assert multiclass_accuracy(model, data, method) >= 70.0
    assert [vid for vid, _ in lefts(out)] == [7]
    out_v = transform_multiclass()([Right((7, data[0][0]))], 1, 1, method, 10, L, 400, True, 50)
    assert lefts(out_v)[0][0] == data[0][0]          # VecId is the vector itself


def test_offline_binary_job_trains_then_predicts():
    data = make_binary(n=100, features=2000, nnz=60, seed=8)
    train = [(x, 1 if y else -1) for x, y in data[:80]]
    test = [(x, 1 if y else -1) for x, y in data[80:]]
    out = pa_binary_classification_offline(train, test, 3, 2, iterations=5, paf_type=1, paf_const=1.0,
                                           pull_limit=300, iteration_wait_time=50, feature_count=2000)
    preds = lefts(out)
    assert len(preds) == 20
    acc = sum(1 for label, pred, _v in preds if label == pred) / 20
    assert acc >= 0.6
    assert len(rights(out)) > 0                       # SimplePSLogic: one record per push


def test_offline_multiclass_job_is_really_multiclass():
    data = make_multi(n=420, features=100, nnz=20)
    out = pa_multi_classification_offline(data[:400], data[400:], 2, 2, iterations=3,
                                          method=PassiveAggressiveOneVersusAll.build_pa_i(5, 0.5),
                                          pull_limit=500, iteration_wait_time=50, feature_count=100)
    preds = lefts(out)
    assert len(preds) == 20 and {p for _l, p, _v in preds} <= set(range(5))
    assert sum(1 for l, p, _v in preds if l == p) / 20 >= 0.5


def test_ndcg_sink_and_success_check(tmp_path):
    assert ndcg_of_rank(1) == pytest.approx(1.0) and ndcg_of_rank(0) == 0.0
    assert ndcg_of_rank(3) == pytest.approx(math.log(2) / math.log(4))
    sink = NDCGSink(period_length=10)
    sink.consume([(1, 5, 3, [(0.9, 5), (0.1, 6)]), (1, 7, 12, [(0.9, 5), (0.1, 7)]), (2, 9, 15, [])])
    assert sink.hits == 2 and sink.count == 3
    assert sink.ndcg == pytest.approx((1.0 + math.log(2) / math.log(3) + 0.0) / 3)
    assert [p[0] for p in sink.per_period()] == [0, 1]
    sink.write_csv(str(tmp_path / "n.csv"))
    sink.write_text(str(tmp_path / "n.txt"))
    assert "all," in (tmp_path / "n.csv").read_text()

    def job():
        try:
            raise SuccessException(42)
        except SuccessException as e:
            raise RuntimeError("job failed") from e

    got = []
    execute_with_success_check(job, got.append)
    assert got == [42]
    with pytest.raises(NoSuccessExceptionReceived):
        execute_with_success_check(lambda: None, got.append)
