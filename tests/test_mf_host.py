"""Host MF: restates ``T/matrix/factorization/PSOfflineMatrixFactorizationTest.scala`` (RMSE < 0.5)
and adds tests for the pieces that had none upstream (online MF, LEMP, top-K, learner+generator)."""
import math
import random

import numpy as np
import pytest

from fps_b200 import Left, Right, last_write_wins, lefts, rights
from fps_b200.models.mf import (COORD, INCR, LC, LENGTH, LI, FactorIsNotANumberException, Rating,
                                SGDUpdater, attach_length, from_string, lemp_top_k, ps_offline_mf,
                                ps_online_learner_and_generator, ps_online_mf, ps_top_k_generator,
                                vector_sum)
from fps_b200.ops import philox


def small_ratings(seed=47, n=100, users=20, items=15):
    rnd = random.Random(seed)
    seen, out = set(), []
    for _ in range(n):
        u, i = rnd.randrange(users), rnd.randrange(items)
        if (u, i) not in seen:
            seen.add((u, i))
            out.append(Rating(u, i, rnd.random()))
    return out


def rmse(ratings, users, items):
    se = 0.0
    for r in ratings:
        se += (r.rating - float(np.dot(users[r.user], items[r.item]))) ** 2
    return math.sqrt(se / len(ratings))


@pytest.mark.parametrize("schedule", ["round_robin", "random"])
def test_offline_mf_reference_acceptance(schedule):
    """Reference config: <=100 ratings, 20x15, rank 15, lr 0.01, 10 epochs, init U[0,1), W=PS=4,
    pullLimit 10 -> training RMSE < 0.5."""
    ratings = small_ratings()
    out = ps_offline_mf(ratings, num_factors=15, range_min=0.0, range_max=1.0, learning_rate=0.01,
                        iterations=10, pull_limit=10, worker_parallelism=4, ps_parallelism=4,
                        iteration_wait_time=50, schedule=schedule, seed=1)
    users = last_write_wins(lefts(out))
    items = last_write_wins(rights(out))
    assert rmse(ratings, users, items) < 0.5
    assert len(lefts(out)) == len(ratings) * 10


def test_offline_mf_negative_sampling_blocks():
    ratings = small_ratings(n=60)
    out = ps_offline_mf(ratings, num_factors=4, learning_rate=0.01, iterations=1,
                        negative_sample_rate=3, worker_parallelism=2, ps_parallelism=2,
                        iteration_wait_time=10)
    assert len(lefts(out)) > len(ratings)      # negatives were trained too


def test_online_mf_runs_and_learns():
    ratings = small_ratings(n=90) * 20
    out = ps_online_mf(ratings, num_factors=8, range_min=0.0, range_max=0.3, learning_rate=0.05,
                       negative_sample_rate=0, worker_parallelism=3, ps_parallelism=2,
                       iteration_wait_time=10)
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    assert rmse(ratings, users, items) < 0.5


def test_online_mf_negative_samples_respect_memory():
    ratings = [Rating(0, i, 1.0) for i in range(6)]
    out = ps_online_mf(ratings, num_factors=2, negative_sample_rate=2, user_memory=3,
                       worker_parallelism=1, ps_parallelism=1, iteration_wait_time=10)
    assert len(lefts(out)) > 6


def test_sgd_updater_and_nan_guard():
    du, dv = SGDUpdater(0.1).delta(1.0, np.array([1.0, 0.0]), np.array([0.5, 2.0]))
    assert np.allclose(du, 0.1 * 0.5 * np.array([0.5, 2.0]))
    assert np.allclose(dv, 0.1 * 0.5 * np.array([1.0, 0.0]))
    du2, _ = SGDUpdater(0.1, lambda_=0.5).delta(1.0, np.array([1.0, 0.0]), np.array([0.5, 2.0]))
    assert np.allclose(du2, du - 0.1 * 0.5 * np.array([1.0, 0.0]))
    with pytest.raises(FactorIsNotANumberException):
        vector_sum([float("nan")], [1.0])


def test_philox_known_answer_and_range():
    w = philox.philox4x32(0, 0, 0, 0, 0, 0)
    assert [int(x) for x in w] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    v = philox.ranged_uniform(np.arange(1000), 10, -0.01, 0.01, seed=7)
    assert v.shape == (1000, 10) and v.min() >= -0.01 and v.max() < 0.01
    assert np.array_equal(v[5], philox.ranged_uniform([5], 10, -0.01, 0.01, seed=7)[0])
    assert not np.array_equal(v, philox.ranged_uniform(np.arange(1000), 10, -0.01, 0.01, seed=8))


def test_pruning_from_string():
    assert from_string("length") == LENGTH()
    assert from_string("coord") == COORD()
    assert from_string("incr:3") == INCR(3)
    assert from_string("lc:2.5") == LC(2.5)
    assert from_string("li:5:2.5") == LI(5, 2.5)
    with pytest.raises(ValueError):
        from_string("nope")


@pytest.mark.parametrize("strategy", [LENGTH(), COORD(), INCR(3), LC(1.2), LI(3, 1.2)])
def test_lemp_topk_is_exact(strategy):
    rng = np.random.default_rng(3)
    k, n = 8, 400
    vecs = rng.normal(size=(n, k)) * rng.uniform(0.1, 3.0, size=(n, 1))
    model = {i: attach_length(vecs[i]) for i in range(n)}
    desc = sorted(((model[i][0], i) for i in range(n)), key=lambda t: -t[0])
    for _ in range(5):
        q = rng.normal(size=k)
        stats = {}
        got = lemp_top_k(attach_length(q), desc, model, 10, 25, strategy, stats)
        want = np.sort(vecs @ q)[::-1][:10]
        assert np.allclose(sorted((s for s, _ in got), reverse=True), want)
        assert stats["scored"] < n          # pruning or early exit did something


def test_topk_generator_job():
    rng = np.random.default_rng(5)
    k, n_items, n_users = 6, 120, 10
    items = rng.normal(size=(n_items, k))
    users = rng.normal(size=(n_users, k))
    model = [Left((u, attach_length(users[u]))) for u in range(n_users)] + \
            [Right((i, attach_length(items[i]))) for i in range(n_items)]
    queries = [Rating(u, 0, 1.0, 100 + u) for u in range(n_users)] + [Rating(999, 0, 1.0, 7)]
    res = ps_top_k_generator(queries, model, num_factors=k, K=5, worker_k=5, bucket_size=10,
                             pruning_algorithm=LI(2, 1.5), worker_parallelism=3, ps_parallelism=2,
                             iteration_wait_time=10, user_memory=-1)
    assert len(res) == len(queries)
    by_ts = {ts: topk for (_item, ts, topk) in res}
    for u in range(n_users):
        want = np.argsort(-(items @ users[u]))[:5]
        assert [i for _s, i in by_ts[100 + u]] == list(want)
    assert by_ts[7] == []         # unknown user -> invalid param -> empty top-K


def test_learner_and_generator_prequential():
    rnd = random.Random(2)
    stream = [Rating(rnd.randrange(12), rnd.randrange(30), 1.0, t) for t in range(300)]
    res = ps_online_learner_and_generator(
        stream, num_factors=5, range_min=-0.1, range_max=0.1, learning_rate=0.1,
        negative_sample_rate=2, K=10, worker_k=10, bucket_size=5, worker_parallelism=3,
        ps_parallelism=2, iteration_wait_time=10, pull_limit=50)
    assert len(res) == len(stream)
    # recommendations never contain items the user has already consumed (seen filter)
    seen = {}
    by_ts = {r[2]: r for r in res}
    for t in range(300):
        user, item, _ts, topk = by_ts[t]
        assert all(i not in seen.get(user, set()) for _s, i in topk)
        seen.setdefault(user, set()).add(item)
    assert any(len(r[3]) > 0 for r in res[50:])
