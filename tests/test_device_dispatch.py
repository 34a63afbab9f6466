"""The ``engine=`` switch of the public entry points (parallel/device_engine.py).

CPU part: registry, limiter unwrapping, and the loud fallback when no GPU engine is available.
GPU part (marked): the same test body through ``engine="host"`` and ``engine="device"`` of the public
entry points only - ``transform``, ``ps_offline_mf``, ``ps_online_mf``, ``transform_binary``."""
import math
import random
import warnings

import numpy as np
import pytest

from fps_b200 import Left, Right, last_write_wins, lefts, rights
from fps_b200.models.mf import Rating, ps_offline_mf, ps_online_mf
from fps_b200.models.mf.vector import vector_sum
from fps_b200.parallel import device_engine as DE
from fps_b200.parameter_server import transform
from fps_b200.worker_logic import add_pull_limiter


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def _ratings(seed=47, n=100, users=20, items=15):
    rnd = random.Random(seed)
    seen, out = set(), []
    for _ in range(n):
        a, b = rnd.randrange(users), rnd.randrange(items)
        if (a, b) not in seen:
            seen.add((a, b)); out.append(Rating(a, b, rnd.random()))
    return out


def _rmse(out, ratings):
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    se = sum((r.rating - float(np.dot(users[r.user], items[r.item]))) ** 2 for r in ratings)
    return math.sqrt(se / len(ratings))


def test_registry_knows_the_builtin_logics_and_update_rules():
    from fps_b200.models.mf.workers import PSOfflineMatrixFactorizationWorker
    w = PSOfflineMatrixFactorizationWorker(4, 0.0, 1.0, 0.01, 0, 128, 1)
    inner, limit = DE.unwrap_limiter(add_pull_limiter(w, 77))
    assert inner is w and limit == 77
    assert DE.unwrap_limiter(w) == (w, 0)
    assert DE.runner_for(w) is not None
    assert DE.param_update_mode(vector_sum) == "add"
    assert DE.param_update_mode(lambda a, b: b) is None

    def overwrite(_old, new):
        return new
    DE.register_param_update(overwrite, "overwrite")
    assert DE.param_update_mode(overwrite) == "overwrite"
    with pytest.raises(ValueError):
        DE.register_param_update(overwrite, "multiply")


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU fallback")
def test_engine_device_without_a_gpu_falls_back_loudly():
    ratings = _ratings()
    with pytest.warns(DE.DeviceFallbackWarning):
        out = ps_offline_mf(ratings, num_factors=15, range_min=0.0, range_max=1.0, learning_rate=0.01,
                            iterations=10, pull_limit=10, worker_parallelism=2, ps_parallelism=2,
                            iteration_wait_time=10, engine="device", seed=1)
    assert _rmse(out, ratings) < 0.5
    with pytest.raises(ValueError):
        ps_offline_mf(ratings, engine="tpu")


def test_unknown_logic_falls_back_with_a_warning_even_on_a_gpu_box():
    from fps_b200.api import WorkerLogic

    class Echo(WorkerLogic):
        def on_recv(self, data, ps):
            ps.pull(data)

        def on_pull_recv(self, param_id, value, ps):
            ps.push(param_id, 1)
            ps.output((param_id, value))

    with pytest.warns(DE.DeviceFallbackWarning):
        out = transform([1, 2, 3, 1], Echo(), lambda i: 10 * i, lambda a, b: a + b, 2, 2, 10, engine="device")
    assert sorted(e.value for e in out if e.is_left)[0][0] == 1
    assert last_write_wins(rights(out)) == {1: 12, 2: 21, 3: 31}


# ---------------------------------------------------------------------------------------------
# GPU: same body, both engines, public entry points only
# ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["host", "device"])
def test_ps_offline_mf_reference_acceptance_on_both_engines(engine):
    ratings = _ratings()
    out = ps_offline_mf(ratings, num_factors=15, range_min=0.0, range_max=1.0, learning_rate=0.01,
                        iterations=10, pull_limit=10, worker_parallelism=4, ps_parallelism=4,
                        iteration_wait_time=10, engine=engine, seed=1)
    assert _rmse(out, ratings) < 0.5
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    assert set(users) == {r.user for r in ratings} and set(items) == {r.item for r in ratings}


@pytest.mark.gpu
def test_transform_with_a_builtin_logic_dispatches_to_the_device_functor():
    """README-signature ``transform`` with the offline MF worker under a pull limiter: the device run
    (conflict-free waves) ends in the same factors as the host engine's serial schedule."""
    from fps_b200.models.mf.factors import RangedRandomFactorInitializerDescriptor
    from fps_b200.models.mf.workers import PSOfflineMatrixFactorizationWorker
    from fps_b200.ops import philox
    from fps_b200.utils.eof import WrapWithEOF, flat_map_with_eof
    rnd = random.Random(9)
    ratings = [Rating(rnd.randrange(30), rnd.randrange(25), rnd.random()) for _ in range(400)]
    res = {}
    for engine in ("host", "device"):
        src = flat_map_with_eof(ratings, WrapWithEOF, 1, lambda key, n: key % n, lambda r: r.user)
        worker = PSOfflineMatrixFactorizationWorker(12, 0.0, 0.5, 0.02, 0, 128, 2, seed=4, shuffle=False)
        init = RangedRandomFactorInitializerDescriptor(12, 0.0, 0.5, 4, philox.STREAM_ITEM).open()
        kw = dict(schedule="feedback_first") if engine == "host" else dict(device_kw=dict(conflict_free=True))
        with warnings.catch_warnings():
            warnings.simplefilter("error", DE.DeviceFallbackWarning)      # a fallback would be a failure here
            out = transform(src, add_pull_limiter(worker, 1), lambda i: init.next_factor(i), vector_sum, 1, 1, 10,
                            engine=engine, **kw)
        res[engine] = (last_write_wins(lefts(out)), last_write_wins(rights(out)))
    (hu, hi), (du, di) = res["host"], res["device"]
    assert set(hu) == set(du) and set(hi) == set(di)
    for k in hu:
        np.testing.assert_allclose(du[k], hu[k], rtol=5e-4, atol=5e-5)
    for k in hi:
        np.testing.assert_allclose(di[k], hi[k], rtol=5e-4, atol=5e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["host", "device"])
def test_ps_online_mf_honours_user_memory_and_learns_on_both_engines(engine):
    rnd = random.Random(3)
    stream = []
    for t in range(6000 if engine == "host" else 30000):
        u = rnd.randrange(200)
        stream.append(Rating(u, (u % 8) * 25 + rnd.randrange(25), 1.0, t))
    out = ps_online_mf(stream, num_factors=8, range_min=-0.1, range_max=0.1, learning_rate=0.2,
                       negative_sample_rate=3, user_memory=16, pull_limit=1600, iteration_wait_time=10,
                       engine=engine, seed=2, **({"batch_size": 500} if engine == "device" else {}))
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    pos = np.mean([users[r.user] @ items[r.item] for r in stream[-500:]])
    neg = np.mean([users[r.user] @ items[(r.item + 100) % 200] for r in stream[-500:] if (r.item + 100) % 200 in items])
    assert pos > neg + 0.1, (engine, pos, neg)


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["host", "device"])
def test_transform_binary_on_both_engines(engine):
    from fps_b200.models.pa import PassiveAggressiveBinaryAlgorithm, SparseVector, transform_binary
    rng = np.random.default_rng(50)
    n_feat = 2000
    w_true = rng.normal(size=n_feat)
    recs, vecs, labels = [], [], []
    for _ in range(300):
        ix = np.sort(rng.choice(n_feat, 40, replace=False))
        vv = rng.normal(size=40)
        v = SparseVector(ix, vv, n_feat)
        vecs.append(v); labels.append(bool(vv @ w_true[ix] > 0))
        recs.append(Left((v, labels[-1])))
    recs = recs * 3 + [Right((k, v)) for k, v in enumerate(vecs[:100])]
    out = transform_binary()(recs, 3, 3, PassiveAggressiveBinaryAlgorithm.build_pa_i(1.0), 10_000, n_feat, True,
                             10, engine=engine)
    by_vec = {e.value[0]: e.value[1] for e in out if e.is_left}      # an unlabelled example is identified by its vector
    preds = {k: by_vec[v] for k, v in enumerate(vecs[:100]) if v in by_vec}
    acc = np.mean([preds[k] == labels[k] for k in range(100)])
    assert acc >= 0.8, (engine, acc)
    assert len(last_write_wins(rights(out))) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["host", "device"])
def test_offline_pa_jobs_on_both_engines(engine):
    """PABinaryClassificationOffline / the genuinely multiclass offline job: epochs over the training set, then
    predictions for the test set, through the same public entry points on both engines."""
    from fps_b200.models.pa import (PassiveAggressiveOneVersusAll, SparseVector, pa_binary_classification_offline,
                                    pa_multi_classification_offline)
    rnd = np.random.default_rng(8)
    F = 500
    w_true = rnd.normal(size=F)
    data = []
    for _ in range(360):
        ix = rnd.choice(F, 40, replace=False)
        vv = rnd.normal(size=40)
        x = SparseVector(ix, vv, F)
        data.append((x, 1 if x.dot(w_true) > 0 else -1))
    out = pa_binary_classification_offline(data[:300], data[300:], 3, 2, iterations=4, paf_type=1, paf_const=1.0,
                                           pull_limit=300, iteration_wait_time=50, feature_count=F, engine=engine)
    preds = lefts(out)
    assert len(preds) == 60 and {p for _l, p, _v in preds} <= {1, -1}
    assert sum(1 for label, pred, _v in preds if label == pred) / 60 >= 0.58      # both engines land near 0.64 on this little set
    assert len(rights(out)) > 0
    Wm = rnd.normal(size=(100, 5))
    multi = []
    for _ in range(420):
        ix = rnd.choice(100, 20, replace=False)
        vv = rnd.normal(size=20)
        multi.append((SparseVector(ix, vv, 100), int(np.argmax(vv @ Wm[ix]))))
    out = pa_multi_classification_offline(multi[:400], multi[400:], 2, 2, iterations=3,
                                          method=PassiveAggressiveOneVersusAll.build_pa_i(5, 0.5),
                                          pull_limit=500, iteration_wait_time=50, feature_count=100, engine=engine)
    preds = lefts(out)
    assert len(preds) == 20 and {p for _l, p, _v in preds} <= set(range(5))
    assert sum(1 for l, p, _v in preds if l == p) / 20 >= 0.35          # chance is 0.2; 20 test records are noisy


@pytest.mark.gpu
def test_user_tensor_logic_runs_on_the_device_engine_through_transform():
    """A user-defined logic written as a tensor program (``TensorWorkerLogic``): pull the ids of the batch,
    push +1 to each and +10 to id 0, output the sum of the pulled rows.  Through ``transform(engine="device")``
    with a pull limit smaller than the batch (the batch is pulled and pushed in slices)."""
    import torch
    from fps_b200.parallel.device_engine import TensorWorkerLogic

    class Count(TensorWorkerLogic):
        def pull_ids(self, batch):
            return batch

        def on_pull_batch(self, batch, ids, rows):
            ones = torch.ones((ids.numel(), 3), device="cuda")
            push_ids = torch.cat([ids, torch.zeros(1, dtype=torch.int64, device="cuda")])
            deltas = torch.cat([ones, torch.full((1, 3), 10.0, device="cuda")])
            return push_ids, deltas, rows[:, :3].sum(dim=0)

    g = torch.Generator().manual_seed(0)
    batches = [torch.randint(1, 50, (200,), generator=g) for _ in range(5)]
    with warnings.catch_warnings():
        warnings.simplefilter("error", DE.DeviceFallbackWarning)
        out = transform(batches, add_pull_limiter(Count(), 64), None, "add", 1, 3, 10, engine="device",
                        device_kw=dict(n_ids=50, dim=3, init_range=(0.0, 0.0), n_shards=3))
    counts = torch.bincount(torch.cat(batches), minlength=50).numpy()
    model = last_write_wins(rights(out))
    slices = sum((len(b) + 63) // 64 for b in batches)
    assert len(lefts(out)) == slices
    for k in range(1, 50):
        if counts[k]:
            np.testing.assert_allclose(model[k], np.full(3, counts[k]), rtol=0, atol=1e-5)
        else:
            assert k not in model                    # lazy table: never touched, never dumped
    np.testing.assert_allclose(model[0], np.full(3, 10.0 * slices), rtol=0, atol=1e-4)
    # the worker outputs are the sums of the rows as pulled: monotone over the job
    sums = [float(v[0]) for v in lefts(out)]
    assert sums[0] >= 0 and sums[-1] > sums[0]


@pytest.mark.gpu
@pytest.mark.parametrize("engine", ["host", "device"])
def test_ps_top_k_generator_on_both_engines(engine):
    """psTopKGenerator through the public entry point: the same model stream and queries, brute-force truth."""
    from fps_b200.models.mf import attach_length, ps_top_k_generator
    from fps_b200.models.mf.pruning import LENGTH
    rng = np.random.default_rng(5)
    k, n_items, n_users = 8, 400, 12
    items = rng.normal(size=(n_items, k)).astype(np.float32)
    users = rng.normal(size=(n_users, k)).astype(np.float32)
    model = ([Left((u, attach_length(users[u]))) for u in range(n_users)] +
             [Right((i, attach_length(items[i]))) for i in range(n_items)])
    queries = [Rating(u, 0, 1.0, 100 + u) for u in range(n_users)]
    res = ps_top_k_generator(queries, model, num_factors=k, K=5, worker_k=5, user_memory=-1,
                             pruning_algorithm=LENGTH(), worker_parallelism=2, ps_parallelism=2,
                             iteration_wait_time=10, engine=engine)
    by_ts = {ts: topk for (_item, ts, topk) in res}
    for u in range(n_users):
        want = np.argsort(-(items.astype(np.float64) @ users[u].astype(np.float64)))[:5]
        got = [i for _s, i in by_ts[100 + u]]
        assert got[0] == want[0] and len(set(got) & set(want.tolist())) >= 4, (engine, u, got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("strategy", ["coord", "li:3:1.5"])
def test_ps_top_k_generator_device_with_candidate_filters(strategy):
    """engine="device", candidate_filters=True: the reference's bucket loop with the strategy's own filters on
    the device (no LENGTH-only note), same answers as brute force."""
    import warnings
    from fps_b200.models.mf import attach_length, ps_top_k_generator
    from fps_b200.models.mf.pruning import from_string
    rng = np.random.default_rng(6)
    k, n_items, n_users = 8, 500, 10
    items = (rng.normal(size=(n_items, k)) * rng.uniform(0.1, 3.0, size=(n_items, 1))).astype(np.float32)
    users = rng.normal(size=(n_users, k)).astype(np.float32)
    model = ([Left((u, attach_length(users[u]))) for u in range(n_users)] +
             [Right((i, attach_length(items[i]))) for i in range(n_items)])
    queries = [Rating(u, 0, 1.0, 100 + u) for u in range(n_users)]
    with warnings.catch_warnings():
        warnings.simplefilter("error", DE.DeviceFallbackWarning)
        res = ps_top_k_generator(queries, model, num_factors=k, K=5, worker_k=5, user_memory=-1, bucket_size=40,
                                 pruning_algorithm=from_string(strategy), worker_parallelism=2, ps_parallelism=2,
                                 iteration_wait_time=10, engine="device", candidate_filters=True)
    by_ts = {ts: topk for (_item, ts, topk) in res}
    for u in range(n_users):
        want = np.argsort(-(items.astype(np.float64) @ users[u].astype(np.float64)))[:5]
        got = [i for _s, i in by_ts[100 + u]]
        assert got[0] == want[0] and len(set(got) & set(want.tolist())) >= 4, (strategy, u, got, want)


@pytest.mark.gpu
def test_learner_generator_entry_point_dispatches_and_announces_pruning():
    from fps_b200.models.mf import ps_online_learner_and_generator
    rnd = random.Random(2)
    stream = [Rating(u := rnd.randrange(40), (u % 4) * 10 + rnd.randrange(10), 1.0, t) for t in range(1500)]
    with pytest.warns(DE.DeviceFallbackWarning, match="LENGTH bucket bound"):      # default strategy is LI(5, 2.5)
        out = ps_online_learner_and_generator(stream, num_factors=8, range_min=-0.1, range_max=0.1, learning_rate=0.1,
                                              negative_sample_rate=3, user_memory=4, K=10, worker_k=10, pull_limit=500,
                                              engine="device", n_users=40, n_items=40, batch_size=50)
    assert len(out) == len(stream)
    assert out.hit_rate(1000) > out.hit_rate(0) - 0.05 and out.hit_rate(1000) > 0.3      # 10 of 40 items = 0.25 by chance


@pytest.mark.gpu
def test_learner_generator_with_candidate_filters_runs_the_strategy_on_the_device():
    import warnings
    from fps_b200.models.mf import ps_online_learner_and_generator
    rnd = random.Random(2)
    stream = [Rating(u := rnd.randrange(40), (u % 4) * 10 + rnd.randrange(10), 1.0, t) for t in range(1200)]
    with warnings.catch_warnings():
        warnings.simplefilter("error", DE.DeviceFallbackWarning)            # LI(5, 2.5) is executed, not announced
        out = ps_online_learner_and_generator(stream, num_factors=8, range_min=-0.1, range_max=0.1, learning_rate=0.1,
                                              negative_sample_rate=3, user_memory=4, K=10, worker_k=10, pull_limit=500,
                                              bucket_size=8, engine="device", candidate_filters=True, n_users=40,
                                              n_items=40, batch_size=50)
    assert len(out) == len(stream)
    assert out.hit_rate(800) > 0.3


@pytest.mark.gpu
def test_lock_logic_on_the_device_is_the_sequential_read_modify_write():
    """LockPSLogicA on both engines (LockPSLogicA.scala:13-46): x <- 0.5 * x + v per record with an overwriting
    update.  With the lock, the cycles on one key run one after the other in arrival order (device: conflict-free
    rounds), so both engines give the sequential result; without it the device loses the earlier of two
    updates of a key in one batch."""
    import torch
    from collections import defaultdict, deque
    from fps_b200.api import WorkerLogic
    from fps_b200.parallel.device_engine import TensorWorkerLogic, conflict_free_rounds
    from fps_b200.server.logics import LockPSLogicA

    rng = np.random.default_rng(3)
    n_keys, n = 12, 300
    keys = rng.integers(0, n_keys, size=n)
    vals = rng.normal(size=n).astype(np.float32)
    want = np.zeros(n_keys)
    for k, v in zip(keys, vals):
        want[k] = 0.5 * want[k] + v

    # the rounds themselves: no key twice in a round, arrival order kept per key
    rounds = conflict_free_rounds(torch.from_numpy(keys).cuda())
    seen_upto = defaultdict(lambda: -1)
    for r in rounds:
        ks = keys[r.cpu().numpy()]
        assert len(set(ks.tolist())) == len(ks)
        for pos, k in zip(r.cpu().numpy().tolist(), ks.tolist()):
            assert pos > seen_upto[k]
            seen_upto[k] = pos
    assert sum(int(r.numel()) for r in rounds) == n

    overwrite = lambda old, new: new                                     # noqa: E731
    DE.register_param_update(overwrite, "overwrite")

    class Decay(TensorWorkerLogic):
        def pull_ids(self, batch):
            return batch[0]

        def on_pull_batch(self, batch, ids, rows):
            return ids, 0.5 * rows[:, :1] + batch[1][:, None], None

    batches = [(torch.from_numpy(keys[s:s + 100]).cuda(), torch.from_numpy(vals[s:s + 100]).cuda())
               for s in range(0, n, 100)]
    kw = dict(n_ids=n_keys, dim=1, init_range=(0.0, 0.0), n_shards=2)
    out = transform(batches, Decay(), ps_logic=LockPSLogicA(lambda _k: 0.0, overwrite), worker_parallelism=1,
                    ps_parallelism=2, iteration_wait_time=10, engine="device", device_kw=kw)
    got = last_write_wins(rights(out))
    np.testing.assert_allclose([float(got[k][0]) for k in range(n_keys)], want, rtol=1e-5, atol=1e-6)

    out_free = transform(batches, Decay(), ps_logic=LockPSLogicA(lambda _k: 0.0, overwrite), worker_parallelism=1,
                         ps_parallelism=2, iteration_wait_time=10, engine="device", device_kw=dict(kw, locking=False))
    free = last_write_wins(rights(out_free))
    assert max(abs(float(free[k][0]) - want[k]) for k in range(n_keys)) > 1e-2        # lost updates

    # host engine, the reference's per-record logic under the same lock logic
    class DecayHost(WorkerLogic):
        def __init__(self):
            self.pending = defaultdict(deque)

        def on_recv(self, data, ps):
            self.pending[data[0]].append(data[1])
            ps.pull(data[0])

        def on_pull_recv(self, param_id, value, ps):
            ps.push(param_id, 0.5 * value + self.pending[param_id].popleft())

    host = transform(list(zip(keys.tolist(), vals.tolist())), DecayHost(),
                     ps_logic=LockPSLogicA(lambda _k: 0.0, overwrite), worker_parallelism=1, ps_parallelism=2,
                     iteration_wait_time=10, engine="host")
    hostm = last_write_wins(rights(host))
    np.testing.assert_allclose([float(hostm[k]) for k in range(n_keys)], want, rtol=1e-5, atol=1e-6)


def test_conflict_free_rounds_schedule():
    """The lock schedule of the device engine (CPU tensors work too): round r holds the records that r earlier
    records of the batch share their key with."""
    import torch
    from fps_b200.parallel.device_engine import conflict_free_rounds
    rounds = conflict_free_rounds(torch.tensor([3, 1, 3, 3, 2, 1, 0]))
    assert [r.tolist() for r in rounds] == [[0, 1, 4, 6], [2, 5], [3]]
    assert conflict_free_rounds(torch.tensor([], dtype=torch.int64)) == []
    one = conflict_free_rounds(torch.tensor([5, 6, 7]))
    assert len(one) == 1 and one[0].tolist() == [0, 1, 2]


def test_pruning_strategy_codes_of_the_c_abi():
    """LEMPPruningStrategy (object or the reference's string form) -> (code, focus set size, switch threshold)."""
    from fps_b200.models.mf import pruning as P
    from fps_b200.ops.topk import _strategy_code
    assert _strategy_code(P.LENGTH()) == (0, 0, 0.0)
    assert _strategy_code("coord") == (1, 0, 0.0)
    assert _strategy_code(P.INCR(7)) == (2, 7, 0.0)
    assert _strategy_code("lc:1.5") == (3, 0, 1.5)
    assert _strategy_code(P.LI(5, 2.5)) == (4, 5, 2.5)
    with pytest.raises(TypeError):
        _strategy_code(3)
    with pytest.raises(ValueError):
        _strategy_code("nearest")


def test_device_pruning_announces_only_what_is_not_executed():
    from fps_b200.models.mf.jobs import device_pruning
    from fps_b200.models.mf.pruning import COORD, LENGTH
    with warnings.catch_warnings():
        warnings.simplefilter("error", DE.DeviceFallbackWarning)
        assert device_pruning(LENGTH()) is True
        assert device_pruning(COORD(), candidate_filters=True) is True
    with pytest.warns(DE.DeviceFallbackWarning, match="candidate_filters=True"):
        assert device_pruning(COORD()) is True
