"""GPU tests of the device substrate and the fused MF kernel against plain fp32/fp64 references."""
import math
import random

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def native():
    from fps_b200.ops import _native
    if not _native.available():
        import build_native
        build_native.build()
    return _native


def test_table_eager_init_matches_philox(native):
    from fps_b200.ops.table import ShardedTable
    for dim, shards, part in [(10, 1, "hash"), (15, 3, "hash"), (64, 4, "range"), (128, 2, "hash"), (1, 3, "range")]:
        t = ShardedTable(1000, dim, n_shards=shards, partition=part, init_range=(-0.01, 0.01), seed=11)
        ids = torch.arange(1000)
        got = t.gather(ids).cpu().numpy()
        want = t.expected_init(np.arange(1000))
        np.testing.assert_allclose(got, want, rtol=0, atol=2e-9)
        did, drows = t.dump()
        assert did.tolist() == list(range(1000))
        np.testing.assert_allclose(drows.cpu().numpy(), want, rtol=0, atol=2e-9)


def test_table_lazy_life_cycle(native):
    """untouched rows are not dumped; first pull -> init(id); push before pull -> value = delta;
    push after pull -> init + delta; overwrite wins over both."""
    from fps_b200.ops.table import ShardedTable, new_flow, read_flow
    t = ShardedTable(100, 6, n_shards=3, partition="hash", init_range=(1.0, 2.0), seed=3, lazy=True)
    flow = new_flow()
    init = t.expected_init(np.arange(100))
    pulled = t.gather(torch.tensor([5, 7]), flow).cpu().numpy()
    np.testing.assert_allclose(pulled, init[[5, 7]], atol=1e-7)
    t.scatter(torch.tensor([7, 9]), torch.ones(2, 6) * 0.5, "add", flow)
    got = t.gather(torch.tensor([7, 9])).cpu().numpy()
    np.testing.assert_allclose(got[0], init[7] + 0.5, atol=1e-6)
    np.testing.assert_allclose(got[1], np.full(6, 0.5), atol=1e-7)       # delta became the value
    t.scatter(torch.tensor([5, 11]), torch.full((2, 6), 3.0), "overwrite")
    got = t.gather(torch.tensor([5, 11])).cpu().numpy()
    np.testing.assert_allclose(got, np.full((2, 6), 3.0), atol=1e-6)
    ids, rows = t.dump()
    assert ids.tolist() == [5, 7, 9, 11]
    s = read_flow(flow)
    assert s["pulls"] == 2 and s["pushes"] == 2 and s["bad_ids"] == 0
    t.gather(torch.tensor([1000, -1]), flow)
    assert read_flow(flow)["bad_ids"] == 2


def test_table_add_norm_column(native):
    from fps_b200.ops.table import ShardedTable
    t = ShardedTable(50, 10, n_shards=2, init_range=(0.0, 0.0), aux_cols=1)
    v = torch.arange(10, dtype=torch.float32)[None, :].repeat(3, 1)
    t.scatter(torch.tensor([3, 4, 40]), v, "add_norm")
    got = t.gather(torch.tensor([3, 40])).cpu().numpy()
    np.testing.assert_allclose(got[:, :10], v[:2].numpy(), atol=1e-6)
    np.testing.assert_allclose(got[:, 10], math.sqrt(float((v[0] ** 2).sum())), rtol=1e-6)


def serial_sgd_oracle(u, i, r, U, V, lr, lam=0.0):
    U, V = U.astype(np.float64).copy(), V.astype(np.float64).copy()
    for a, b, y in zip(u, i, r):
        e = y - U[a] @ V[b]
        du = lr * (e * V[b] - lam * U[a])
        dv = lr * (e * U[a] - lam * V[b])
        U[a] += du
        V[b] += dv
    return U, V


@pytest.mark.parametrize("dim,shards,lazy", [(10, 1, False), (15, 4, True), (64, 3, False), (128, 2, True), (200, 1, False)])
def test_mf_step_matches_serial_oracle_on_conflict_free_schedule(native, dim, shards, lazy):
    from fps_b200.models.mf.device import DeviceMF, conflict_free_waves
    rng = np.random.default_rng(dim)
    nu, ni, n = 300, 200, 3000
    u = rng.integers(0, nu, n).astype(np.int32)
    i = rng.integers(0, ni, n).astype(np.int32)
    r = rng.random(n).astype(np.float32)
    mf = DeviceMF(nu, ni, dim, 0.0, 0.3, 0.05, lambda_=0.01, lazy=lazy, n_shards=shards, seed=5)
    U0 = mf.users.expected_init(np.arange(nu))
    V0 = mf.items.expected_init(np.arange(ni))
    du, di, dr = (torch.from_numpy(x).cuda() for x in (u, i, r))
    order_u, order_i, order_r = [], [], []
    for wu, wi, wv in conflict_free_waves(du, di, dr):
        mf.step(wu, wi, wv)
        order_u += wu.tolist(); order_i += wi.tolist(); order_r += wv.tolist()
    torch.cuda.synchronize()
    U, V = serial_sgd_oracle(order_u, order_i, order_r, U0, V0, 0.05, 0.01)
    s = mf.stats()
    assert s["records"] == n and s["bad_ids"] == 0
    got_v = mf.items.gather(torch.arange(ni)).cpu().numpy()
    got_u = mf.users.gather(torch.arange(nu)).cpu().numpy()
    np.testing.assert_allclose(got_v, V, rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(got_u, U, rtol=2e-4, atol=2e-5)
    # predict == fp64 dot of the trained factors
    p = mf.predict(du[:100], di[:100]).cpu().numpy()
    np.testing.assert_allclose(p, np.einsum("ij,ij->i", U[u[:100]], V[i[:100]]), rtol=1e-3, atol=1e-4)


def test_offline_mf_device_reference_acceptance(native):
    """Same tiny config as the reference's own test: RMSE < 0.5, Hogwild mode."""
    from fps_b200.models.mf import Rating
    from fps_b200.models.mf.device import ps_offline_mf_device
    from fps_b200 import last_write_wins, lefts, rights
    rnd = random.Random(47)
    seen, ratings = set(), []
    for _ in range(100):
        a, b = rnd.randrange(20), rnd.randrange(15)
        if (a, b) not in seen:
            seen.add((a, b)); ratings.append(Rating(a, b, rnd.random()))
    out = ps_offline_mf_device(ratings, num_factors=15, range_min=0.0, range_max=1.0, learning_rate=0.01,
                               iterations=10, pull_limit=10, n_shards=4, seed=1)
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    se = sum((r.rating - float(users[r.user] @ items[r.item])) ** 2 for r in ratings)
    assert math.sqrt(se / len(ratings)) < 0.5


def test_device_agrees_with_host_engine_oracle(native):
    """The same job on the host engine (fp64, serial schedule) and on the device
    (conflict-free waves) ends in the same factors."""
    from fps_b200.models.mf import Rating, ps_offline_mf
    from fps_b200.models.mf.device import ps_offline_mf_device
    from fps_b200 import last_write_wins, lefts, rights
    rnd = random.Random(9)
    ratings = [Rating(rnd.randrange(30), rnd.randrange(25), rnd.random()) for _ in range(400)]
    kw = dict(num_factors=12, range_min=0.0, range_max=0.5, learning_rate=0.02, iterations=2, seed=4)
    host = ps_offline_mf(ratings, worker_parallelism=1, ps_parallelism=1, iteration_wait_time=10,
                         pull_limit=1, shuffle=False, schedule="feedback_first", **kw)
    dev = ps_offline_mf_device(ratings, shuffle=False, conflict_free=True, pull_limit=0, n_shards=1, **kw)
    hu, hi = last_write_wins(lefts(host)), last_write_wins(rights(host))
    du, di = last_write_wins(lefts(dev)), last_write_wins(rights(dev))
    assert set(hu) == set(du) and set(hi) == set(di)
    for k in hu:
        np.testing.assert_allclose(du[k], hu[k], rtol=5e-4, atol=5e-5)
    for k in hi:
        np.testing.assert_allclose(di[k], hi[k], rtol=5e-4, atol=5e-5)


def test_pull_limiter_is_a_device_credit_counter(native):
    from fps_b200.models.mf.device import DeviceMF
    g = torch.Generator().manual_seed(1)
    n = 200_000
    u = torch.randint(0, 5000, (n,), generator=g, dtype=torch.int32).cuda()
    i = torch.randint(0, 3000, (n,), generator=g, dtype=torch.int32).cuda()
    r = torch.rand(n, generator=g).cuda()
    for limit in (10, 500, 1600):
        mf = DeviceMF(5000, 3000, 32, 0.0, 0.1, 0.01, pull_limit=limit)
        mf.step(u, i, r)
        torch.cuda.synchronize()
        s = mf.stats()
        assert s["records"] == n and s["in_flight"] == 0
        assert 0 < s["max_in_flight"] <= limit, s
        # allowances are handed out per CTA and the grid is sized to fit the limit: nobody has to wait
        assert s["limiter_stalls"] == 0
    mf = DeviceMF(5000, 3000, 32, 0.0, 0.1, 0.01, pull_limit=0)
    mf.step(u, i, r)
    assert mf.stats()["max_in_flight"] == 0       # unlimited: limiter not engaged


def test_nan_guard_counts_instead_of_poisoning(native):
    from fps_b200.models.mf import FactorIsNotANumberException
    from fps_b200.models.mf.device import DeviceMF
    mf = DeviceMF(10, 10, 8, 0.0, 0.1, 0.01)
    u = torch.tensor([1, 2], dtype=torch.int32).cuda()
    i = torch.tensor([3, 4], dtype=torch.int32).cuda()
    r = torch.tensor([float("nan"), 1.0]).cuda()
    mf.step(u, i, r)
    with pytest.raises(FactorIsNotANumberException):
        mf.stats()
    assert torch.isfinite(mf.items.gather(torch.arange(10))).all()


@pytest.mark.parametrize("dim,shards", [(16, 1), (64, 3), (128, 2)])
def test_combine_variant_equals_oracle_and_combines_hot_keys(native, dim, shards):
    """Sorted, run-carrying kernel: exact on a conflict-free schedule; on a skewed batch it pulls
    far fewer item rows than it processes records and still learns like the direct kernel."""
    from fps_b200.models.mf.device import DeviceMF, conflict_free_waves
    rng = np.random.default_rng(dim)
    nu, ni, n = 400, 150, 3000
    u = rng.integers(0, nu, n).astype(np.int32)
    i = rng.integers(0, ni, n).astype(np.int32)
    r = rng.random(n).astype(np.float32)
    mf = DeviceMF(nu, ni, dim, 0.0, 0.3, 0.05, lambda_=0.01, n_shards=shards, seed=5, variant="combine")
    U0 = mf.users.expected_init(np.arange(nu))
    V0 = mf.items.expected_init(np.arange(ni))
    du, di, dr = (torch.from_numpy(x).cuda() for x in (u, i, r))
    ou, oi, orr = [], [], []
    for wu, wi, wv in conflict_free_waves(du, di, dr):
        mf.step(wu, wi, wv)
        # inside a wave every item is distinct, so the kernel's sorted order is a valid serial order
        order = torch.argsort(wi, stable=True)
        ou += wu[order].tolist(); oi += wi[order].tolist(); orr += wv[order].tolist()
    torch.cuda.synchronize()
    U, V = serial_sgd_oracle(ou, oi, orr, U0, V0, 0.05, 0.01)
    s = mf.stats()
    assert s["records"] == n and s["bad_ids"] == 0 and s["combined_rows"] == n
    np.testing.assert_allclose(mf.items.gather(torch.arange(ni)).cpu().numpy(), V, rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(mf.users.gather(torch.arange(nu)).cpu().numpy(), U, rtol=2e-4, atol=2e-5)

    # skewed Hogwild batch: compare with the direct kernel
    g = torch.Generator().manual_seed(3)
    m = 200_000
    uu = torch.randint(0, 20_000, (m,), generator=g, dtype=torch.int32).cuda()
    ii = (torch.rand(m, generator=g) ** 3 * 5_000).int().cuda()
    rr = (0.5 + 0.3 * torch.sin(uu.float()) * torch.cos(ii.float())).cuda()
    res = {}
    for variant in ("direct", "combine"):
        m2 = DeviceMF(20_000, 5_000, dim, 0.0, 0.5 / dim ** 0.5, 0.01, n_shards=shards, seed=9, variant=variant,
                      pull_limit=1600)
        for _ in range(8):
            m2.step(uu, ii, rr)
        torch.cuda.synchronize()
        res[variant] = (m2.rmse(uu, ii, rr), m2.stats(check_nan=False))
        assert res[variant][1]["nan_count"] == 0, (variant, res[variant][1])
    assert res["combine"][1]["records"] == 8 * m and res["combine"][1]["max_in_flight"] <= 1600
    assert res["combine"][1]["combined_rows"] < 0.8 * 8 * m           # hot keys really were combined
    assert abs(res["combine"][0] - res["direct"][0]) < 0.05, res
    assert res["combine"][0] < 0.4


def test_streaming_kernel_idle_timeout_and_results(native):
    """N2: persistent ring consumer exits by itself after iteration_wait_ms of idleness."""
    from fps_b200.models.mf.device import DeviceMF
    from fps_b200.models.mf.stream import OnlineStream
    mf = DeviceMF(2000, 1000, 16, 0.0, 0.2, 0.05, pull_limit=1600)
    g = torch.Generator().manual_seed(2)
    n = 50_000
    u = torch.randint(0, 2000, (n,), generator=g, dtype=torch.int32)
    i = torch.randint(0, 1000, (n,), generator=g, dtype=torch.int32)
    r = torch.rand(n, generator=g)
    before = mf.rmse(u.cuda(), i.cuda(), r.cuda())
    st = OnlineStream(mf, capacity=1 << 14, iteration_wait_ms=300, max_lifetime_ms=20_000)
    st.start()
    for k in range(0, n, 5000):
        st.feed(u[k:k + 5000], i[k:k + 5000], r[k:k + 5000])
    reason = st.join()
    s = mf.stats()
    assert reason == 1, reason                      # idle timeout, not the lifetime cap
    assert s["records"] == n and s["ring_done"] == n
    assert s["max_in_flight"] <= 1600
    assert mf.rmse(u.cuda(), i.cuda(), r.cuda()) < before


def test_streaming_consumer_is_rearmed_after_a_producer_gap(native):
    """A producer pause longer than iteration_wait_ms ends the consumer kernel (idle timeout); the next
    feed() re-arms it with ring_head / ring_done kept, so nothing is lost or processed twice."""
    import time
    from fps_b200.models.mf.device import DeviceMF
    from fps_b200.models.mf.stream import OnlineStream
    mf = DeviceMF(2000, 1000, 16, 0.0, 0.2, 0.05, pull_limit=0)
    g = torch.Generator().manual_seed(4)
    n = 30_000
    u = torch.randint(0, 2000, (n,), generator=g, dtype=torch.int32)
    i = torch.randint(0, 1000, (n,), generator=g, dtype=torch.int32)
    r = torch.rand(n, generator=g)
    st = OnlineStream(mf, capacity=1 << 13, iteration_wait_ms=100, max_lifetime_ms=20_000)
    st.feed(u[:10_000], i[:10_000], r[:10_000])
    time.sleep(0.6)                                  # > iteration_wait_ms: the kernel leaves
    st.feed(u[10_000:], i[10_000:], r[10_000:])      # 20 000 records through an 8 192-slot ring
    reason = st.join()
    s = mf.stats()
    assert st.relaunches >= 1
    assert reason == 1 and s["records"] == n and s["ring_done"] == n and s["nan_count"] == 0


# ---- tile variant: own bucketing + shared-memory sort / pre-combine + TMA-staged rows ---------------
@pytest.mark.parametrize("dim,shards", [(10, 1), (16, 3), (64, 2), (100, 1), (128, 2), (128, 3), (200, 1)])
def test_tiles_variant_equals_oracle_on_conflict_free_schedule(native, dim, shards):
    """Register path (dim < 125 or > 128) and the cp.async.bulk ring path (dim 100 / 128) against the
    serial fp64 rule; inside a wave every item is distinct, so item order is a valid serial order."""
    from fps_b200.models.mf.device import DeviceMF, conflict_free_waves
    rng = np.random.default_rng(dim + shards)
    nu, ni, n = 400, 150, 3000
    u = rng.integers(0, nu, n).astype(np.int32)
    i = rng.integers(0, ni, n).astype(np.int32)
    r = rng.random(n).astype(np.float32)
    mf = DeviceMF(nu, ni, dim, 0.0, 0.3, 0.05, lambda_=0.01, n_shards=shards, seed=5, variant="tiles")
    U0 = mf.users.expected_init(np.arange(nu))
    V0 = mf.items.expected_init(np.arange(ni))
    du, di, dr = (torch.from_numpy(x).cuda() for x in (u, i, r))
    ou, oi, orr = [], [], []
    for wu, wi, wv in conflict_free_waves(du, di, dr):
        mf.step(wu, wi, wv)
        order = torch.argsort(wi, stable=True)
        ou += wu[order].tolist(); oi += wi[order].tolist(); orr += wv[order].tolist()
    torch.cuda.synchronize()
    U, V = serial_sgd_oracle(ou, oi, orr, U0, V0, 0.05, 0.01)
    s = mf.stats()
    assert s["records"] == n and s["bad_ids"] == 0
    assert s["pulls"] == n and s["pushes"] == n          # every run has length one here
    np.testing.assert_allclose(mf.items.gather(torch.arange(ni)).cpu().numpy(), V, rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(mf.users.gather(torch.arange(nu)).cpu().numpy(), U, rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize("dim", [64, 128])
def test_tiles_single_walker_is_the_serial_rule_even_with_repeated_users(native, dim):
    """pull_limit=1 leaves one walker: the whole batch is one serial pass in item order.  Five users
    repeat all the time, so every prefetched user row is stale by the time it is used - the reload on
    a clash must make the result equal to the serial rule anyway; runs of equal items are carried."""
    from fps_b200.models.mf.device import DeviceMF
    rng = np.random.default_rng(dim)
    nu, ni, n = 5, 250, 750          # one tile (<= 768 records): the pass order is plain item order
    u = rng.integers(0, nu, n).astype(np.int32)
    i = np.repeat(rng.permutation(ni).astype(np.int32), 3)           # every item three times, adjacent
    r = rng.random(n).astype(np.float32)
    # equal items keep their input order only if the bucketing happens to; make the three records of an
    # item identical in (user, rating) so that any order among them is the same serial schedule
    u = np.repeat(u[::3], 3); r = np.repeat(r[::3], 3)
    mf = DeviceMF(nu, ni, dim, 0.0, 0.3, 0.05, lambda_=0.01, n_shards=1, seed=7, variant="tiles", pull_limit=1)
    U0 = mf.users.expected_init(np.arange(nu))
    V0 = mf.items.expected_init(np.arange(ni))
    mf.step(*(torch.from_numpy(x).cuda() for x in (u, i, r)))
    torch.cuda.synchronize()
    order = np.argsort(i, kind="stable")
    U, V = serial_sgd_oracle(u[order], i[order], r[order], U0, V0, 0.05, 0.01)
    s = mf.stats()
    # limit 1 = one record in flight: the lone walker publishes and refreshes after every record of a run
    assert s["records"] == n and s["pulls"] == ni + n and s["max_in_flight"] == 1 and s["in_flight"] == 0
    np.testing.assert_allclose(mf.items.gather(torch.arange(ni)).cpu().numpy(), V, rtol=5e-4, atol=5e-5)
    np.testing.assert_allclose(mf.users.gather(torch.arange(nu)).cpu().numpy(), U, rtol=5e-4, atol=5e-5)


@pytest.mark.parametrize("dim,shards", [(16, 1), (64, 3), (128, 2)])
def test_tiles_variant_combines_hot_keys_and_learns_like_direct(native, dim, shards):
    from fps_b200.models.mf.device import DeviceMF
    g = torch.Generator().manual_seed(3)
    m = 200_000
    uu = torch.randint(0, 20_000, (m,), generator=g, dtype=torch.int32).cuda()
    ii = (torch.rand(m, generator=g) ** 3 * 5_000).int().cuda()
    rr = (0.5 + 0.3 * torch.sin(uu.float()) * torch.cos(ii.float())).cuda()
    ev = torch.cuda.Event()
    ev.record()
    res = {}
    for variant, limit in (("direct", 0), ("tiles", 0), ("tiles", 1600), ("tiles", 100_000)):
        m2 = DeviceMF(20_000, 5_000, dim, 0.0, 0.5 / dim ** 0.5, 0.01, n_shards=shards, seed=9, variant=variant,
                      pull_limit=limit)
        for k in range(8):
            m2.step(uu, ii, rr, ready_event=ev if k % 2 else None)      # both orderings of the bucketing
        torch.cuda.synchronize()
        st = m2.stats(check_nan=False)
        res[(variant, limit)] = (m2.rmse(uu, ii, rr), st)
        assert st["nan_count"] == 0 and st["records"] == 8 * m and st["bad_ids"] == 0, (variant, limit, st)
        if limit:
            assert 0 < st["max_in_flight"] <= limit and st["in_flight"] == 0, st
    t = res[("tiles", 0)][1]
    assert t["pulls"] < 0.8 * 8 * m and t["pushes"] < 0.8 * 8 * m          # hot keys really were combined
    assert t["pulls"] == t["combined_rows"]
    # a hot item's run is walked serially with a refresh every 32 records, which is a more careful schedule
    # than the direct kernel's thousands of concurrent stale updates: it learns a little slower per step
    # (Hogwild runs are not repeatable: the bounds are loose on purpose, exactness is tested above)
    for key, (rmse, _) in res.items():
        assert abs(rmse - res[("direct", 0)][0]) < 0.15 and rmse < 0.45, {k: v[0] for k, v in res.items()}


def test_tiles_variant_counts_bad_ids_and_handles_odd_sizes(native):
    from fps_b200.models.mf.device import DeviceMF
    mf = DeviceMF(50, 40, 128, 0.0, 0.1, 0.01, variant="tiles")
    for n in (1, 31, 33, 1025, 2500):
        g = torch.Generator().manual_seed(n)
        u = torch.randint(0, 50, (n,), generator=g, dtype=torch.int32)
        i = torch.randint(0, 40, (n,), generator=g, dtype=torch.int32)
        u[0] = 50          # outside the user table
        if n > 1:
            i[1] = -3      # outside the item table
        mf.reset_stats()
        mf.step(u.cuda(), i.cuda(), torch.rand(n, generator=g).cuda())
        torch.cuda.synchronize()
        s = mf.stats()
        assert s["bad_ids"] == min(n, 2) and s["records"] == n - min(n, 2), (n, s)
    assert torch.isfinite(mf.items.gather(torch.arange(40))).all()


def test_step_from_host_pipelines_the_tile_path(native):
    from fps_b200.models.mf.device import DeviceMF
    g = torch.Generator().manual_seed(5)
    n = 50_000
    batches = []
    for _ in range(3):
        u = torch.randint(0, 3000, (n,), generator=g, dtype=torch.int32).pin_memory()
        i = torch.randint(0, 2000, (n,), generator=g, dtype=torch.int32).pin_memory()
        r = torch.rand(n, generator=g).pin_memory()
        batches.append((u, i, r))
    mf = DeviceMF(3000, 2000, 128, 0.0, 0.1, 0.02, variant="tiles")
    first = None
    for k in range(12):
        mf.step_from_host(*batches[k % 3])
    last = mf.flush_host_steps()
    s = mf.stats()
    assert s["records"] == 12 * n and s["nan_count"] == 0
    assert 0 < last < 0.5


def test_negative_sampler_respects_user_memory_and_pool(native):
    """K4 on the device: negatives come from the items seen so far, never from the user's last
    ``user_memory`` items; the positive closes its block; slots that cannot be filled are skipped."""
    from fps_b200.models.mf.device import DeviceMF, NegativeSampler
    from fps_b200.ops import _native as N
    mf = DeviceMF(50, 400, 8, 0.0, 0.1, 0.05, n_shards=3)
    rate, mem = 4, 6
    sm = NegativeSampler(mf, rate, mem, seed=3)
    rng = np.random.default_rng(0)
    history = {u: [] for u in range(50)}
    pool = set()
    for step in range(6):
        n = 50                                          # one record per user and batch: ring order = record order
        u = rng.permutation(50).astype(np.int32)
        i = rng.integers(0, 40 + 60 * step, n).astype(np.int32)
        r = rng.random(n).astype(np.float32)
        ou, oi, orr = (t.cpu().numpy() for t in sm.expand(*(torch.from_numpy(x).cuda() for x in (u, i, r))))
        pool |= set(i.tolist())
        for a, b in zip(u.tolist(), i.tolist()):
            history[a].append(b)
        assert sm.items_seen() == len(pool)
        # draw-major layout: slot d * n + k
        blocks_u, blocks_i, blocks_r = ou.reshape(rate + 1, n).T, oi.reshape(rate + 1, n).T, orr.reshape(rate + 1, n).T
        np.testing.assert_array_equal(blocks_u[:, rate], u)
        np.testing.assert_array_equal(blocks_i[:, rate], i)
        np.testing.assert_array_equal(blocks_r[:, rate], r)
        filled = 0
        for k in range(n):
            # the ring holds the user's last `mem` items as of the end of this batch (batch granularity)
            recent = set(history[int(u[k])][-mem:])
            for d in range(rate):
                if blocks_u[k, d] == N.SKIP_ID:
                    continue
                filled += 1
                assert blocks_u[k, d] == u[k] and blocks_r[k, d] == 0.0
                assert int(blocks_i[k, d]) in pool and int(blocks_i[k, d]) not in recent
        assert filled > 0.9 * n * rate
    # skipped slots are neither trained nor counted as bad ids
    ou, oi, orr = sm.expand(*(torch.from_numpy(x).cuda() for x in (u, i, r)))
    ou[::7] = N.SKIP_ID
    for variant in ("direct", "tiles"):
        m2 = DeviceMF(50, 400, 8, 0.0, 0.1, 0.05, n_shards=3, variant=variant)
        m2.step(ou, oi, orr)
        s = m2.stats()
        assert s["bad_ids"] == 0 and s["records"] == int((ou != N.SKIP_ID).sum()), (variant, s)


def test_ids_in_the_padding_of_the_last_shard_are_bad_ids(native):
    """n_ids % n_shards != 0: the slots past the last id exist in memory but are not parameters."""
    from fps_b200.ops.table import ShardedTable, new_flow, read_flow
    for part in ("hash", "range"):
        t = ShardedTable(10, 4, n_shards=3, partition=part, init_range=(1.0, 1.0))
        flow = new_flow()
        t.scatter(torch.tensor([9, 10, 11]), torch.ones(3, 4), "add", flow)
        s = read_flow(flow)
        assert s["bad_ids"] == 2 and s["pushes"] == 1, (part, s)
        ids, _ = t.dump()
        assert ids.tolist() == list(range(10))
