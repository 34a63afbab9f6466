"""GPU tests of the device job front-ends that compose kernels: top-K generator, prequential
learner+generator, model dump / load round trip (checkpoint-resume), P2P probes on one GPU."""
import random

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from fps_b200 import Left, Right, last_write_wins, lefts, rights  # noqa: E402
from fps_b200.models.mf import Rating, attach_length  # noqa: E402


def test_topk_generator_device_matches_brute_force():
    from fps_b200.models.mf.device_topk import ps_top_k_generator_device
    rng = np.random.default_rng(5)
    k, n_items, n_users = 16, 3000, 40
    items = rng.normal(size=(n_items, k)).astype(np.float32)
    users = rng.normal(size=(n_users, k)).astype(np.float32)
    model = [Left((u, attach_length(users[u]))) for u in range(n_users)] + \
            [Right((i, attach_length(items[i]))) for i in range(n_items)]
    queries = [Rating(u, 0, 1.0, 100 + u) for u in range(n_users)] + [Rating(n_users + 5, 0, 1.0, 7)]
    res = ps_top_k_generator_device(queries, model, num_factors=k, K=10, worker_k=10, user_memory=-1,
                                    n_users=n_users + 10)
    by_ts = {ts: topk for (_item, ts, topk) in res}
    for u in range(n_users):
        want = np.argsort(-(items.astype(np.float64) @ users[u].astype(np.float64)))[:10]
        got = [i for _s, i in by_ts[100 + u]]
        assert len(set(got) & set(want.tolist())) >= 9
        assert got[0] == want[0]
    assert by_ts[7] == []                       # unknown user -> empty recommendation


def test_learner_generator_device_prequential():
    from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
    from fps_b200.utils.evaluation import NDCGSink
    rnd = random.Random(2)
    # users have a favourite group of items; the model should learn to rank them
    stream = []
    for t in range(3000):
        u = rnd.randrange(50)
        item = (u % 5) * 20 + rnd.randrange(20)
        stream.append(Rating(u, item, 1.0, t))
    out, lg = ps_online_learner_and_generator_device(
        stream, n_users=50, n_items=100, batch_size=100, num_factors=8, range_min=-0.1, range_max=0.1,
        learning_rate=0.1, negative_sample_rate=3, K=10, worker_k=10, user_memory=4, pull_limit=500)
    assert len(out) == len(stream)
    early = NDCGSink().consume(out[200:700]).hit_rate
    late = NDCGSink().consume(out[-500:]).hit_rate
    assert late > early + 0.1 and late > 0.2, (early, late)      # random ranking would hit 10/100 = 0.1
    s = lg.stats()
    assert s["records"] > len(stream) and s["max_in_flight"] <= 500 and s["nan_count"] == 0


def test_dump_load_round_trip_resumes_training():
    """Checkpoint = the output stream; resume = load() (transform_with_double_model_load)."""
    from fps_b200.models.mf.device import DeviceMF
    g = torch.Generator().manual_seed(0)
    n = 20_000
    u = torch.randint(0, 500, (n,), generator=g, dtype=torch.int32).cuda()
    i = torch.randint(0, 300, (n,), generator=g, dtype=torch.int32).cuda()
    r = torch.rand(n, generator=g).cuda()
    a = DeviceMF(500, 300, 12, 0.0, 0.2, 0.05, lazy=True, n_shards=3)
    for _ in range(3):
        a.step(u, i, r)
    ckpt = a.dump()
    b = DeviceMF(500, 300, 12, -9.0, 9.0, 0.05, lazy=True, n_shards=2, seed=99)   # different init and sharding
    b.load(ckpt)
    ua, ia = last_write_wins(lefts(ckpt)), last_write_wins(rights(ckpt))
    ckpt_b = b.dump()
    ub, ib = last_write_wins(lefts(ckpt_b)), last_write_wins(rights(ckpt_b))
    assert ua.keys() == ub.keys() and ia.keys() == ib.keys()
    for k in ua:
        np.testing.assert_allclose(ub[k], ua[k], rtol=0, atol=1e-6)
    for k in ia:
        np.testing.assert_allclose(ib[k], ia[k], rtol=0, atol=1e-6)
    assert abs(a.rmse(u, i, r) - b.rmse(u, i, r)) < 1e-5


@pytest.mark.parametrize("streaming", [False, True])
def test_online_mf_device_job_with_negative_sampling(streaming):
    """Device twin of ps_online_mf: implicit positives + sampled negatives, micro-batched launches or
    the persistent ring consumer; positives must end up scored above random pairs."""
    from fps_b200.models.mf.device import ps_online_mf_device
    rnd = random.Random(3)
    stream = []
    for t in range(30000):
        u = rnd.randrange(200)
        stream.append(Rating(u, (u % 8) * 25 + rnd.randrange(25), 1.0, t))
    out, mf = ps_online_mf_device(stream, num_factors=8, range_min=-0.1, range_max=0.1, learning_rate=0.2,
                                  negative_sample_rate=3, batch_size=500, streaming=streaming,
                                  iteration_wait_time=500, return_model=True)
    s = mf.stats()
    assert s["records"] > len(stream) and s["nan_count"] == 0       # negatives were trained too
    assert s["max_in_flight"] <= 1600
    if streaming:
        assert s["exit_reason"] == 1 and s["ring_done"] == s["records"]
    users, items = last_write_wins(lefts(out)), last_write_wins(rights(out))
    pos = np.mean([users[r.user] @ items[r.item] for r in stream[-500:]])
    neg = np.mean([users[r.user] @ items[(r.item + 100) % 200] for r in stream[-500:] if (r.item + 100) % 200 in items])
    assert pos > neg + 0.1 and pos > 3 * abs(neg), (pos, neg)


def test_probes_run_on_one_gpu():
    import ctypes as C
    from fps_b200.ops import _native as N
    buf = N.DeviceBuffer(64 << 20)
    t = buf.tensor((64 << 18,))
    t.zero_()
    ms = C.c_float()
    for mode in (0, 1, 2):
        N.check(N.lib().fps_p2p_probe(buf.ptr, buf.ptr, 64 << 20, mode, 2, C.byref(ms), N.current_stream()))
        assert ms.value > 0
    torch.cuda.synchronize()
    assert float(t[:4].sum()) != 0.0            # stores / reds really landed


def test_streaming_checkpoint_round_trip_is_bit_exact(tmp_path):
    """save() / restore(): chunked pinned transfers straight from the dump kernel and into the scatter
    kernel, different sharding and initialisation on the way back, rows equal bit for bit."""
    from fps_b200.models.mf.device import DeviceMF
    g = torch.Generator().manual_seed(0)
    n = 60_000
    u = torch.randint(0, 30_000, (n,), generator=g, dtype=torch.int32).cuda()
    i = torch.randint(0, 20_000, (n,), generator=g, dtype=torch.int32).cuda()
    r = torch.rand(n, generator=g).cuda()
    for lazy in (True, False):
        a = DeviceMF(30_000, 20_000, 24, 0.0, 0.2, 0.05, lazy=lazy, n_shards=3)
        for _ in range(2):
            a.step(u, i, r)
        prefix = str(tmp_path / f"ckpt_{lazy}")
        counts = a.save(prefix, chunk_rows=4096)
        b = DeviceMF(30_000, 20_000, 24, -9.0, 9.0, 0.05, lazy=lazy, n_shards=2, seed=99)
        restored = b.restore(prefix)
        assert restored == counts
        if lazy:
            assert counts["users"] == int(torch.unique(u).numel()) and counts["items"] == int(torch.unique(i).numel())
        else:
            assert counts == {"users": 30_000, "items": 20_000}
        ids_u, ids_i = torch.unique(u).long(), torch.unique(i).long()
        assert torch.equal(a.users.gather(ids_u), b.users.gather(ids_u))
        assert torch.equal(a.items.gather(ids_i), b.items.gather(ids_i))
        assert abs(a.rmse(u, i, r) - b.rmse(u, i, r)) < 1e-6


def test_every_update_stream_folds_to_the_on_close_dump():
    """emit="every_update": one Right((item, vec)) per push and one Left((user, vec)) per rating in a device
    append buffer; folded last-write-wins it is the same model as the on-close dump."""
    from fps_b200.models.mf.device import DeviceMF
    from fps_b200.utils.checkpoint import fold
    g = torch.Generator().manual_seed(1)
    n = 3000
    mf = DeviceMF(200, 150, 12, 0.0, 0.2, 0.05, lazy=True, n_shards=2, emit="every_update", emit_capacity=2500)
    total = 0
    for step in range(4):
        u = torch.randint(0, 200, (n,), generator=g, dtype=torch.int32).cuda()
        i = torch.randint(0, 150, (n,), generator=g, dtype=torch.int32).cuda()
        mf.step(u, i, torch.rand(n, generator=g).cuda())
        total += n
    stream = mf.drain_updates(as_stream=True)
    assert sum(e.is_left for e in stream) == total and sum(e.is_right for e in stream) == total
    users_f, items_f = fold(stream)
    users_c, items_c = fold(mf.dump())
    assert users_f.keys() == users_c.keys() and items_f.keys() == items_c.keys()
    for k in users_c:
        np.testing.assert_array_equal(users_f[k], users_c[k])
    for k in items_c:
        np.testing.assert_array_equal(items_f[k], items_c[k])
    (uu, _), (ii, _) = mf.drain_updates()
    assert len(uu) == 0 and len(ii) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("lazy", [False, True])
def test_locked_read_modify_write_loses_no_update(lazy):
    """N13, the device form of LockPSLogicA (per-row lock word, system-scope CAS): 200 000 non-atomic
    read - add 1 - write cycles on 37 keys spread over 3 shards.  Under the lock every cycle counts; the same
    kernel without the lock loses updates, which is what the lock logic exists to prevent."""
    import torch
    from fps_b200.ops.table import ShardedTable, new_flow, read_flow
    n_keys, n, dim = 37, 200_000, 5
    g = torch.Generator().manual_seed(9)
    ids = torch.randint(0, n_keys, (n,), generator=g)
    ones = torch.ones((n, dim))
    counts = torch.bincount(ids, minlength=n_keys).float()

    t = ShardedTable(n_keys, dim, n_shards=3, partition="hash", init_range=(0.0, 0.0), aux_cols=1, lazy=lazy)
    flow = new_flow()
    t.locked_axpby(ids, 1.0, ones, flow=flow)
    torch.cuda.synchronize()
    rows = t.gather(torch.arange(n_keys))
    assert torch.equal(rows[:, :dim].cpu(), counts[:, None].expand(n_keys, dim))
    assert torch.all(rows[:, dim] == 0)                         # every lock word released
    st = read_flow(flow)
    assert st["pushes"] == n and st["bad_ids"] == 0
    if lazy:
        live, _ = t.dump()
        assert sorted(live.tolist()) == sorted(torch.nonzero(counts).flatten().tolist())

    # a decaying update is order dependent, but bounded: x <- 0.5 x + 1 stays in [1, 2)
    t.locked_axpby(ids[:5000], 0.5, ones[:5000])
    x = t.gather(torch.arange(n_keys))[:, :dim]
    assert torch.all((x >= 1.0) & (x < counts.max() + 1))

    free = ShardedTable(n_keys, dim, n_shards=3, partition="hash", init_range=(0.0, 0.0), aux_cols=1)
    free.locked_axpby(ids, 1.0, ones, use_lock=False)
    torch.cuda.synchronize()
    lost = counts.sum() - free.gather(torch.arange(n_keys))[:, 0].cpu().sum()
    assert lost > 0, "the unprotected cycle was expected to lose updates under this contention"

    with pytest.raises(ValueError):
        ShardedTable(8, 4, n_shards=1).locked_axpby(ids[:4] % 8, 1.0, ones[:4, :4])        # no aux column for the lock


@pytest.mark.gpu
def test_learner_generator_step_replayed_from_a_cuda_graph():
    """cuda_graph=True: the step sequence is captured after two eager steps of a batch size and replayed; the
    sampler's step tag lives on the device, so replays keep drawing fresh negatives and the job learns as the
    eager one does."""
    import random
    import torch
    from fps_b200.models.mf import Rating
    from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
    rnd = random.Random(2)
    stream = [Rating(u := rnd.randrange(40), (u % 4) * 10 + rnd.randrange(10), 1.0, t) for t in range(2000)]
    kw = dict(n_users=40, n_items=40, batch_size=50, num_factors=8, range_min=-0.1, range_max=0.1, learning_rate=0.1,
              negative_sample_rate=3, user_memory=4, K=10, worker_k=10, pull_limit=500)
    out_g, lg_g = ps_online_learner_and_generator_device(stream, cuda_graph=True, **kw)
    out_e, lg_e = ps_online_learner_and_generator_device(stream, cuda_graph=False, **kw)
    assert 50 in lg_g._graphs and not lg_e._graphs                 # 40 batches of 50: 2 eager, 38 replayed
    assert lg_g.batches == lg_e.batches == 40
    assert lg_g.stats()["records"] == lg_e.stats()["records"]     # same number of SGD records (negatives included)
    assert len(out_g) == len(out_e) == len(stream)
    hg, he = out_g.hit_rate(1000), out_e.hit_rate(1000)
    assert hg > 0.3 and abs(hg - he) < 0.12, (hg, he)
    assert int(lg_g._tag_dev.item()) == 40                          # the tag advanced on every replay


@pytest.mark.gpu
def test_learner_generator_accepts_columnar_input():
    """(users, items, ratings, timestamps) arrays instead of Rating objects: same job, same records."""
    import random
    from fps_b200.models.mf import Rating
    from fps_b200.models.mf.device_topk import ps_online_learner_and_generator_device
    rnd = random.Random(5)
    stream = [Rating(u := rnd.randrange(30), (u % 3) * 10 + rnd.randrange(10), 1.0, 100 + t) for t in range(600)]
    kw = dict(n_users=30, n_items=30, batch_size=64, num_factors=8, range_min=-0.1, range_max=0.1, learning_rate=0.1,
              negative_sample_rate=2, user_memory=4, K=5, worker_k=5, pull_limit=500, seed=3)
    out_a, _ = ps_online_learner_and_generator_device(stream, **kw)
    cols = (np.array([r.user for r in stream]), np.array([r.item for r in stream]), np.array([r.rating for r in stream]),
            np.array([r.timestamp for r in stream]))
    out_b, _ = ps_online_learner_and_generator_device(cols, **kw)
    assert len(out_a) == len(out_b) == 600
    assert out_b[7][:3] == (stream[7].user, stream[7].item, stream[7].timestamp)
    assert abs(out_a.hit_rate(300) - out_b.hit_rate(300)) < 0.15
