"""GPU tests of the tcgen05 LEMP top-K kernel, its CUDA-core twin and the merge kernel against a
plain fp32 / fp64 PyTorch reference."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def reference_topk(q, items, k):
    s = q.double() @ items.double().T
    v, i = torch.topk(s, k, dim=1)
    return v, i


def check_against_reference(out_s, out_i, q, items, k, ids=None, rtol=2e-5, atol=2e-5):
    ref_v, ref_i = reference_topk(q, items, k)
    got_s = out_s.double()
    # scores: sorted descending and equal to the reference's
    assert torch.all(got_s[:, :-1] >= got_s[:, 1:])
    torch.testing.assert_close(got_s, ref_v, rtol=rtol, atol=atol)
    # ids: every returned id really has the returned score (robust to ties / near-ties)
    idx = out_i.long()
    if ids is not None:
        inv = torch.empty(int(ids.max()) + 1, dtype=torch.long, device=ids.device)
        inv[ids.long()] = torch.arange(ids.numel(), device=ids.device)
        idx = inv[idx]
    true_scores = torch.gather(q.double() @ items.double().T, 1, idx)
    torch.testing.assert_close(true_scores, got_s, rtol=rtol, atol=atol)
    # and (almost) all of them are in the reference set
    same = (idx.sort(dim=1).values == ref_i.sort(dim=1).values).float().mean().item()
    assert same > 0.98, same


@pytest.mark.parametrize("dim,n_q,n_items,k", [(10, 37, 1000, 10), (64, 200, 5000, 75), (128, 128, 4096, 100),
                                                (128, 300, 20011, 50), (32, 1, 70, 5)])
def test_tensor_core_topk_matches_reference(dim, n_q, n_items, k):
    from fps_b200.ops.topk import lemp_topk, sort_items_by_length
    g = torch.Generator(device="cuda").manual_seed(dim + n_q)
    q = torch.randn(n_q, dim, device="cuda", generator=g)
    items = torch.randn(n_items, dim, device="cuda", generator=g) * (0.2 + 2.0 * torch.rand(n_items, 1, device="cuda", generator=g))
    it_s, len_s, ids_s = sort_items_by_length(items)
    stats = torch.zeros(4, device="cuda", dtype=torch.int64)
    out_s, out_i = lemp_topk(q, it_s, k, item_len=len_s, item_ids=ids_s, use_pruning=True, stats=stats)
    torch.cuda.synchronize()
    check_against_reference(out_s, out_i, q, items, k)
    st = stats.tolist()
    assert 0 < st[0] <= st[1]
    # without pruning: same answer, every tile scored
    stats.zero_()
    out_s2, out_i2 = lemp_topk(q, it_s, k, item_ids=ids_s, use_pruning=False, stats=stats)
    check_against_reference(out_s2, out_i2, q, items, k)
    assert stats[0].item() == stats[1].item()


@pytest.mark.parametrize("wide", [False, True])
@pytest.mark.parametrize("dim,n_q,n_items,k", [(10, 37, 1000, 10), (64, 200, 5000, 75), (128, 300, 20011, 100),
                                                (96, 129, 70, 5), (128, 1024, 100_000, 50)])
def test_packed_tma_pipeline_matches_reference(dim, n_q, n_items, k, wide):
    """Pre-packed tiles + TMA + warp-specialised tcgen05 pipeline: same answers as the reference,
    with and without the LEMP early exit."""
    from fps_b200.ops.topk import PackedItems, lemp_topk_packed, sort_items_by_length
    g = torch.Generator(device="cuda").manual_seed(dim * 7 + n_q)
    q = torch.randn(n_q, dim, device="cuda", generator=g)
    items = torch.randn(n_items, dim, device="cuda", generator=g) * (0.2 + 2.0 * torch.rand(n_items, 1, device="cuda", generator=g) ** 3)
    it_s, len_s, ids_s = sort_items_by_length(items)
    packed = PackedItems(it_s, len_s, ids_s, wide=wide)      # 64-item tiles / 128-item tiles with K-block stages
    stats = torch.zeros(4, device="cuda", dtype=torch.int64)
    for pruning in (True, False):
        stats.zero_()
        out_s, out_i = lemp_topk_packed(q, packed, k, use_pruning=pruning, stats=stats)
        torch.cuda.synchronize()
        check_against_reference(out_s, out_i, q, items, k)
        st = stats.tolist()
        assert 0 < st[0] <= st[1]
        if not pruning:
            assert st[0] == st[1]
    # repeated calls on the same packed image are stable
    a, _ = lemp_topk_packed(q, packed, k)
    b, _ = lemp_topk_packed(q, packed, k)
    torch.testing.assert_close(a, b)


def test_length_pruning_skips_tiles_on_skewed_lengths():
    """LEMP's point: with a long-tailed length distribution most buckets are never scored."""
    from fps_b200.ops.topk import lemp_topk, sort_items_by_length
    g = torch.Generator(device="cuda").manual_seed(0)
    dim, n_items = 64, 200_000
    lens = torch.rand(n_items, 1, device="cuda", generator=g) ** 8 * 10 + 0.01
    items = torch.nn.functional.normalize(torch.randn(n_items, dim, device="cuda", generator=g), dim=1) * lens
    q = torch.randn(256, dim, device="cuda", generator=g)
    it_s, len_s, ids_s = sort_items_by_length(items)
    stats = torch.zeros(4, device="cuda", dtype=torch.int64)
    out_s, out_i = lemp_topk(q, it_s, 20, item_len=len_s, item_ids=ids_s, stats=stats)
    check_against_reference(out_s, out_i, q, items, 20)
    scored, total = stats[0].item(), stats[1].item()
    assert scored < 0.5 * total, (scored, total)


def test_3xtf32_keeps_fp32_fidelity():
    """Plain TF32 would err at ~1e-3 relative; the split keeps scores at fp32 level."""
    from fps_b200.ops.topk import lemp_topk
    g = torch.Generator(device="cuda").manual_seed(5)
    q = torch.randn(128, 128, device="cuda", generator=g)
    items = torch.randn(2048, 128, device="cuda", generator=g)
    out_s, _ = lemp_topk(q, items, 16, use_pruning=False)
    ref_v, _ = reference_topk(q, items, 16)
    err = ((out_s.double() - ref_v).abs() / ref_v.abs().clamp_min(1.0)).max().item()
    assert err < 5e-6, err


def test_simt_twin_matches_reference():
    from fps_b200.ops.topk import topk_simt
    g = torch.Generator(device="cuda").manual_seed(9)
    q = torch.randn(150, 48, device="cuda", generator=g)
    items = torch.randn(3000, 48, device="cuda", generator=g)
    out_s, out_i = topk_simt(q, items, 30)
    check_against_reference(out_s, out_i, q, items, 30)


def test_merge_with_seen_filter():
    from fps_b200.ops.topk import topk_merge
    g = torch.Generator(device="cuda").manual_seed(2)
    n_q, parts, k_in = 9, 4, 75
    ids = torch.stack([torch.randperm(100000, device="cuda", generator=g)[: parts * k_in] for _ in range(n_q)]).int()
    scores = torch.randn(n_q, parts * k_in, device="cuda", generator=g)
    seen = torch.full((n_q, 8), -1, device="cuda", dtype=torch.int32)
    seen[:, 0] = ids[torch.arange(n_q), scores.argmax(dim=1)]       # hide every query's best item
    out_s, out_i = topk_merge(scores.view(n_q, parts, k_in), ids.view(n_q, parts, k_in), 100, seen)
    masked = scores.clone()
    masked[torch.arange(n_q), scores.argmax(dim=1)] = -float("inf")
    ref_v, ref_idx = torch.topk(masked, 100, dim=1)
    torch.testing.assert_close(out_s, ref_v)
    assert torch.equal(out_i.long(), torch.gather(ids.long(), 1, ref_idx))
    assert not (out_i == seen[:, :1]).any()


def test_in_place_table_scoring_masks_dead_slots():
    """``lemp_topk_table``: the worker-resident table is scored where it lies; slots whose live id is
    negative never appear, whatever their score."""
    from fps_b200.ops.topk import lemp_topk_table
    g = torch.Generator(device="cuda").manual_seed(11)
    n_rows, stride, dim, k = 5000, 12, 10, 20
    table = torch.randn(n_rows, stride, device="cuda", generator=g)
    table[:, dim:] = 7.0                                   # padding columns must not contribute
    live = torch.arange(n_rows, device="cuda", dtype=torch.int32) * 3 + 1      # slot -> id
    dead = torch.rand(n_rows, device="cuda", generator=g) < 0.4
    live[dead] = -1
    table[dead] *= 50.0                                    # dead slots would win every query if scored
    q = torch.randn(300, dim, device="cuda", generator=g)
    s, ids, _ = lemp_topk_table(q, table, dim, k, live)
    ref = q.double() @ table[:, :dim].double().T
    ref[:, dead] = -float("inf")
    want_s, want_slot = ref.topk(k, dim=1)
    torch.testing.assert_close(s.double(), want_s, rtol=2e-5, atol=2e-5)
    assert (ids >= 0).all()
    same = (ids.long() == live[want_slot].long()).float().mean().item()
    assert same > 0.99, same
    # nothing live at all -> empty lists
    s0, i0, _ = lemp_topk_table(q, table, dim, k, torch.full_like(live, -1))
    assert (i0 == -1).all() and torch.isinf(s0).all()


def test_wide_kernel_fill_plan_covers_query_tile_boundaries(monkeypatch):
    """FPS_TOPK_FILL=1: CTAs take equal shares of virtual splits and work in segments across query-tile
    boundaries (re-staging queries in TMEM); the static plan flag is read once per process, so this runs the
    check in a child process."""
    import os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch; sys.path.insert(0, %r)\n"
        "from fps_b200.ops.topk import PackedItems, lemp_topk_packed, sort_items_by_length\n"
        "g = torch.Generator(device='cuda').manual_seed(3)\n"
        "for n_q in (1500, 4000):\n"
        "    q = torch.randn(n_q, 96, device='cuda', generator=g)\n"
        "    items = torch.randn(60000, 96, device='cuda', generator=g) * (0.2 + 2 * torch.rand(60000, 1, device='cuda', generator=g) ** 3)\n"
        "    it_s, len_s, ids_s = sort_items_by_length(items)\n"
        "    packed = PackedItems(it_s, len_s, ids_s, wide=True)\n"
        "    ref = (q.double() @ items.double().T).topk(50, dim=1)\n"
        "    for prune in (False, True):\n"
        "        s, i = lemp_topk_packed(q, packed, 50, use_pruning=prune)\n"
        "        torch.cuda.synchronize()\n"
        "        assert torch.allclose(s.double(), ref.values, rtol=2e-5, atol=2e-5)\n"
        "        assert (i.long() == ref.indices).float().mean() > 0.99\n"
        "print('FILL_OK')\n" % root)
    env = dict(os.environ, FPS_TOPK_FILL="1")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=200, env=env)
    assert r.returncode == 0 and "FILL_OK" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def _skewed_problem(n_q, n_items, dim, seed):
    torch = pytest.importorskip("torch")
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(n_q, dim, device="cuda", generator=g)
    lens = torch.rand(n_items, 1, device="cuda", generator=g) ** 3 * 4 + 0.05
    items = torch.nn.functional.normalize(torch.randn(n_items, dim, device="cuda", generator=g), dim=1) * lens
    from fps_b200.ops.topk import sort_items_by_length
    return (q,) + sort_items_by_length(items)


@pytest.mark.parametrize("strategy", ["length", "coord", "incr:3", "lc:1.3", "li:3:1.3"])
def test_device_candidate_filters_match_host_predicates(strategy):
    """lengthPruning / coordPruning / incrPruning on the device against the host restatement
    (LEMPPruningFunctions.scala:10-90), bucket by bucket as PSTopKGeneratorWorker.scala:82-96 picks them."""
    torch = pytest.importorskip("torch")
    from fps_b200.models.mf import pruning as P
    from fps_b200.ops.topk import lemp_keep_mask
    n_q, n_items, dim, bucket = 6, 400, 16, 48
    q, it, ln, ids = _skewed_problem(n_q, n_items, dim, 11)
    scores = q.double() @ it.double().T
    theta = scores.topk(40, dim=1).values[:, -1].float()          # the 40th best score of every query
    keep = lemp_keep_mask(q, it, ln, theta, strategy, bucket).cpu().numpy()
    strat = P.from_string(strategy)
    qn, itn, lnn = q.double().cpu().numpy(), it.double().cpu().numpy(), ln.double().cpu().numpy()
    want = np.zeros_like(keep)
    for a in range(n_q):
        user = (float(np.linalg.norm(qn[a])), qn[a])
        focus, fset = P.focus_coordinate(qn[a]), P.focus_set(qn[a], strat.num_focus)
        for b0 in range(0, n_items, bucket):
            b1 = min(n_items, b0 + bucket)
            pred = P.bucket_filter(strat, user, float(theta[a]), lnn[b0], lnn[b1 - 1], focus, fset)
            for i in range(b0, b1):
                want[a, i] = pred((i, (lnn[i], itn[i])))
    mism = int((keep != want).sum())
    assert mism <= max(2, keep.size // 500), f"{strategy}: {mism} of {keep.size} predicates differ"
    if strategy != "length":
        assert 0 < want.sum() < want.size              # the filter does prune and does keep
    # safe filters: nothing that beats theta is ever dropped
    beats = (scores.cpu().numpy() > theta.cpu().numpy()[:, None] * (1 + 1e-6) + 1e-6)
    assert not (beats & ~keep).any()


@pytest.mark.parametrize("strategy", ["length", "coord", "incr:4", "lc:1.5", "li:5:2.5"])
def test_filtered_retrieval_is_exact_and_counts_like_the_host(strategy):
    torch = pytest.importorskip("torch")
    from fps_b200.models.mf import pruning as P
    from fps_b200.models.mf.workers import lemp_top_k
    from fps_b200.ops.topk import lemp_topk_filtered
    n_q, n_items, dim, k, bucket = 24, 3000, 24, 20, 64
    q, it, ln, ids = _skewed_problem(n_q, n_items, dim, 12)
    stats = torch.zeros(4, device="cuda", dtype=torch.int64)
    s, i = lemp_topk_filtered(q, it, ln, k, strategy, bucket, item_ids=ids, stats=stats)
    ref = (q.double() @ it.double().T).topk(k, dim=1)
    assert torch.equal(ids[ref.indices].sort(dim=1).values.int(), i.sort(dim=1).values)
    assert torch.allclose(s, ref.values.float(), rtol=1e-5, atol=1e-5)
    seen, pruned, scored, skipped = stats.tolist()
    assert seen == pruned + scored and scored < n_q * n_items            # the search stops early / prunes
    # the host engine's loop over the same buckets
    strat = P.from_string(strategy)
    itn, lnn, idn = it.double().cpu().numpy(), ln.double().cpu().numpy(), ids.cpu().numpy()
    model = {int(idn[j]): (float(lnn[j]), itn[j]) for j in range(n_items)}
    desc = [(float(lnn[j]), int(idn[j])) for j in range(n_items)]
    host = {}
    for a in range(n_q):
        qa = q[a].double().cpu().numpy()
        lemp_top_k((float(np.linalg.norm(qa)), qa), desc, model, k, bucket, strat, stats=host)
    assert abs(host.get("scored", 0) - scored) <= 0.03 * host["scored"] + 64, (host, stats.tolist())
    assert abs(host.get("pruned", 0) - pruned) <= 0.03 * max(host.get("pruned", 0), 1) + 64, (host, stats.tolist())


def test_in_place_scorer_through_the_packed_pipeline_matches_the_masked_kernel():
    """InPlaceScorer (pack with the live mask as an extra column + fps_lemp_topk_wide) against brute force over
    the live slots, including a query batch that spans query tiles, fewer live items than k, and rank 127."""
    from fps_b200.ops.topk import InPlaceScorer, lemp_topk_table
    g = torch.Generator(device="cuda").manual_seed(21)
    for n_rows, dim, n_q, k, p_live in [(5000, 10, 300, 75, 0.6), (700, 127, 40, 20, 0.5), (300, 16, 10, 50, 0.1)]:
        stride = ((dim + 3) // 4) * 4
        rows = torch.randn(n_rows, stride, device="cuda", generator=g)
        live = torch.rand(n_rows, device="cuda", generator=g) < p_live
        live_ids = torch.where(live, torch.arange(n_rows, device="cuda", dtype=torch.int32) + 1000, torch.full((n_rows,), -1, device="cuda", dtype=torch.int32))
        q = torch.randn(n_q, dim, device="cuda", generator=g)
        sc = InPlaceScorer(n_rows, dim)
        s, i = sc.topk(q, rows, k, live_ids)
        full = q.double() @ rows[:, :dim].double().T
        full[:, ~live] = -float("inf")
        n_live = int(live.sum())
        kk = min(k, n_live)
        ref_v, ref_i = full.topk(kk, dim=1)
        torch.testing.assert_close(s[:, :kk].double(), ref_v, rtol=2e-5, atol=2e-5)
        assert torch.equal(i[:, :kk].sort(dim=1).values, (ref_i + 1000).int().sort(dim=1).values) or \
            (i[:, :kk].sort(dim=1).values == (ref_i + 1000).int().sort(dim=1).values).float().mean() > 0.99
        if kk < k:
            assert torch.all(i[:, kk:] == -1) and torch.all(torch.isinf(s[:, kk:]))
        # second call after the table changed: the image is re-packed
        rows[:, :dim] *= 0.5
        s2, _ = sc.topk(q, rows, k, live_ids)
        torch.testing.assert_close(s2[:, :kk].double(), ref_v * 0.5, rtol=2e-5, atol=2e-5)
