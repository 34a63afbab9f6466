"""Drivers (reference G5-G7), checkpoint text / npz round trips, observability helpers - all on CPU."""
import os
import random
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "drivers"))

from fps_b200 import Left, Right, WorkerLogic, add_pull_limiter, last_write_wins, lefts, rights, transform  # noqa: E402
from fps_b200.utils import checkpoint  # noqa: E402
from fps_b200.utils.observability import engine_report, flow_report, nvtx_range  # noqa: E402


def make_log(path, n_users=15, n_items=40, days=6, per_day=60, sep=" "):
    rnd = random.Random(4)
    t = 1_000_000
    with open(path, "w") as f:
        for d in range(days):
            for _ in range(per_day):
                u = rnd.randrange(n_users)
                i = (u % 4) * 10 + rnd.randrange(10)
                t += rnd.randrange(1, 2 * 86400 // per_day)
                f.write(sep.join(map(str, (t, u, i))) + "\n")


def test_checkpoint_text_and_npz_round_trip(tmp_path):
    stream = [Left((3, np.array([0.5, -1.25]))), Right((7, np.array([2.0, 3.0]))), Left((3, np.array([1.0, 2.0]))),
              Right((9, 4.5))]
    nl, nr = checkpoint.save_stream(stream, str(tmp_path / "u.map"), str(tmp_path / "i.map"))
    assert (nl, nr) == (1, 2)
    assert (tmp_path / "u.map").read_text().splitlines() == ["3;1.0", "3;2.0"]      # last write wins, one line per coordinate
    back = checkpoint.load_stream(str(tmp_path / "u.map"), str(tmp_path / "i.map"))
    l, r = checkpoint.fold(back)
    assert np.array_equal(l[3], [1.0, 2.0]) and np.array_equal(r[7], [2.0, 3.0]) and r[9].tolist() == [4.5]
    checkpoint.save_npz(str(tmp_path / "m.npz"), stream[:3])
    l2, r2 = checkpoint.fold(checkpoint.load_npz(str(tmp_path / "m.npz")))
    assert np.allclose(l2[3], [1.0, 2.0]) and np.allclose(r2[7], [2.0, 3.0])


def test_resume_from_text_checkpoint_continues_the_model(tmp_path):
    """dump -> id;value files -> transform_with_double_model_load equivalent on the host MF job."""
    from fps_b200.models.mf import Rating, ps_offline_mf
    rnd = random.Random(1)
    ratings = [Rating(rnd.randrange(12), rnd.randrange(9), rnd.random()) for _ in range(150)]
    out = ps_offline_mf(ratings, num_factors=4, range_min=0.0, range_max=0.4, learning_rate=0.05, iterations=3,
                        worker_parallelism=2, ps_parallelism=2, iteration_wait_time=10)
    checkpoint.save_stream(out, str(tmp_path / "u"), str(tmp_path / "i"))
    users, items = checkpoint.fold(checkpoint.load_stream(str(tmp_path / "u"), str(tmp_path / "i")))
    u0, i0 = last_write_wins(lefts(out)), last_write_wins(rights(out))
    assert users.keys() == u0.keys() and items.keys() == i0.keys()
    assert all(np.allclose(users[k], u0[k]) for k in u0) and all(np.allclose(items[k], i0[k]) for k in i0)


def test_driver_online_mf_implicit(tmp_path):
    import online_mf_implicit as drv
    log = tmp_path / "lastfm.txt"
    make_log(str(log))
    rc = drv.main([str(log), str(tmp_path / "UserVector.map"), str(tmp_path / "ItemVector.map"),
                   "--num-factors", "4", "--learning-rate", "0.1", "--negative-sample-rate", "2",
                   "--worker-parallelism", "2", "--ps-parallelism", "2", "--iteration-wait-time", "10"])
    assert rc == 0
    users = checkpoint.read_id_value(str(tmp_path / "UserVector.map"))
    items = checkpoint.read_id_value(str(tmp_path / "ItemVector.map"))
    assert len(users) == 15 and all(len(v) == 4 for v in users.values()) and len(items) > 20


def test_driver_learner_generator_ndcg_and_session_tools(tmp_path):
    import online_learner_generator_ndcg as drv
    import session_tools
    log = tmp_path / "week_all.csv"
    make_log(str(log), sep=",")
    out = tmp_path / "ndcg.csv"
    rc = drv.main([str(log), str(out), "--num-factors", "4", "--learning-rate", "0.2", "--negative-sample-rate", "2",
                   "--K", "10", "--worker-k", "10", "--bucket-size", "5", "--worker-parallelism", "2",
                   "--ps-parallelism", "2", "--pull-limit", "50"])
    assert rc == 0
    lines = out.read_text().splitlines()
    assert lines[0] == "period,ndcg,hit_rate,count" and lines[-1].startswith("all,") and len(lines) >= 4
    # notebook 1: split by days
    slog = tmp_path / "data_session"
    make_log(str(slog))
    stats = session_tools.split(str(slog), str(tmp_path / "ds"), train_days=4, test_days=2)
    assert stats["train_events"] > 0 and stats["test_users"] > 0 and stats["users_in_both"] <= stats["test_users"]
    train = np.loadtxt(str(tmp_path / "ds_train"), dtype=np.int64)
    test = np.loadtxt(str(tmp_path / "ds_test"), dtype=np.int64)
    assert train.shape[1] == 3 and test.shape[1] == 2 and np.all(np.diff(test[:, 0]) >= 0)
    # notebook 2: brute-force top-k recall / precision from dumped vectors
    rng = np.random.default_rng(0)
    checkpoint.write_id_value(str(tmp_path / "U.map"), {u: rng.normal(size=3) for u in range(15)})
    checkpoint.write_id_value(str(tmp_path / "I.map"), {i: rng.normal(size=3) for i in range(40)})
    res = session_tools.evaluate(str(tmp_path / "U.map"), str(tmp_path / "I.map"), str(tmp_path / "ds_train"),
                                 str(tmp_path / "ds_test"), k=5)
    assert res["active_users"] > 0 and 0.0 <= res["recall"] <= 1.0 and 0.0 <= res["precision"] <= 1.0


def test_observability_reports():
    class L(WorkerLogic):
        def on_recv(self, d, ps):
            ps.pull(d)

        def on_pull_recv(self, pid, v, ps):
            ps.push(pid, 1)

    with nvtx_range("host.transform"):
        out, eng = transform(range(40), add_pull_limiter(L(), 4), lambda i: 0, lambda a, b: a + b, 2, 2, 10,
                             schedule="input_first", return_engine=True)
    rep = engine_report(eng)
    assert rep["input_records"] == 40 and rep["max_in_flight"] == 4 and rep["stalled_pulls"] > 0
    fr = flow_report({"records": 100, "pulls": 100, "pushes": 100, "combined_rows": 40, "remote_pulls": 30,
                      "loss_sum": 25.0, "max_in_flight": 64}, link_row_bytes=512)
    assert fr["combine_ratio"] == 2.5 and fr["remote_fraction"] == 0.75 and fr["mean_loss"] == 0.25
    assert fr["nvlink_bytes_each_way"] == 30 * 512


def test_fold_accepts_a_generator():
    from fps_b200 import Left, Right
    from fps_b200.utils.checkpoint import fold
    stream = [Left((1, 1.0)), Right((7, 2.0)), Left((1, 3.0)), Right((8, 4.0))]
    lefts_, rights_ = fold(iter(stream))
    assert lefts_ == {1: 3.0} and rights_ == {7: 2.0, 8: 4.0}
