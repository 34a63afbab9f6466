"""Property tests of the host engine (hypothesis).  The reference declares ScalaCheck but never
writes a property; these pin the invariants every schedule must keep."""
import numpy as np
from hypothesis import given, settings, strategies as st

from fps_b200 import (WorkerLogic, add_pull_limiter, last_write_wins, rights, transform,
                      transform_with_model_load)
from fps_b200.utils.observability import engine_report

SCHEDULES = st.sampled_from(["round_robin", "random", "input_first", "feedback_first"])


class PullThenPush(WorkerLogic):
    """Every record pulls its key; every answer pushes +1 and reports what it saw."""

    def on_recv(self, data, ps):
        ps.pull(data)

    def on_pull_recv(self, pid, value, ps):
        ps.push(pid, 1)
        ps.output((pid, value))


@settings(max_examples=25, deadline=None)
@given(keys=st.lists(st.integers(0, 30), min_size=1, max_size=120), W=st.integers(1, 5), S=st.integers(1, 5),
       limit=st.integers(1, 12), schedule=SCHEDULES, seed=st.integers(0, 5))
def test_counts_are_schedule_independent_and_limiter_holds(keys, W, S, limit, schedule, seed):
    """Whatever the interleaving: the final value of a key is init + number of records with that key,
    every record is answered exactly once, and unanswered pulls per worker never exceed the limit."""
    out, eng = transform(keys, add_pull_limiter(PullThenPush(), limit), lambda i: 100 * i, lambda a, b: a + b,
                         W, S, 0, schedule=schedule, seed=seed, return_engine=True)
    final = last_write_wins(rights(out))
    counts = {}
    for k in keys:
        counts[k] = counts.get(k, 0) + 1
    assert final == {k: 100 * k + c for k, c in counts.items()}
    answers = [e.value for e in out if e.is_left]
    assert len(answers) == len(keys)
    assert engine_report(eng)["max_in_flight"] <= limit
    for pid, value in answers:                 # a read is stale at worst: between init and the final value
        assert 100 * pid <= value <= 100 * pid + counts[pid]


@settings(max_examples=15, deadline=None)
@given(n=st.integers(1, 40), W=st.integers(1, 4), S=st.integers(1, 4), rounds=st.integers(1, 3),
       schedule=SCHEDULES, seed=st.integers(0, 3))
def test_model_load_exactness_for_any_parallelism(n, W, S, rounds, schedule, seed):
    """`FlinkSimpleStackTest` "model load" generalised: ids preloaded with 10*i, each pulled and pushed +1
    `rounds` times -> exactly 10*i + rounds, for every worker / shard count and schedule."""
    model = [(i, 10 * i) for i in range(n)]
    data = [i for _ in range(rounds) for i in range(n)]
    from fps_b200.server.logics import SimplePSLogic
    out = transform_with_model_load(model)(data, PullThenPush(), SimplePSLogic(lambda i: -999, lambda a, b: a + b),
                                           worker_parallelism=W, ps_parallelism=S, iteration_wait_time=0,
                                           schedule=schedule, seed=seed)
    final = last_write_wins(rights(out))
    assert final == {i: 10 * i + rounds for i in range(n)}


@settings(max_examples=20, deadline=None)
@given(ids=st.lists(st.integers(0, 10_000), min_size=1, max_size=50), n_shards=st.integers(1, 7),
       n_ids=st.integers(1, 20_000))
def test_partition_formulas_round_trip(ids, n_shards, n_ids):
    """hash: owner = id % S, slot = id // S; range: owner = id // ceil(n/S), slot = id - owner*div -
    (owner, slot) identifies the id, owners are in range (the formulas of SURVEY §2.9)."""
    from fps_b200.server.logics import range_partition_div
    div = range_partition_div(max(n_ids, max(ids) + 1), n_shards)
    for i in ids:
        ho, hs = i % n_shards, i // n_shards
        assert hs * n_shards + ho == i and 0 <= ho < n_shards
        ro = i // div
        rs = i - ro * div
        assert ro * div + rs == i and 0 <= ro < n_shards and 0 <= rs < div


@settings(max_examples=10, deadline=None)
@given(seed=st.integers(0, 1000), dim=st.integers(1, 9))
def test_philox_init_is_a_pure_function_of_seed_stream_id(seed, dim):
    from fps_b200.ops import philox
    ids = np.array([0, 1, 7, 123456, 2**31 - 1], dtype=np.int64)
    a = philox.ranged_uniform(ids, dim, -0.5, 0.25, seed, philox.STREAM_ITEM)
    b = philox.ranged_uniform(ids[::-1].copy(), dim, -0.5, 0.25, seed, philox.STREAM_ITEM)[::-1]
    assert np.array_equal(a, b) and a.shape == (5, dim)
    assert np.all(a >= -0.5) and np.all(a < 0.25)
    other = philox.ranged_uniform(ids, dim, -0.5, 0.25, seed, philox.STREAM_USER)
    assert not np.array_equal(a, other)
