"""Integration tests on the host engine: restates ``T/FlinkParameterServerTest.scala``,
``T/FlinkSimpleStackTest.scala`` (incl. "model load" == 10*i+3), ``T/FlinkCombinationStackTest.scala``,
``T/utils/FlinkEOFTest.scala`` and ``T/utils/FlinkSleepBlockerTest.scala``."""
import random
import time
from collections import deque

import pytest

from fps_b200 import (Left, ParameterServerLogic, PSReceiver, PSSender, Right, WorkerLogic,
                      WorkerReceiver, WorkerSender, add_pull_limiter, last_write_wins, lefts,
                      rights, transform, transform_generic, transform_with_double_model_load,
                      transform_with_model_load)
from fps_b200.client import (CombinationWorkerSender, CountClientSender, MultipleWorkerReceiver,
                             TimerClientSender)
from fps_b200.common import all_should_send, any_should_send
from fps_b200.parallel.streams import DataStream
from fps_b200.server import (CombinationPSSender, CountPSSender, MultiplePSReceiver, SimplePSLogic,
                             TimerPSSender)
from fps_b200.utils import EOFHandler, block, flat_map_with_eof

SCHEDULES = ["round_robin", "random", "input_first", "feedback_first"]


class PullThenPush(WorkerLogic):
    def on_recv(self, data, ps):
        ps.pull(data)

    def on_pull_recv(self, pid, val, ps):
        ps.push(pid, 1)


class DictPS(ParameterServerLogic):
    def __init__(self):
        self.params = {}

    def on_pull_recv(self, pid, widx, ps):
        ps.answer_pull(pid, self.params.setdefault(pid, 0), widx)

    def on_push_recv(self, pid, delta, ps):
        self.params[pid] = self.params.setdefault(pid, 0) + delta

    def close(self, ps):
        for kv in self.params.items():
            ps.output(kv)


@pytest.mark.parametrize("schedule", SCHEDULES)
def test_model_load_exact(schedule):
    n = 50
    model = DataStream.from_collection([(i, i * 10) for i in range(n)], 2)
    data = [i for i in range(n) for _ in range(3)]
    random.Random(1).shuffle(data)
    out = transform_with_model_load(model)(
        DataStream.from_collection(data, 3), PullThenPush(), DictPS(),
        lambda m: m.param_id % 3, lambda m: m.worker_partition_index, 4, 3, 4000,
        schedule=schedule, seed=3)
    assert not lefts(out)
    assert sorted(rights(out)) == [(i, i * 10 + 3) for i in range(n)]


def test_model_load_tolerates_empty_model_partition():
    out = transform_with_model_load([(0, 5)])(
        [0, 0], PullThenPush(), DictPS(), None, None, 4, 2, 100)
    assert sorted(rights(out)) == [(0, 7)]


@pytest.mark.parametrize("schedule", SCHEDULES)
def test_simple_stack_array_params(schedule):
    """P = list of floats; every id pulled once per record, pushed +[1,1]; SimplePSLogic emits
    one Right per push, folded last-write-wins."""
    class L(WorkerLogic):
        def on_recv(self, d, ps):
            ps.pull(d)

        def on_pull_recv(self, pid, val, ps):
            ps.output((pid, list(val)))
            ps.push(pid, [1.0, 1.0])

    data = [i % 7 for i in range(70)]
    out = transform(data, L(), lambda i: [float(i), 0.0],
                    lambda a, b: [x + y for x, y in zip(a, b)], 4, 3, 100, schedule=schedule)
    final = last_write_wins(rights(out))
    assert final == {i: [i + 10.0, 10.0] for i in range(7)}
    assert len(lefts(out)) == 70 and len(rights(out)) == 70


def test_custom_wire_format_generic_transform():
    """Fully custom wire types (tuples / strings), P = deque of ints - the shape of the
    reference's "flink mock PS" smoke test, with assertions added."""
    class WS(WorkerSender):
        def on_pull(self, pid, collect, part):
            collect(("pull", part, pid))

        def on_push(self, pid, delta, collect, part):
            collect(("push", part, pid, tuple(delta)))

    class PR(PSReceiver):
        def on_worker_msg(self, msg, on_pull, on_push):
            if msg[0] == "pull":
                on_pull(msg[2], msg[1])
            else:
                on_push(msg[2], deque(msg[3]))

    class PSS(PSSender):
        def on_pull_answer(self, pid, value, widx, collect):
            collect(f"{widx}:{pid}:{','.join(map(str, value))}")

    class WR(WorkerReceiver):
        def on_pull_answer_recv(self, msg, handler):
            from fps_b200 import PullAnswer
            w, pid, vals = msg.split(":")
            handler(PullAnswer(int(pid), deque(int(v) for v in vals.split(",") if v)))

    class L(WorkerLogic):
        def on_recv(self, d, ps):
            ps.pull(d)

        def on_pull_recv(self, pid, val, ps):
            ps.push(pid, deque([len(val)]))

    out = transform_generic(
        list(range(12)) * 2, L(), SimplePSLogic(lambda i: deque([i]), lambda a, b: deque(list(a) + list(b))),
        lambda m: m[2], lambda m: int(m.split(":")[0]), 3, 2,
        WR(), WS(), PR(), PSS(), 100)
    final = last_write_wins(rights(out))
    assert set(final) == set(range(12))
    assert all(len(v) == 3 and v[0] == k for k, v in final.items())


def test_bad_feedback_key_raises():
    with pytest.raises(RuntimeError, match="partition ID"):
        transform([1], PullThenPush(), ps_logic=DictPS(), w_in_partition=lambda m: 99,
                  worker_parallelism=2, ps_parallelism=1, iteration_wait_time=10)


def test_combination_stack_batched_both_ways():
    """Client AND(count 4, timer), server OR(count, timer); realtime engine."""
    ws = CombinationWorkerSender(all_should_send, [CountClientSender(4), TimerClientSender(0.05)])
    pss = CombinationPSSender(any_should_send, [CountPSSender(3), TimerPSSender(0.05)])
    data = [i % 5 for i in range(40)]
    out, eng = transform_generic(
        data, PullThenPush(), DictPS(), lambda arr: arr[0].param_id if False else 0,
        lambda arr: arr[0].worker_partition_index, 2, 1,
        MultipleWorkerReceiver(), ws, MultiplePSReceiver(), pss, 600, return_engine=True)
    assert eng.realtime
    assert sorted(rights(out)) == [(i, 8) for i in range(5)]
    # batching really happened
    assert eng.stats.worker_to_ps_msgs < 80


def test_pull_limiter_in_engine_bounds_in_flight():
    limited = add_pull_limiter(PullThenPush(), 3)
    out, eng = transform(list(range(100)), limited, ps_logic=DictPS(), worker_parallelism=2,
                         ps_parallelism=2, iteration_wait_time=10, schedule="input_first",
                         return_engine=True)
    assert sorted(rights(out)) == [(i, 1) for i in range(100)]
    for lg in eng.worker_logics:
        assert lg.max_in_flight == 3 and lg.stalled_pulls > 0 and lg.pull_counter == 0


def test_eof_barrier():
    src_par, per_src = 7, 5
    src = DataStream.from_partitions([[i] * per_src for i in range(src_par)])

    class H(EOFHandler):
        def open(self, idx, par):
            self.idx, self.par, self.sum, self.eof = idx, par, 0, False

        def flat_map(self, v):
            assert not self.eof and v % self.par == self.idx
            self.sum += v
            return ()

        def on_eof(self):
            self.eof = True
            return (f"EOF {self.sum}",)

    for down in (1, 3):
        src = DataStream.from_partitions([[i] * per_src for i in range(src_par)])
        got = flat_map_with_eof(src, H, down, lambda k, n: k % n, lambda x: x).collect()
        assert len(got) == down
        assert sum(int(g.split()[1]) for g in got) == sum(range(src_par)) * per_src


def test_eof_barrier_empty_source_ok():
    src = DataStream.from_partitions([[], [1, 2]])

    class H(EOFHandler):
        def flat_map(self, v):
            return (v,)

        def on_eof(self):
            return ("EOF",)

    assert flat_map_with_eof(src, H, 1, lambda k, n: 0, lambda x: x).collect() == [1, 2, "EOF"]


def test_sleep_blocker():
    t0 = time.time()
    first = next(iter(block([1, 2, 3], 150).partitions()[0]))
    dt = time.time() - t0
    assert first == 1 and 0.9 * 0.15 <= dt <= 1.5 * 0.15 + 0.2
    t0 = time.time()
    assert block([1, 2, 3], 50).collect() == [1, 2, 3]
    assert time.time() - t0 < 0.3      # sleeps once, not per element


def test_double_model_load_places_rows_and_buffers_pulls():
    class TopWorker(WorkerLogic):
        def __init__(self):
            self.model = {}

        def update_model(self, pid, p):
            self.model[pid] = p

        def on_recv(self, q, ps):
            ps.pull(q)

        def on_pull_recv(self, pid, val, ps):
            ps.output((pid, val, sorted(self.model)))

    model = [Left((u, 100 + u)) for u in range(6)] + [Right((i, -i)) for i in range(8)]
    out, eng = transform_with_double_model_load(model, worker_owner=lambda i, w: i % w)(
        DataStream.from_collection([0, 1, 2, 3, 4, 5], 1).broadcast(2), TopWorker(),
        SimplePSLogic(lambda i: -1, lambda old, new: new), None, None, 2, 3, 100,
        return_engine=True)
    res = lefts(out)
    assert len(res) == 12
    for pid, val, items in res:
        assert val == 100 + pid          # loaded value, never the -1 initialiser
    owned = sorted(tuple(lg.inner.model) for lg in eng.worker_logics)
    assert owned == [(0, 2, 4, 6), (1, 3, 5, 7)]
