"""Restates ``T/WorkerLogicTest.scala:8-47`` plus the blocking limiter and the future API
(untested upstream)."""
import threading
import time

from fps_b200 import (ParameterServerClient, WorkerLogic, WorkerLogicWithFuture,
                      add_blocking_pull_limiter, add_pull_limiter, transform, lefts)


class CountingClient(ParameterServerClient):
    def __init__(self):
        self.pulls, self.pushes, self.outs = [], [], []

    def pull(self, i):
        self.pulls.append(i)

    def push(self, i, d):
        self.pushes.append((i, d))

    def output(self, o):
        self.outs.append(o)


class PullOnRecv(WorkerLogic):
    def on_recv(self, data, ps):
        ps.pull(data)

    def on_pull_recv(self, pid, val, ps):
        pass


def test_pull_limiter_limits_pulls():
    limited = add_pull_limiter(PullOnRecv(), 10)
    ps = CountingClient()
    for i in range(20):
        limited.on_recv(i, ps)
    assert len(ps.pulls) == 10
    for i in range(5):
        limited.on_pull_recv(i, 0.0, ps)
    assert len(ps.pulls) == 15
    limited.on_recv(20, ps)
    assert len(ps.pulls) == 15
    for i in range(5, 21):
        limited.on_pull_recv(i, 0.0, ps)
    assert len(ps.pulls) == 21
    # FIFO re-issue order and invariant
    assert ps.pulls == list(range(21))
    assert limited.max_in_flight <= 10


def test_pull_limiter_passes_push_and_output_through():
    class L(WorkerLogic):
        def on_recv(self, d, ps):
            ps.push(d, 1.0)
            ps.output(d)

        def on_pull_recv(self, *a):
            pass

    limited = add_pull_limiter(L(), 1)
    ps = CountingClient()
    for i in range(5):
        limited.on_recv(i, ps)
    assert len(ps.pushes) == 5 and len(ps.outs) == 5


def test_blocking_pull_limiter_blocks_second_thread():
    limited = add_blocking_pull_limiter(PullOnRecv(), 3)
    ps = CountingClient()
    done = threading.Event()

    def feeder():
        for i in range(6):
            limited.on_recv(i, ps)
        done.set()

    t = threading.Thread(target=feeder, daemon=True)
    t.start()
    time.sleep(0.2)
    assert len(ps.pulls) == 3 and not done.is_set()
    for i in range(3):
        limited.on_pull_recv(i, 0.0, ps)
    assert done.wait(2.0)
    assert len(ps.pulls) == 6
    assert limited.max_in_flight <= 3


def test_future_client_sends_pull_and_matches_fifo():
    got = []

    class F(WorkerLogicWithFuture):
        def on_data_recv(self, data, ps):
            fut = ps.pull(data % 2)
            fut.on_complete(lambda ans, d=data: got.append((d, ans)))
            fut.on_complete(lambda ans, d=data: ps.output((d, ans[1])))

    out = transform([0, 1, 2, 3], F(), lambda i: 100 + i, lambda a, b: a + b, 1, 2, 0)
    assert sorted(lefts(out)) == [(0, 100), (1, 101), (2, 100), (3, 101)]
    assert len(got) == 4
