"""Restates ``T/server/SimplePSLogicTest.scala``, ``LockPSLogicATest.scala`` (7 cases),
``LockPSLogicBTest.scala`` (8 cases) and adds RangePSLogicWithClose (untested upstream)."""
import pytest

from fps_b200 import ParameterServer, RuntimeContext
from fps_b200.server import (LockPSLogicA, LockPSLogicB, RangePSLogicWithClose, SimplePSLogic,
                             SimplePSLogicWithClose)


class MockPS(ParameterServer):
    def __init__(self):
        self.answers, self.outs = [], []

    def answer_pull(self, i, v, w):
        self.answers.append((i, v, w))

    def output(self, o):
        self.outs.append(o)


def test_simple_ps_logic_init_and_push():
    logic = SimplePSLogic(lambda i: i - 19, lambda a, b: a + b)
    ps = MockPS()
    logic.on_pull_recv(42, 0, ps)
    assert logic.params[42] == 23
    assert ps.answers == [(42, 23, 0)]
    logic.on_push_recv(42, 0, ps)
    assert ps.outs == [(42, 23)]
    # push to an absent id stores the delta itself
    logic.on_push_recv(7, 5, ps)
    assert logic.params[7] == 5 and ps.outs[-1] == (7, 5)


def test_simple_ps_logic_with_close_dumps_at_close():
    logic = SimplePSLogicWithClose(lambda i: 0, lambda a, b: a + b)
    ps = MockPS()
    logic.on_pull_recv(1, 0, ps)
    logic.on_push_recv(1, 4, ps)
    logic.on_push_recv(2, 9, ps)
    assert ps.outs == []
    logic.close(ps)
    assert sorted(ps.outs) == [(1, 4), (2, 9)]


def test_by_name_factories_rematerialised_per_open():
    made = []

    def factory():
        made.append(1)
        return lambda i: len(made)

    logic = SimplePSLogic(factory, lambda: (lambda a, b: a + b), by_name=True)
    logic.open({}, RuntimeContext(0, 1))
    assert len(made) == 2


@pytest.mark.parametrize("n,feature_count", [(3, 10), (4, 12), (3, 500000), (8, 47236)])
def test_range_ps_logic_slots(n, feature_count):
    import math
    div = math.ceil(feature_count / n)
    shards = []
    for s in range(n):
        lg = RangePSLogicWithClose(feature_count, lambda i: 0.0, lambda a, b: a + b)
        lg.open({}, RuntimeContext(s, n))
        shards.append(lg)
    assert sum(len(s.params) for s in shards) == feature_count
    ps = MockPS()
    for fid in (0, div - 1, div, feature_count - 1):
        sh = shards[fid // div]
        sh.on_pull_recv(fid, 1, ps)
        sh.on_push_recv(fid, 2.5, ps)
    for sh in shards:
        sh.close(ps)
    assert sorted(o[0] for o in ps.outs) == sorted({0, div - 1, div, feature_count - 1})
    assert all(o[1] == 2.5 for o in ps.outs)
    with pytest.raises(IndexError):
        shards[0].on_pull_recv(feature_count + 5, 0, ps)


def test_range_push_before_pull_stores_delta():
    lg = RangePSLogicWithClose(4, lambda i: 100.0, lambda a, b: a + b)
    lg.open({}, RuntimeContext(0, 1))
    ps = MockPS()
    lg.on_push_recv(2, 7.0, ps)
    lg.on_pull_recv(2, 0, ps)
    assert ps.answers == [(2, 7.0, 0)]


@pytest.mark.parametrize("cls", [LockPSLogicA, LockPSLogicB])
class TestLockCommon:
    def mk(self, cls):
        return cls(lambda i: i * 2, lambda a, b: a + b), MockPS()

    def test_push_before_pull_throws(self, cls):
        lg, ps = self.mk(cls)
        with pytest.raises(RuntimeError):
            lg.on_push_recv(1, 1, ps)

    def test_init(self, cls):
        lg, ps = self.mk(cls)
        lg.on_pull_recv(21, 3, ps)
        assert ps.answers == [(21, 42, 3)]
        assert lg.params[21].locked

    def test_update_and_emit(self, cls):
        lg, ps = self.mk(cls)
        lg.on_pull_recv(21, 3, ps)
        lg.on_push_recv(21, 8, ps)
        assert ps.outs == [(21, 50)]
        assert not lg.params[21].locked

    def test_second_pull_queued_not_answered(self, cls):
        lg, ps = self.mk(cls)
        lg.on_pull_recv(21, 3, ps)
        lg.on_pull_recv(21, 4, ps)
        assert len(ps.answers) == 1
        assert list(lg.params[21].queue) == [4]

    def test_push_releases_lock(self, cls):
        lg, ps = self.mk(cls)
        lg.on_pull_recv(21, 3, ps)
        lg.on_push_recv(21, 1, ps)
        lg.on_pull_recv(21, 4, ps)
        assert ps.answers[-1] == (21, 43, 4)

    def test_push_with_waiter_answers_new_value_keeps_lock(self, cls):
        lg, ps = self.mk(cls)
        lg.on_pull_recv(21, 3, ps)
        lg.on_pull_recv(21, 4, ps)
        lg.on_push_recv(21, 8, ps)
        assert ps.answers[-1] == (21, 50, 4)
        assert lg.params[21].locked and not lg.params[21].queue


def test_lock_a_keeps_duplicate_waiters():
    lg, ps = LockPSLogicA(lambda i: 0, lambda a, b: a + b), MockPS()
    lg.on_pull_recv(1, 0, ps)
    lg.on_pull_recv(1, 5, ps)
    lg.on_pull_recv(1, 5, ps)
    assert len(lg.params[1].queue) == 2


def test_lock_b_collapses_duplicate_waiters():
    lg, ps = LockPSLogicB(lambda i: 0, lambda a, b: a + b), MockPS()
    lg.on_pull_recv(1, 0, ps)
    lg.on_pull_recv(1, 5, ps)
    lg.on_pull_recv(1, 5, ps)
    lg.on_pull_recv(1, 6, ps)
    assert list(lg.params[1].queue) == [5, 6]


def test_lock_b_combined_push_sequence():
    lg, ps = LockPSLogicB(lambda i: 10, lambda a, b: a + b), MockPS()
    lg.on_pull_recv(1, 0, ps)
    lg.on_pull_recv(1, 1, ps)
    lg.on_pull_recv(1, 1, ps)
    lg.on_push_recv(1, 1, ps)     # answers worker 1 with 11, still locked
    lg.on_push_recv(1, 1, ps)     # queue empty -> unlock
    assert ps.answers == [(1, 10, 0), (1, 11, 1)]
    assert ps.outs == [(1, 11), (1, 12)]
    assert not lg.params[1].locked
