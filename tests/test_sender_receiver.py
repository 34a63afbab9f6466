"""Restates ``T/SenderReceiverTest.scala:41-205`` (intervals shortened from 5 s to 0.4 s)."""
import time

import pytest

from fps_b200.client import (CombinationWorkerSender, CountClientSender, MultipleWorkerReceiver,
                             SimpleWorkerReceiver, SimpleWorkerSender, TimerClientSender)
from fps_b200.common import all_should_send, any_should_send
from fps_b200.entities import PSToWorker, PullAnswer
from fps_b200.server import (CombinationPSSender, CountPSSender, MultiplePSReceiver,
                             SimplePSReceiver, SimplePSSender)

INTERVAL = 0.4


class Sink:
    def __init__(self):
        self.pull_map, self.push_map = {}, {}

    def one(self, m):
        if m.msg.is_left:
            self.pull_map[m.worker_partition_index] = m.msg.value
        else:
            self.push_map[m.worker_partition_index] = m.msg.value

    def many(self, arr):
        for m in arr:
            self.one(m)


def eventually(cond, timeout):
    end = time.time() + timeout
    while time.time() < end:
        if cond():
            return True
        time.sleep(0.02)
    return cond()


def test_simple_client_sender():
    s, sender = Sink(), SimpleWorkerSender()
    for i in range(1, 4):
        sender.on_pull(i, s.one, i)
    for i in range(4, 11):
        sender.on_push(i, float(i), s.one, i)
    assert sorted(s.pull_map) == [1, 2, 3]
    assert sorted(s.push_map) == list(range(4, 11))


def test_counter_client_sender():
    s, send_after = Sink(), 3
    sender = CombinationWorkerSender(lambda cs: cs[0].should_send(), [CountClientSender(send_after)])
    for i in range(1, send_after):
        sender.on_pull(i, s.many, i)
    assert not s.pull_map and not s.push_map
    sender.on_pull(send_after, s.many, send_after)
    assert len(s.pull_map) == send_after
    for i in range(1, send_after + 1):
        sender.on_push(i, float(i), s.many, i)
    assert len(s.push_map) == send_after
    sender.on_push(send_after + 1, 0.0, s.many, send_after + 1)
    assert len(s.push_map) == send_after


def test_timer_client_sender():
    s = Sink()
    sender = CombinationWorkerSender(lambda cs: cs[0].should_send(), [TimerClientSender(INTERVAL)])
    try:
        sender.on_pull(1, s.many, 1)
        assert len(s.pull_map) == 0
        assert eventually(lambda: len(s.pull_map) == 1, 2 * INTERVAL + 1)
        s.pull_map.clear()
        for i in range(1, 6):
            sender.on_pull(i, s.many, i)
            sender.on_push(i, float(i), s.many, i)
        assert eventually(lambda: len(s.pull_map) == 5 and len(s.push_map) == 5, 2 * INTERVAL + 1)
    finally:
        sender.shutdown()


def test_combined_counter_or_timer():
    s, limit = Sink(), 5
    sender = CombinationWorkerSender(any_should_send, [CountClientSender(limit), TimerClientSender(INTERVAL)])
    try:
        for i in range(1, limit):
            sender.on_push(i, float(i), s.many, i)
        assert len(s.push_map) == 0
        sender.on_push(limit, float(limit), s.many, limit)
        assert len(s.push_map) == limit
        sender.on_pull(1, s.many, 1)
        assert len(s.pull_map) == 0
        assert eventually(lambda: len(s.pull_map) == 1, 2 * INTERVAL + 1)
    finally:
        sender.shutdown()


def test_combined_counter_and_timer_sticky_flags():
    """The subtle case: 1 pull + (limit - 1) pushes are delivered."""
    s, limit = Sink(), 5
    sender = CombinationWorkerSender(all_should_send, [CountClientSender(limit), TimerClientSender(INTERVAL)])
    try:
        sender.on_pull(1, s.many, 1)
        time.sleep(INTERVAL + 0.3)
        assert len(s.pull_map) == 0
        for i in range(1, limit + 1):
            sender.on_push(i, float(i), s.many, i)
        assert eventually(lambda: len(s.pull_map) == 1 and len(s.push_map) == limit - 1, 2)
        time.sleep(0.1)
        assert len(s.push_map) == limit - 1
    finally:
        sender.shutdown()


def test_ps_side_senders_and_receivers():
    got = []
    SimplePSSender().on_pull_answer(3, 1.5, 2, got.append)
    assert got == [PSToWorker(2, PullAnswer(3, 1.5))]
    batches = []
    snd = CombinationPSSender(any_should_send, [CountPSSender(2)])
    snd.on_pull_answer(1, 1.0, 0, batches.append)
    assert batches == []
    snd.on_pull_answer(2, 2.0, 0, batches.append)
    assert len(batches) == 1 and len(batches[0]) == 2
    answers = []
    MultipleWorkerReceiver().on_pull_answer_recv(batches[0], answers.append)
    assert [a.param_id for a in answers] == [1, 2]
    SimpleWorkerReceiver().on_pull_answer_recv(got[0], answers.append)
    assert answers[-1] == PullAnswer(3, 1.5)

    pulls, pushes, wire = [], [], []
    ws = SimpleWorkerSender()
    ws.on_pull(9, wire.append, 4)
    ws.on_push(9, 0.5, wire.append, 4)
    SimplePSReceiver().on_worker_msg(wire[0], lambda i, w: pulls.append((i, w)), lambda i, d: pushes.append((i, d)))
    MultiplePSReceiver().on_worker_msg(wire, lambda i, w: pulls.append((i, w)), lambda i, d: pushes.append((i, d)))
    assert pulls == [(9, 4), (9, 4)] and pushes == [(9, 0.5)]
