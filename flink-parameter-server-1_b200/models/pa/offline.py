"""Offline (multi-epoch) passive-aggressive training followed by test-set prediction.

Parity: ``M/passive/aggressive/classification/binary/PABinaryClassificationOffline.scala:47-387``
(and its copy ``classification/multi/PAMultiClassificationOffline.scala``, which upstream is
still the *binary* job - SURVEY Q7; ``pa_multi_classification_offline`` here is a genuine
multiclass job built on the one-vs-all / cost-based rules).

Protocol kept from the reference: training and test sources each end with an EOF per worker;
a worker stores its training vectors, on the training EOF runs ``iterations`` epochs - pull
every active feature of a vector, update when all answers are in, push the deltas - with the
pull limit counted in *features*; when the last epoch has drained and the test EOF has been
seen it switches to prediction.  Differences: epochs are driven by a state machine on the
engine thread instead of a second thread + busy-wait (``:230-294``); predictions are emitted as
worker outputs ``(label, prediction, vector)`` instead of only being logged (``:353-358``).
"""
from __future__ import annotations

import random
from collections import deque
from typing import Deque, Dict, List, Optional

import numpy as np

from ...api import ParameterServerClient, WorkerLogic
from ...entities import EOF, Left, Right
from ...parallel.streams import DataStream, as_stream
from ...parameter_server import transform
from ...server import SimplePSLogic
from ...worker_logic import add_pull_limiter
from .algorithm import (PassiveAggressiveBinaryAlgorithm, PassiveAggressiveFilter,
                        PassiveAggressiveFilterI, PassiveAggressiveFilterII,
                        PassiveAggressiveMulticlassAlgorithm, PassiveAggressiveOneVersusAll)
from .sparse import LegacySparseVector, SparseVector

_TRAIN_EOF = ("train_eof",)
_TEST_EOF = ("test_eof",)


def _to_sparse(vec, feature_count: int = 0) -> SparseVector:
    if isinstance(vec, SparseVector):
        return vec
    if isinstance(vec, LegacySparseVector):
        d = vec.get_values()
        keys = sorted(d)
        return SparseVector(keys, [d[k] for k in keys], max(feature_count, (max(keys) + 1) if keys else 0))
    raise TypeError(type(vec))


class _Job:
    __slots__ = ("vec", "label", "answers", "train")

    def __init__(self, vec, label, train):
        self.vec, self.label, self.train, self.answers = vec, label, train, {}


class PAOfflineWorker(WorkerLogic):
    def __init__(self, method, iterations: int, shuffle: bool = True, seed: int = 0):
        self.method, self.iterations, self.shuffle, self.seed = method, iterations, shuffle, seed
        self.train: List[tuple] = []
        self.test: List[tuple] = []
        self.train_eof = self.test_eof = False
        self.epoch = 0
        self.outstanding = 0
        self.predicting = False
        self.waiting: Dict[int, Deque[_Job]] = {}
        self._rng = random.Random(seed)
        self.predictions: List[tuple] = []

    def open(self, ctx=None):
        self._rng = random.Random((self.seed << 8) ^ (ctx.index_of_this_subtask if ctx else 0))

    def _issue(self, vec, label, train, ps):
        job = _Job(vec, label, train)
        self.outstanding += 1
        if vec.active_size == 0:
            self._complete(job, ps)
            return
        for k in vec.indices.tolist():
            self.waiting.setdefault(k, deque()).append(job)
            ps.pull(k)

    def _start_epoch(self, ps):
        self.epoch += 1
        data = list(self.train)
        if self.shuffle:
            self._rng.shuffle(data)
        if not data:
            self._advance(ps)
            return
        for vec, label in data:
            self._issue(vec, label, True, ps)

    def _advance(self, ps):
        if self.outstanding:
            return
        if self.epoch < self.iterations:
            self._start_epoch(ps)
        elif self.test_eof and not self.predicting:
            self.predicting = True
            for vec, label in self.test:
                self._issue(vec, label, False, ps)

    def on_recv(self, data, ps):
        if data is _TRAIN_EOF:
            self.train_eof = True
            self._advance(ps)
        elif data is _TEST_EOF:
            self.test_eof = True
            if self.train_eof:
                self._advance(ps)
        elif data.is_left:
            if self.train_eof:
                raise RuntimeError("training record after the training EOF")
            self.train.append(data.value)
        else:
            self.test.append(data.value)

    def on_pull_recv(self, param_id, value, ps):
        q = self.waiting[param_id]
        job = q.popleft()
        if not q:
            del self.waiting[param_id]
        job.answers[param_id] = value
        if len(job.answers) == job.vec.active_size:
            self._complete(job, ps)

    def _complete(self, job, ps):
        if job.train:
            for i, d in self.method.delta(job.vec, job.answers, job.label):
                ps.push(i, d)
        else:
            pred = self.method.predict(job.vec, job.answers)
            self.predictions.append((job.label, pred))
            ps.output((job.label, pred, job.vec))
        self.outstanding -= 1
        if self.train_eof:
            self._advance(ps)


def _offline_job_device(training_src, test_src, ps_parallelism, method, iterations, pull_limit, seed, shuffle,
                        feature_count, batch_size: int = 4096, return_model: bool = False):
    """The offline job on the device engine (``PABinaryClassificationOffline.scala:47-387``): the training
    set is turned into CSR mini-batches once and stays on the device, every epoch launches the fused
    gather / margin / scatter kernel over the batches (order reshuffled per epoch), then the test set is
    scored with the final weights.  The feature-counted limiter (``:230-249``) is the kernels' credit
    pool.  Output: ``Left((label, prediction, vector))`` per test record + the on-close weight dump."""
    import torch
    from ...parallel.device_context import get_context
    from .device import DevicePA, method_to_device, to_csr
    inner = getattr(method, "inner", method)
    binary_pm1 = inner is not method                      # +1 / -1 labels of the binary offline job
    train = [(_to_sparse(v, feature_count), l) for v, l in as_stream(training_src, 1).collect()]
    test = [(_to_sparse(v, feature_count), l) for v, l in as_stream(test_src, 1).collect()]
    if not feature_count:
        feature_count = 1 + max([int(v.indices.max()) for v, _ in train + test if v.active_size] + [0])
    kw = method_to_device(inner)
    ctx = get_context()
    shards = None if ctx.distributed else max(1, min(int(ps_parallelism), 16))
    pa = DevicePA(feature_count, label_count=kw["label_count"], algorithm=kw["algorithm"], variant=kw["variant"],
                  aggressiveness=kw["aggressiveness"], pull_limit=pull_limit, cost=kw["cost"], n_shards=shards)

    def batches(records):
        out = []
        for s in range(0, len(records), batch_size):
            part = records[s:s + batch_size]
            indptr, indices, values = to_csr([v for v, _ in part])
            lab = [l for _, l in part]
            if kw["algorithm"] == "binary":
                lab_t = torch.tensor([1 if (l is True or l > 0) else -1 for l in lab], dtype=torch.int32).cuda()
            else:
                lab_t = torch.tensor([int(l) for l in lab], dtype=torch.int32).cuda()
            out.append((torch.from_numpy(indptr).cuda(), torch.from_numpy(indices).cuda(),
                        torch.from_numpy(values).cuda(), lab_t, part))
        return out

    rnd = random.Random(seed)
    train_b = batches(train)
    for _ in range(iterations):
        order = list(range(len(train_b)))
        if shuffle:
            rnd.shuffle(order)
        for k in order:
            indptr, indices, values, lab, _part = train_b[k]
            pa.step(indptr, indices, values, lab)
    out = []
    for indptr, indices, values, _lab, part in batches(test):
        pred = pa.step(indptr, indices, values, None, predict_only=True).cpu().tolist()
        for (vec, label), p in zip(part, pred):
            if kw["algorithm"] == "binary":
                p = (1 if p > 0 else -1) if binary_pm1 else (p > 0)
            out.append(Left((label, p, vec)))
    torch.cuda.synchronize()
    ctx.barrier()
    out = out + pa.dump()
    return (out, pa) if return_model else out


def _offline_job(training_src, test_src, worker_parallelism, ps_parallelism, method, init, add,
                 iterations, pull_limit, iteration_wait_time, seed, shuffle, feature_count,
                 engine: str = "host", **engine_kw):
    if engine not in ("host", "device", "auto"):
        raise ValueError("engine must be 'host', 'device' or 'auto'")
    if engine != "host":
        from ...parallel import device_engine
        ok, why = device_engine.device_ready()
        if ok:
            return _offline_job_device(training_src, test_src, ps_parallelism, method, iterations, pull_limit, seed,
                                       shuffle, feature_count, **engine_kw)
        device_engine.fallback(why)
    rnd = random.Random(seed)
    W = worker_parallelism

    def place(records, wrap):
        parts = [[] for _ in range(W)]
        for vec, label in records:
            parts[rnd.randrange(W)].append(wrap((_to_sparse(vec, feature_count), label)))
        return parts

    train_parts = place(list(as_stream(training_src, 1).collect()), Left)
    test_parts = place(list(as_stream(test_src, 1).collect()), Right)
    stream = DataStream([train_parts[w] + [_TRAIN_EOF] + test_parts[w] + [_TEST_EOF] for w in range(W)])
    worker = add_pull_limiter(PAOfflineWorker(method, iterations, shuffle, seed), pull_limit)
    return transform(stream, worker, ps_logic=SimplePSLogic(init, add), worker_parallelism=W,
                     ps_parallelism=ps_parallelism, iteration_wait_time=iteration_wait_time, **engine_kw)


class _BoolAdapter(PassiveAggressiveBinaryAlgorithm):
    """Labels of the offline job are +1 / -1 (``binary/PassiveAggressiveFilter.scala``)."""

    def __init__(self, inner):
        self.inner = inner

    def delta(self, x, model, label):
        return self.inner.delta(x, model, label > 0)

    def predict(self, x, model):
        return 1 if self.inner.predict(x, model) else -1


def pa_binary_classification_offline(training_src, test_src, worker_parallelism: int, ps_parallelism: int,
                                     iterations: int, paf_type: int = 0, paf_const: float = 1.0,
                                     pull_limit: int = 10000, iteration_wait_time: int = 10000, *,
                                     seed: int = 0, shuffle: bool = True, feature_count: int = 0,
                                     **engine_kw):
    """``paf_type`` 0 = PA, 1 = PA-I, 2 = PA-II; records are ``(vector, label in {+1,-1})``."""
    algo = {0: PassiveAggressiveBinaryAlgorithm.build_pa(),
            1: PassiveAggressiveBinaryAlgorithm.build_pa_i(paf_const),
            2: PassiveAggressiveBinaryAlgorithm.build_pa_ii(paf_const)}[paf_type]
    return _offline_job(training_src, test_src, worker_parallelism, ps_parallelism, _BoolAdapter(algo),
                        lambda _i: 0.0, lambda a, b: a + b, iterations, pull_limit, iteration_wait_time,
                        seed, shuffle, feature_count, **engine_kw)


def pa_multi_classification_offline(training_src, test_src, worker_parallelism: int, ps_parallelism: int,
                                    iterations: int, method: PassiveAggressiveMulticlassAlgorithm,
                                    pull_limit: int = 10000, iteration_wait_time: int = 10000, *,
                                    seed: int = 0, shuffle: bool = True, feature_count: int = 0, **engine_kw):
    L = method.label_count
    return _offline_job(training_src, test_src, worker_parallelism, ps_parallelism, method,
                        lambda _i: np.zeros(L), lambda a, b: np.asarray(a) + np.asarray(b), iterations,
                        pull_limit, iteration_wait_time, seed, shuffle, feature_count, **engine_kw)
