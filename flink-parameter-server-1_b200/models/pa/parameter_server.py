"""Online passive-aggressive classification on the parameter server (host engine).

Parity: ``M/passive/aggressive/PassiveAggressiveParameterServer.scala`` -
``transformBinary`` (``:210-237``), ``transformMulticlass`` (``:170-182``),
``transformMulticlassWithLongId`` (``:126-138``), shared ``transformGeneric`` (``:239-368``),
``rangePartitionerPS`` (``:372-384``).

Input records: ``Left((vector, label))`` = labelled -> train; ``Right((long_id, vector))`` =
unlabelled -> predict.  Output: ``Left((vec_id, predicted_label))`` and, at shutdown,
``Right((feature_id, param))`` for every touched feature.

Worker protocol: pull every active feature of the vector, collect the answers per example,
then either push one delta per feature or output a prediction.  Always behind a pull limiter
(counted in features).
"""
from __future__ import annotations

from collections import deque
from typing import Callable, Deque, Dict, List, Optional

import numpy as np

from ...api import ParameterServerClient, WorkerLogic
from ...parameter_server import transform_generic, transform_with_model_load
from ...client import SimpleWorkerReceiver, SimpleWorkerSender
from ...server import (RangePSLogicWithClose, SimplePSLogicWithClose, SimplePSReceiver, SimplePSSender,
                       range_partition_div)
from ...worker_logic import add_pull_limiter
from .algorithm import (PassiveAggressiveAlgorithm, binary_initializer, multiclass_initializer)
from .sparse import SparseVector


def range_partitioner_ps(feature_count: int, ps_parallelism: int) -> Callable:
    size = range_partition_div(feature_count, ps_parallelism)
    return lambda msg: abs(msg.param_id) // size


def hash_partitioner_ps(ps_parallelism: int) -> Callable:
    return lambda msg: abs(msg.param_id) % ps_parallelism


class _Pending:
    __slots__ = ("record", "vector", "answers")

    def __init__(self, record, vector):
        self.record, self.vector, self.answers = record, vector, []


class PAWorkerLogic(WorkerLogic):
    """``PassiveAggressiveParameterServer.scala:283-338``."""

    def __init__(self, method: PassiveAggressiveAlgorithm, long_id: bool):
        self.method, self.long_id = method, long_id
        self.param_waiting_queue: Dict[int, Deque[_Pending]] = {}

    @staticmethod
    def _vector(rec) -> SparseVector:
        return rec.value[0] if rec.is_left else rec.value[1]

    def _vec_id(self, rec):
        if rec.is_right and self.long_id:
            return rec.value[0]
        return self._vector(rec)

    def on_recv(self, data, ps: ParameterServerClient) -> None:
        vec = self._vector(data)
        pend = _Pending(data, vec)
        if vec.active_size == 0:
            self._finish(pend, ps)
            return
        for k in vec.indices.tolist():
            self.param_waiting_queue.setdefault(k, deque()).append(pend)
            ps.pull(k)

    def on_pull_recv(self, param_id: int, model_value, ps: ParameterServerClient) -> None:
        q = self.param_waiting_queue[param_id]
        pend = q.popleft()
        pend.answers.append((param_id, model_value))
        if len(pend.answers) == pend.vector.active_size:
            self._finish(pend, ps)
        if not q:
            del self.param_waiting_queue[param_id]

    def _finish(self, pend: _Pending, ps: ParameterServerClient) -> None:
        model = dict(pend.answers)
        rec = pend.record
        if rec.is_left:
            for i, v in self.method.delta(pend.vector, model, rec.value[1]):
                ps.push(i, v)
        else:
            ps.output((self._vec_id(rec), self.method.predict(pend.vector, model)))


def _transform_generic(model, init, add, input_source, worker_parallelism, ps_parallelism, method,
                       pull_limit, feature_count, range_partitioning, iteration_wait_time, long_id,
                       engine: str = "host", **engine_kw):
    if engine not in ("host", "device", "auto"):
        raise ValueError("engine must be 'host', 'device' or 'auto'")
    if engine != "host":
        # the device twin: CSR mini-batches through the fused gather / margin / scatter kernel
        from ...parallel import device_engine
        ok, why = device_engine.device_ready()
        if ok:
            from ...parallel.device_context import get_context
            from .device import run_pa_method_device
            records = input_source.collect() if hasattr(input_source, "collect") else list(input_source)
            warm = None
            if model is not None:
                warm = [(e.value if hasattr(e, "value") else e) for e in
                        (model.collect() if hasattr(model, "collect") else list(model))]
            shards = None if get_context().distributed else max(1, min(int(ps_parallelism), 16))
            out, _pa = run_pa_method_device(records, method, feature_count=feature_count, pull_limit=pull_limit,
                                            range_partitioning=range_partitioning, model=warm, long_id=long_id,
                                            n_shards=shards, **engine_kw)
            return out
        device_engine.fallback(why)
    if range_partitioning:
        server_logic = RangePSLogicWithClose(feature_count, init, add)
        partitioner = range_partitioner_ps(feature_count, ps_parallelism)
    else:
        server_logic = SimplePSLogicWithClose(init, add)
        partitioner = hash_partitioner_ps(ps_parallelism)
    worker = add_pull_limiter(PAWorkerLogic(method, long_id), pull_limit)
    w_in = lambda m: m.worker_partition_index  # noqa: E731
    if model is not None:
        return transform_with_model_load(model)(
            input_source, worker, server_logic, partitioner, w_in, worker_parallelism, ps_parallelism,
            iteration_wait_time, **engine_kw)
    return transform_generic(input_source, worker, server_logic, partitioner, w_in, worker_parallelism,
                             ps_parallelism, SimpleWorkerReceiver(), SimpleWorkerSender(),
                             SimplePSReceiver(), SimplePSSender(), iteration_wait_time, **engine_kw)


def transform_binary(model=None):
    def run(input_source, worker_parallelism: int, ps_parallelism: int, passive_aggressive_method,
            pull_limit: int, feature_count: int, range_partitioning: bool, iteration_wait_time: int,
            **engine_kw):
        return _transform_generic(model, binary_initializer, lambda a, b: a + b, input_source,
                                  worker_parallelism, ps_parallelism, passive_aggressive_method, pull_limit,
                                  feature_count, range_partitioning, iteration_wait_time, False, **engine_kw)
    return run


def _multiclass(model, long_id: bool):
    def run(input_source, worker_parallelism: int, ps_parallelism: int, passive_aggressive_method,
            pull_limit: int, label_count: int, feature_count: int, range_partitioning: bool,
            iteration_wait_time: int, **engine_kw):
        return _transform_generic(model, multiclass_initializer(label_count),
                                  lambda a, b: np.asarray(a) + np.asarray(b), input_source,
                                  worker_parallelism, ps_parallelism, passive_aggressive_method, pull_limit,
                                  feature_count, range_partitioning, iteration_wait_time, long_id, **engine_kw)
    return run


def transform_multiclass(model=None):
    return _multiclass(model, False)


def transform_multiclass_with_long_id(model=None):
    return _multiclass(model, True)
