"""The public parameter-server entry points on the host engine.

Parity with ``M/FlinkParameterServer.scala``:
  * ``transform(training_data, worker_logic, param_init, param_update, W, S, wait)``   - ``:62-77``
  * ``transform(..., ps_logic=...)`` (hash partitioning, simple wire)               - ``:108-149``
  * ``transform_generic`` (12 arguments, every component pluggable)                  - ``:195-336``
  * ``transform_with_model_load(model)(...)``                                       - ``:377-566``
  * ``transform_with_double_model_load(model)(...)``                                - ``:607-816``

All return the materialised output stream: a list of ``Left(worker_out)`` / ``Right(ps_out)``
(the reference's ``DataStream[Either[WOut, PSOut]]``), which doubles as the model-dump format.
``return_engine=True`` also hands back the engine (stats, per-subtask logic instances).
"""
from __future__ import annotations

from typing import Any, Callable, Iterable, List, Optional

from .api import (ParameterServer, ParameterServerClient, ParameterServerLogic, RuntimeContext,
                  WorkerLogic)
from .client import SimpleWorkerReceiver, SimpleWorkerSender
from .entities import EOF, Left, PSToWorker, PullAnswer, Push, Right, WorkerToPS
from .parallel.engine import HostEngine
from .parallel.streams import DataStream, as_stream
from .server import SimplePSLogic, SimplePSReceiver, SimplePSSender


def hash_func(x) -> int:
    """``abs(x.hashCode)`` (``:126``); for ints ``hashCode`` is the value itself."""
    return abs(x if isinstance(x, int) else hash(x))


def default_param_partitioner(msg: WorkerToPS) -> int:
    """``hash(paramId)``; the engine applies the final ``% psParallelism`` (``:128-134``)."""
    return hash_func(msg.param_id)


def default_w_in_partition(msg: PSToWorker) -> int:
    return msg.worker_partition_index


def transform_generic(training_data, worker_logic: WorkerLogic, ps_logic: ParameterServerLogic,
                      param_partitioner, w_in_partition, worker_parallelism: int,
                      ps_parallelism: int, worker_receiver, worker_sender, ps_receiver, ps_sender,
                      iteration_wait_time: int, *, return_engine: bool = False, **engine_kw):
    eng = HostEngine(training_data, worker_logic, ps_logic, param_partitioner, w_in_partition,
                     worker_parallelism, ps_parallelism, worker_receiver, worker_sender,
                     ps_receiver, ps_sender, iteration_wait_time, **engine_kw)
    out = eng.run()
    return (out, eng) if return_engine else out


def transform(training_data, worker_logic: WorkerLogic, param_init=None, param_update=None,
              worker_parallelism: int = 1, ps_parallelism: int = 1, iteration_wait_time: int = 10000,
              *, ps_logic: Optional[ParameterServerLogic] = None, param_partitioner=None,
              w_in_partition=None, by_name: bool = False, return_engine: bool = False,
              engine: str = "host", device_kw: Optional[dict] = None, **engine_kw):
    """README signature.  Give either ``param_init`` + ``param_update`` (builds a
    ``SimplePSLogic``) or a ready ``ps_logic``.

    ``engine="device"`` runs the job on the GPU engine when the worker logic (under its pull limiter,
    if any) has a registered device functor and ``param_update`` a device update mode
    (``parallel/device_engine.py``); otherwise it says so in a ``DeviceFallbackWarning`` and runs
    here.  ``device_kw`` are passed to the functor (e.g. ``n_users``, ``feature_count``)."""
    if engine not in ("host", "device", "auto"):
        raise ValueError("engine must be 'host', 'device' or 'auto'")
    if engine != "host":
        from .parallel import device_engine
        handled, out = device_engine.try_transform(
            training_data, worker_logic, param_init, param_update, worker_parallelism, ps_parallelism,
            iteration_wait_time, ps_logic=ps_logic, **(device_kw or {}))
        if handled:
            return out
    if ps_logic is None:
        if param_init is None or param_update is None:
            raise ValueError("transform needs param_init and param_update, or ps_logic")
        ps_logic = SimplePSLogic(param_init, param_update, by_name=by_name)
    return transform_generic(
        training_data, worker_logic, ps_logic,
        param_partitioner or default_param_partitioner,
        w_in_partition or default_w_in_partition,
        worker_parallelism, ps_parallelism,
        SimpleWorkerReceiver(), SimpleWorkerSender(), SimplePSReceiver(), SimplePSSender(),
        iteration_wait_time, return_engine=return_engine, **engine_kw)


# --- warm start ----------------------------------------------------------------------------

class _Param:
    """``Parameter(id, param)`` of the double-load protocol / ``Right(param)`` of the single."""
    __slots__ = ("id", "param")

    def __init__(self, id_, param):
        self.id, self.param = id_, param


class _ModelWorkerData:
    __slots__ = ("id", "param")

    def __init__(self, id_, param):
        self.id, self.param = id_, param


class _Training:
    __slots__ = ("data",)

    def __init__(self, data):
        self.data = data


class _EOFMark:
    pass


_EOF = _EOFMark()


def _model_with_eof(model, worker_parallelism: int, wrap) -> DataStream:
    """Rebalance the model to the workers and append one EOF per worker subtask.  The reference
    throws when a model partition is empty (``:416``); here an empty partition just yields its
    EOF."""
    parts = as_stream(model, 1).rebalance(worker_parallelism).partitions()

    def with_eof(p):
        for rec in p:
            yield wrap(rec)
        yield _EOF

    return DataStream([with_eof(p) for p in parts])


class _UnwrapClient(ParameterServerClient):
    def __init__(self, ps: ParameterServerClient):
        self.ps = ps

    def pull(self, param_id: int) -> None:
        self.ps.pull(param_id)

    def push(self, param_id: int, delta_update) -> None:
        self.ps.push(param_id, _Param(param_id, delta_update))

    def output(self, out) -> None:
        self.ps.output(out)


class _LoadWorker(WorkerLogic):
    def __init__(self, inner: WorkerLogic, ps_parallelism: int):
        self.inner = inner
        self.S = ps_parallelism
        self.received_eof = False
        self.data_buffer: List[Any] = []

    @property
    def needs_realtime(self):
        return bool(getattr(self.inner, "needs_realtime", False))

    def open(self, runtime_context=None):
        self.inner.open(runtime_context)

    def close(self):
        self.inner.close()

    def on_recv(self, rec, ps):
        if rec is _EOF:
            self.received_eof = True
            for ps_idx in range(self.S):      # X6: notify every shard
                ps.push(ps_idx, _EOF)
            wrapped = _UnwrapClient(ps)
            buf, self.data_buffer = self.data_buffer, []
            for dp in buf:
                self.inner.on_recv(dp, wrapped)
        elif isinstance(rec, _Param):
            ps.push(rec.id, rec)
        elif isinstance(rec, _ModelWorkerData):
            self.inner.update_model(rec.id, rec.param)
        else:
            if self.received_eof:
                self.inner.on_recv(rec.data, _UnwrapClient(ps))
            else:
                self.data_buffer.append(rec.data)

    def on_pull_recv(self, param_id, value, ps):
        if isinstance(value, _Param):
            self.inner.on_pull_recv(param_id, value.param, _UnwrapClient(ps))
        elif not self.allow_eof_answers:
            raise RuntimeError("PS should not send EOF pull answers")

    allow_eof_answers = False


class _WrapPS(ParameterServer):
    def __init__(self, ps):
        self.ps = ps

    def answer_pull(self, param_id, value, worker_partition_index):
        self.ps.answer_pull(param_id, _Param(param_id, value), worker_partition_index)

    def output(self, out):
        self.ps.output(out)


class _LoadPS(ParameterServerLogic):
    def __init__(self, inner: ParameterServerLogic, worker_parallelism: int, keepalive: bool):
        self.inner = inner
        self.W = worker_parallelism
        self.eof_count_down = worker_parallelism
        self.pull_buffer: List[tuple] = []
        self.keepalive = keepalive

    def open(self, parameters=None, runtime_context=None):
        self.inner.open(parameters, runtime_context)

    def close(self, ps):
        self.inner.close(_WrapPS(ps))

    def on_pull_recv(self, param_id, worker_partition_index, ps):
        if self.eof_count_down == 0:
            self.inner.on_pull_recv(param_id, worker_partition_index, _WrapPS(ps))
        else:
            self.pull_buffer.append((param_id, worker_partition_index))

    def on_push_recv(self, param_id, delta, ps):
        if delta is _EOF:
            self.eof_count_down -= 1
            if self.eof_count_down == 0:
                buf, self.pull_buffer = self.pull_buffer, []
                for pid, widx in buf:
                    self.inner.on_pull_recv(pid, widx, _WrapPS(ps))
        else:
            if self.keepalive and self.eof_count_down > 0:
                # keeps the iteration timer alive during a long load (``:780-783``)
                ps.answer_pull(param_id, _EOF, param_id % self.W)
            self.inner.on_push_recv(param_id, delta.param, _WrapPS(ps))


def _wrapped_partitioners(param_partitioner, w_in_partition):
    def wpp(msg: WorkerToPS) -> int:
        inner = msg.msg
        if inner.is_left:
            return param_partitioner(msg)
        push = inner.value
        if push.delta is _EOF:
            return push.param_id            # EOFs are addressed by shard index itself (``:489``)
        return param_partitioner(WorkerToPS(msg.worker_partition_index,
                                            Right(Push(push.param_id, push.delta.param))))

    def wwp(msg: PSToWorker) -> int:
        if msg.msg.param is _EOF:
            return msg.worker_partition_index
        return w_in_partition(PSToWorker(msg.worker_partition_index,
                                         PullAnswer(msg.msg.param_id, msg.msg.param.param)))
    return wpp, wwp


def transform_with_model_load(model):
    """Warm start: ``transform_with_model_load(model)(training_data, worker_logic, ps_logic, ...)``.
    ``model`` is an iterable / stream of ``(id, P)``.  Contract: the PS logic must accept a push
    before any pull (``:342-344``); ``SimplePSLogic`` and friends store the pushed value."""

    def run(training_data, worker_logic: WorkerLogic, ps_logic: ParameterServerLogic,
            param_partitioner=None, w_in_partition=None, worker_parallelism: int = 1,
            ps_parallelism: int = 1, iteration_wait_time: int = 10000, *,
            return_engine: bool = False, **engine_kw):
        pp, wp = _wrapped_partitioners(param_partitioner or default_param_partitioner,
                                       w_in_partition or default_w_in_partition)
        model_s = _model_with_eof(model, worker_parallelism, lambda kv: _Param(kv[0], kv[1]))
        train_s = as_stream(training_data, worker_parallelism).map(_Training)
        return transform_generic(
            model_s.union(train_s), _LoadWorker(worker_logic, ps_parallelism),
            _LoadPS(ps_logic, worker_parallelism, keepalive=False), pp, wp,
            worker_parallelism, ps_parallelism,
            SimpleWorkerReceiver(), SimpleWorkerSender(), SimplePSReceiver(), SimplePSSender(),
            iteration_wait_time, return_engine=return_engine, **engine_kw)

    return run


def transform_with_double_model_load(model, worker_owner: Optional[Callable[[int, int], int]] = None):
    """Warm start of both model halves.  ``model`` yields ``Left((id, P))`` -> pushed to the PS,
    ``Right((id, P))`` -> ``worker_logic.update_model(id, P)``.

    The reference places worker-side rows by round-robin ``rebalance``.  Pass
    ``worker_owner(id, W) -> worker`` to place them by key instead (what the top-K and
    learner+generator jobs actually need: the worker that owns an item must hold it)."""

    def run(training_data, worker_logic, ps_logic: ParameterServerLogic,
            param_partitioner=None, w_in_partition=None, worker_parallelism: int = 1,
            ps_parallelism: int = 1, iteration_wait_time: int = 10000, *,
            return_engine: bool = False, **engine_kw):
        if not hasattr(worker_logic, "update_model"):
            raise TypeError("double model load needs a worker logic with update_model(id, param) "
                            "(BaseMFWorkerLogic)")
        pp, wp = _wrapped_partitioners(param_partitioner or default_param_partitioner,
                                       w_in_partition or default_w_in_partition)

        def wrap(e):
            k, v = e.value
            return _Param(k, v) if e.is_left else _ModelWorkerData(k, v)

        W = worker_parallelism
        if worker_owner is None:
            model_s = _model_with_eof(model, W, wrap)
        else:
            def route(e):
                return worker_owner(e.value[0], W) if e.is_right else hash_func(e.value[0])
            parts = as_stream(model, 1).partition_custom(route, W).partitions()

            def with_eof(p):
                for rec in p:
                    yield wrap(rec)
                yield _EOF
            model_s = DataStream([with_eof(p) for p in parts])
        train_s = as_stream(training_data, W).map(_Training)
        lw = _LoadWorker(worker_logic, ps_parallelism)
        lw.allow_eof_answers = True
        return transform_generic(
            model_s.union(train_s), lw, _LoadPS(ps_logic, W, keepalive=True), pp, wp,
            W, ps_parallelism,
            SimpleWorkerReceiver(), SimpleWorkerSender(), SimplePSReceiver(), SimplePSSender(),
            iteration_wait_time, return_engine=return_engine, **engine_kw)

    return run
