"""The device engine behind the public API.

``transform(training_data, worker_logic, param_init, param_update, W, S, wait, engine="device")``
(``M/FlinkParameterServer.scala:62-77``, ``:108-149``) and the named jobs
(``ps_offline_mf``, ``ps_online_mf``, ``ps_top_k_generator``, ``ps_online_learner_and_generator``,
``transform_binary``, ``transform_multiclass``) accept ``engine="host" | "device" | "auto"``.

On the device a worker logic is not Python code but a *device functor* - a fused kernel that does
pull, compute and push for a whole batch.  This module is the registry that maps the built-in
``WorkerLogic`` classes to their functors and the known ``param_update`` callables to the update
modes the tables implement, so a new functor can be plugged in without touching the front-ends:

    @register_device_logic(MyWorkerLogic)
    def run_my_logic(job: DeviceJob) -> list: ...

    register_param_update(my_update_fn, "add")        # "add" | "overwrite" | "add_norm"

A logic or an update rule with no device twin falls back to the host engine **with a warning**
(``DeviceFallbackWarning``), never silently; ``engine="device"`` with ``strict=True`` raises instead.
"""
from __future__ import annotations

import warnings
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple


class DeviceFallbackWarning(UserWarning):
    """The requested job ran on the host engine because it has no device twin."""


@dataclass
class DeviceJob:
    """Everything a device runner gets: the unwrapped logic, the limiter's limit, the materialised
    input records and the arguments of ``transform``."""
    logic: Any
    pull_limit: int
    records: List[Any]
    param_init: Optional[Callable] = None
    param_update: Optional[Callable] = None
    update_mode: Optional[str] = None
    ps_logic: Any = None
    worker_parallelism: int = 1
    ps_parallelism: int = 1
    iteration_wait_time: int = 10000
    extra: Dict[str, Any] = field(default_factory=dict)


_RUNNERS: Dict[type, Callable[[DeviceJob], list]] = {}
_PARAM_UPDATES: List[Tuple[Callable, str]] = []
UPDATE_MODES = ("add", "overwrite", "add_norm")


def register_device_logic(logic_cls: type):
    def deco(fn: Callable[[DeviceJob], list]):
        _RUNNERS[logic_cls] = fn
        return fn
    return deco


def register_param_update(fn: Callable, mode: str) -> None:
    if mode not in UPDATE_MODES:
        raise ValueError(f"update mode must be one of {UPDATE_MODES}")
    _PARAM_UPDATES.append((fn, mode))


def param_update_mode(fn: Optional[Callable]) -> Optional[str]:
    """The table update mode a ``param_update`` callable stands for, or ``None``."""
    if fn is None:
        return None
    if isinstance(fn, str):
        return fn if fn in UPDATE_MODES else None
    tagged = getattr(fn, "device_update_mode", None)
    if tagged in UPDATE_MODES:
        return tagged
    for known, mode in _PARAM_UPDATES:
        if fn is known:
            return mode
    return None


def unwrap_limiter(worker_logic) -> Tuple[Any, int]:
    """``add_pull_limiter(logic, n)`` -> ``(logic, n)``; a bare logic -> ``(logic, 0)``."""
    limit = 0
    seen = 0
    while hasattr(worker_logic, "inner") and seen < 8:
        limit = int(getattr(worker_logic, "pull_limit", limit) or limit)
        worker_logic = worker_logic.inner
        seen += 1
    return worker_logic, limit


def device_ready() -> Tuple[bool, str]:
    try:
        import torch
    except Exception as e:      # pragma: no cover
        return False, f"torch is not importable ({e})"
    if not torch.cuda.is_available():
        return False, "no CUDA device"
    from ..ops import _native
    if not _native.available():
        return False, f"{_native.LIB_PATH} has not been built"
    return True, ""


def runner_for(logic) -> Optional[Callable[[DeviceJob], list]]:
    _ensure_builtin_runners()
    for cls in type(logic).__mro__:
        if cls in _RUNNERS:
            return _RUNNERS[cls]
    return None


def _materialise(training_data) -> list:
    if hasattr(training_data, "collect"):
        return training_data.collect()
    return list(training_data)


def fallback(reason: str, strict: bool = False) -> None:
    if strict:
        raise NotImplementedError(f"engine='device': {reason}")
    warnings.warn(f"engine='device': {reason}; running on the host engine", DeviceFallbackWarning, stacklevel=3)


def try_transform(training_data, worker_logic, param_init, param_update, worker_parallelism: int,
                  ps_parallelism: int, iteration_wait_time: int, ps_logic=None, strict: bool = False,
                  **extra) -> Tuple[bool, Optional[list]]:
    """Run the job on the device engine if its logic has a registered functor.  Returns
    ``(handled, output)``; ``handled`` False = the caller should run the host engine (a
    ``DeviceFallbackWarning`` has been issued)."""
    ok, why = device_ready()
    if not ok:
        fallback(why, strict)
        return False, None
    logic, limit = unwrap_limiter(worker_logic)
    run = runner_for(logic)
    if run is None:
        fallback(f"no device functor is registered for {type(logic).__name__} "
                 f"(register one with parallel.device_engine.register_device_logic)", strict)
        return False, None
    if ps_logic is not None and param_update is None:
        param_update = getattr(ps_logic, "param_update", None) or getattr(ps_logic, "update", None)
        param_init = param_init or getattr(ps_logic, "param_init", None) or getattr(ps_logic, "init", None)
    mode = param_update_mode(param_update)
    if param_update is not None and mode is None:
        fallback(f"param_update {getattr(param_update, '__name__', param_update)!r} has no device update mode "
                 f"(known: {UPDATE_MODES}; tag it with register_param_update)", strict)
        return False, None
    job = DeviceJob(logic, limit, _materialise(training_data), param_init, param_update, mode, ps_logic,
                    worker_parallelism, ps_parallelism, iteration_wait_time, dict(extra))
    return True, run(job)


# ---------------------------------------------------------------------------------------------
# built-in functors
# ---------------------------------------------------------------------------------------------
_builtin_done = False


def _ensure_builtin_runners() -> None:
    """Registers the functors of the reference's own worker logics (imported lazily: the model
    packages import ``parameter_server``, which imports this module)."""
    global _builtin_done
    if _builtin_done:
        return
    _builtin_done = True
    from ..models.mf import vector as mf_vector
    from ..models.mf.workers import (PSOfflineMatrixFactorizationWorker, PSOnlineMatrixFactorizationWorker)
    from ..models.pa.parameter_server import PAWorkerLogic

    register_param_update(mf_vector.vector_sum, "add")

    def _ratings(records):
        out = []
        for r in records:
            if hasattr(r, "is_right"):          # Right(Rating) / Left(EOF) wrappers of flat_map_with_eof
                if r.is_right:
                    out.append(r.value)
            else:
                out.append(r)
        return out

    def _n_shards(job: DeviceJob):
        from .device_context import get_context
        return None if get_context().distributed else max(1, min(job.ps_parallelism, 16))

    @register_device_logic(PSOfflineMatrixFactorizationWorker)
    def _offline_mf(job: DeviceJob) -> list:
        from ..models.mf.device import ps_offline_mf_device
        w = job.logic
        d = w.factor_init_desc
        if job.update_mode not in (None, "add"):
            raise NotImplementedError("the MF functor applies param_update = add")
        return ps_offline_mf_device(_ratings(job.records), d.num_factors, d.range_min, d.range_max,
                                    w.factor_update.learning_rate, w.negative_sample_rate,
                                    getattr(w, "user_memory", 128), w.iterations, job.pull_limit,
                                    seed=w.seed, shuffle=w.shuffle, lambda_=w.factor_update.lambda_,
                                    n_shards=_n_shards(job), **job.extra)

    @register_device_logic(PSOnlineMatrixFactorizationWorker)
    def _online_mf(job: DeviceJob) -> list:
        from ..models.mf.device import ps_online_mf_device
        w = job.logic
        d = w.factor_init_desc
        if job.update_mode not in (None, "add"):
            raise NotImplementedError("the MF functor applies param_update = add")
        return ps_online_mf_device(_ratings(job.records), d.num_factors, d.range_min, d.range_max,
                                   w.factor_update.learning_rate, w.negative_sample_rate, w.user_memory,
                                   job.pull_limit, seed=w.seed, **job.extra)

    @register_device_logic(PAWorkerLogic)
    def _pa(job: DeviceJob) -> list:
        from ..models.pa.device import run_pa_method_device
        feature_count = job.extra.pop("feature_count", None) or getattr(job.ps_logic, "feature_count", None)
        if not feature_count:
            raise ValueError("the PA functor needs feature_count (pass it to transform(..., feature_count=n))")
        out, _pa_model = run_pa_method_device(job.records, job.logic.method, feature_count=feature_count,
                                              pull_limit=job.pull_limit, long_id=job.logic.long_id,
                                              n_shards=_n_shards(job), **job.extra)
        return out


# ---------------------------------------------------------------------------------------------
# generic device runtime for user logics written as tensor programs
# ---------------------------------------------------------------------------------------------
class TensorWorkerLogic:
    """A user worker logic for the device engine: the per-record callbacks of ``WorkerLogic``
    (``M/WorkerLogic.scala:23-58``) restated per *batch*, with tensors instead of messages.

    For every input batch the engine (1) asks ``pull_ids(batch)`` which parameters the batch needs,
    (2) pulls those rows from their owner shards with one gather kernel (P2P loads over NVLink for
    remote shards; lazy ``param_init`` applies), (3) calls ``on_pull_batch(batch, ids, rows)``, any
    torch program, which returns ``(push_ids, deltas, outputs)``, (4) applies ``param_update`` to
    ``push_ids`` on the owner shards with one scatter kernel (``add`` = remote ``red.add``,
    ``overwrite``, ``add_norm``) and appends ``outputs`` (a tensor or ``None``) to the worker side of
    the output stream.  Semantics are the asynchronous ones: no barrier between ranks, every rank
    runs its own batches; ``pull_limit`` caps the rows pulled per gather (a larger batch is pulled,
    processed and pushed in slices)."""

    def open(self, table, ctx) -> None:        # table: ops.table.ShardedTable
        pass

    def pull_ids(self, batch):
        raise NotImplementedError

    def on_pull_batch(self, batch, ids, rows):
        raise NotImplementedError

    def close(self) -> None:
        pass


def conflict_free_rounds(ids):
    """The device form of the per-key lock of ``LockPSLogicA`` (``M/server/LockPSLogicA.scala:13-46``): a
    pull of a key that is locked waits until the push of the pull before it has been applied.  For a batch of
    pulls that is a schedule in rounds - record ``j`` runs in round ``r`` if ``r`` earlier records of the
    batch asked for the same key - and inside a round no key occurs twice, so gather -> program -> scatter of
    a round is a set of independent read-modify-write cycles.  Returns the list of index tensors (positions in
    arrival order) of the rounds, computed on the device (stable sort + run starts)."""
    import torch
    n = int(ids.numel())
    if n == 0:
        return []
    order = torch.argsort(ids, stable=True)
    s = ids[order]
    idx = torch.arange(n, device=ids.device)
    start = torch.ones(n, dtype=torch.bool, device=ids.device)
    start[1:] = s[1:] != s[:-1]
    first = torch.cummax(torch.where(start, idx, torch.zeros_like(idx)), 0).values
    rank = torch.empty_like(idx)
    rank[order] = idx - first
    n_rounds = int(rank.max().item()) + 1
    if n_rounds == 1:
        return [idx]
    return [torch.nonzero(rank == r).flatten() for r in range(n_rounds)]


def _index_batch(batch, index, n: int):
    """``_slice_batch`` with an index tensor."""
    import torch
    if isinstance(batch, torch.Tensor):
        return batch[index.to(batch.device)] if batch.dim() and batch.shape[0] == n else batch
    if isinstance(batch, (tuple, list)):
        return type(batch)(_index_batch(b, index, n) for b in batch)
    if isinstance(batch, dict):
        return {k: _index_batch(v, index, n) for k, v in batch.items()}
    return batch


def run_tensor_logic(batches, logic: TensorWorkerLogic, n_ids: int, dim: int, *, init_range=(0.0, 0.0),
                     update_mode: str = "add", pull_limit: int = 0, lazy: bool = True, partition: str = "hash",
                     seed: int = 0, n_shards=None, aux_cols: int = 0, return_table: bool = False,
                     locking: bool = False):
    """The device twin of ``transform`` for a ``TensorWorkerLogic``.  Returns the output stream:
    ``Left(output tensor)`` per batch that produced one, then ``Right((id, row))`` for every live row of
    the local shard (on-close form).

    ``locking=True`` is the device twin of ``LockPSLogicA`` / ``LockPSLogicB``: every batch is run in
    conflict-free rounds (``conflict_free_rounds``), so the read-modify-write cycles on one key happen one
    after the other in arrival order and the result of a rank's stream is the sequential one - whatever
    ``param_update`` is.  The lock is the rank's: cycles of different ranks on one key still interleave
    (asynchronous semantics), as they do between the reference's workers whenever the lock is free."""
    import torch
    from ..entities import Left, Right
    from ..ops import philox
    from ..ops.table import ShardedTable, new_flow
    from .device_context import get_context
    if update_mode not in UPDATE_MODES:
        raise ValueError(f"update_mode must be one of {UPDATE_MODES}")
    ctx = get_context()
    table = ShardedTable(n_ids, dim, n_shards=n_shards, partition=partition, init_range=init_range, seed=seed,
                         stream_tag=philox.STREAM_ITEM, lazy=lazy, aux_cols=aux_cols, ctx=ctx)
    flow = new_flow()
    logic.open(table, ctx)
    out = []
    def run_part(batch, ids):
        step = int(pull_limit) if pull_limit and pull_limit > 0 else max(int(ids.numel()), 1)
        for s in range(0, max(int(ids.numel()), 1), step):
            part = ids[s:s + step]
            sub = batch if step >= ids.numel() else _slice_batch(batch, s, s + step, ids.numel())
            rows = table.gather(part, flow)
            push_ids, deltas, produced = logic.on_pull_batch(sub, part, rows)
            if push_ids is not None and push_ids.numel():
                table.scatter(push_ids, deltas, update_mode, flow)
            if produced is not None:
                out.append(Left(produced))

    for batch in batches:
        ids = logic.pull_ids(batch)
        ids = ids.to(device="cuda", dtype=torch.int64)
        if not locking:
            run_part(batch, ids)
            continue
        rounds = conflict_free_rounds(ids)
        for index in rounds:
            if len(rounds) == 1:
                run_part(batch, ids)
            else:
                run_part(_index_batch(batch, index, int(ids.numel())), ids[index])
    logic.close()
    torch.cuda.synchronize()
    ctx.barrier()
    ids, rows = table.dump()
    rows_np = rows.cpu().numpy()
    out.extend(Right((int(k), v)) for k, v in zip(ids.tolist(), rows_np))
    return (out, table, flow) if return_table else out


def _slice_batch(batch, lo: int, hi: int, n: int):
    """Slices every tensor of a batch (tensor, tuple / list / dict of tensors) whose first dimension is
    the batch length; anything else is passed through."""
    import torch
    if isinstance(batch, torch.Tensor):
        return batch[lo:hi] if batch.dim() and batch.shape[0] == n else batch
    if isinstance(batch, (tuple, list)):
        return type(batch)(_slice_batch(b, lo, hi, n) for b in batch)
    if isinstance(batch, dict):
        return {k: _slice_batch(v, lo, hi, n) for k, v in batch.items()}
    return batch


@register_device_logic(TensorWorkerLogic)
def _run_tensor_job(job: DeviceJob) -> list:
    kw = dict(job.extra)
    n_ids, dim = kw.pop("n_ids", None), kw.pop("dim", None)
    if n_ids is None or dim is None:
        raise ValueError("a TensorWorkerLogic job needs device_kw={'n_ids': ..., 'dim': ...}")
    if "locking" not in kw:           # a lock logic on the server side asks for the lock schedule
        from ..server.logics import LockPSLogicA
        kw["locking"] = isinstance(job.ps_logic, LockPSLogicA)
    return run_tensor_logic(job.records, job.logic, n_ids, dim, update_mode=job.update_mode or "add",
                            pull_limit=job.pull_limit, **kw)
