"""Passive-aggressive classification on the device engine (``csrc/pa.cu``).

Parity: ``PassiveAggressiveParameterServer.transformBinary / transformMulticlass``
(``M/passive/aggressive/PassiveAggressiveParameterServer.scala:126-237``): weights live in a
sharded HBM table (range or hash partitioned with the reference's formulas), a CSR mini-batch
of examples is one kernel launch: pull of every active feature, margin / tau, push of the
deltas as ``red.add`` on the owner shard.  Labelled examples train, unlabelled ones predict.
The dump at ``close`` lists only touched features (``RangePSLogicWithClose.scala:42-46``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, List, Optional, Sequence, Tuple

import numpy as np

from ...entities import Left, Right
from ...ops import _native as N
from ...ops import philox
from ...ops.table import ShardedTable, new_flow, read_flow
from ...utils.observability import nvtx_range
from .sparse import SparseVector

MULTI = {"binary": 0, "ova": 1, "pb": 2, "ml": 3}
VARIANT = {"pa": 0, "pa_i": 1, "pa_ii": 2}
REPLICATE_BELOW_BYTES = 16 << 20


def to_csr(vectors: Sequence[SparseVector]):
    """``(indptr int64, indices int32, values fp32)`` numpy arrays of a list of sparse vectors."""
    nnz = np.fromiter((v.active_size for v in vectors), dtype=np.int64, count=len(vectors))
    indptr = np.zeros(len(vectors) + 1, dtype=np.int64)
    np.cumsum(nnz, out=indptr[1:])
    indices = np.concatenate([v.indices for v in vectors]).astype(np.int32) if len(vectors) else np.zeros(0, np.int32)
    values = np.concatenate([v.values for v in vectors]).astype(np.float32) if len(vectors) else np.zeros(0, np.float32)
    return indptr, indices, values


class DevicePA:
    def __init__(self, feature_count: int, *, label_count: int = 1, algorithm: str = "binary",
                 variant: str = "pa", aggressiveness: float = 1.0, range_partitioning: bool = True,
                 pull_limit: int = 0, lazy: bool = True, cost: Optional[np.ndarray] = None,
                 n_shards: Optional[int] = None, ctx=None, replicated: Optional[bool] = None):
        """``replicated`` (several GPUs): keep the whole weight table on every GPU and push through the
        NVLS multicast alias instead of sharding it.  Default: replicate tables up to
        ``REPLICATE_BELOW_BYTES`` - RCV1-sized binary weights are 189 KB, and as 4-byte rows behind
        NVLink they made 8 GPUs slower than one (0.51 vs 0.88 G examples/s)."""
        import torch
        from ...parallel.device_context import get_context
        self.ctx = ctx or get_context()
        self.algorithm = algorithm
        self.feature_count, self.label_count = feature_count, (1 if algorithm == "binary" else label_count)
        if replicated is None:
            replicated = self.ctx.distributed and feature_count * self.label_count * 4 <= REPLICATE_BELOW_BYTES
        # multiclass rows of 256 bytes and more are padded to whole 128-byte lines: a 400-byte row (100 classes)
        # straddles 4-5 lines wherever it starts, a 512-byte one covers exactly 4 (measured on config 4: 0.675 ->
        # see profiles/r2_pa/README.md; FPS_PA_ALIGN_ROWS=0 keeps the packed layout)
        pad = 0
        if algorithm != "binary" and self.label_count >= 64 and os.environ.get("FPS_PA_ALIGN_ROWS", "1") != "0":
            pad = (-self.label_count) % 32
        self.table = ShardedTable(feature_count, self.label_count, n_shards=n_shards,
                                  partition="range" if range_partitioning else "hash",
                                  init_range=(0.0, 0.0), lazy=lazy, ctx=self.ctx,
                                  scalar=(algorithm == "binary"), replicated=replicated, aux_cols=pad)
        self.params = N.FpsPaParams()
        self.params.variant = VARIANT[variant]
        self.params.C = aggressiveness
        self.params.multiclass = MULTI[algorithm]
        self.params.n_classes = self.label_count
        self.params.pull_limit = int(pull_limit)
        self.params.local_shard = 0 if self.table.replicated else (self.ctx.rank if self.ctx.distributed else -1)
        self.flow = new_flow()
        self.cost = None if cost is None else torch.as_tensor(np.asarray(cost, dtype=np.float32)).cuda().contiguous()
        self.launches = 0

    def step(self, indptr, indices, values, labels=None, predict_only: bool = False):
        """One fused launch over a CSR batch on the device.  Returns the predictions made with
        the pulled (pre-update) weights: +1/-1 (binary) or the class index."""
        import torch
        n = indptr.numel() - 1
        pred = torch.empty(n, device="cuda", dtype=torch.int32)
        self.params.predict_only = int(predict_only or labels is None)
        with nvtx_range("fps.pa.step"):
            N.check(N.lib().fps_pa_step_cost(self.table.view_ref(), N.ptr(indptr), N.ptr(indices), N.ptr(values),
                                             N.ptr(labels), n, C.byref(self.params), N.ptr(self.cost),
                                             N.ptr(pred), N.ptr(self.flow), N.current_stream()), "fps_pa_step")
            if not self.params.predict_only:
                self.table.flush_deltas()    # replicated table: this launch's summed deltas -> the other replicas
        self.launches += 1
        return pred

    def step_vectors(self, vectors: Sequence[SparseVector], labels: Optional[Sequence] = None):
        import torch
        indptr, indices, values = to_csr(vectors)
        lab = None
        if labels is not None:
            if self.algorithm == "binary":
                lab = torch.tensor([1 if (l is True or l > 0) else -1 for l in labels], dtype=torch.int32).cuda()
            else:
                lab = torch.tensor(list(labels), dtype=torch.int32).cuda()
        return self.step(torch.from_numpy(indptr).cuda(), torch.from_numpy(indices).cuda(),
                         torch.from_numpy(values).cuda(), lab)

    def stats(self) -> dict:
        return read_flow(self.flow)

    def dump(self) -> list:
        """``Right((feature_id, param))`` for every touched feature of the local shard."""
        ids, rows = self.table.dump()
        rows = rows[:, : self.label_count].cpu().numpy().astype(np.float64)        # padding columns stay inside
        if self.algorithm == "binary":
            return [Right((k, float(v[0]))) for k, v in zip(ids.tolist(), rows)]
        return [Right((k, v)) for k, v in zip(ids.tolist(), rows)]

    def load(self, model: Iterable[Tuple[int, object]]) -> None:
        """Warm start (``transformWithModelLoad``) from ``(feature_id, param)`` pairs."""
        import torch
        model = list(model)
        if not model:
            return
        ids = torch.tensor([k for k, _ in model], dtype=torch.int64)
        vals = torch.from_numpy(np.stack([np.atleast_1d(np.asarray(v, dtype=np.float32)) for _, v in model]))
        self.table.load(ids, vals)


def transform_binary_device(input_source, passive_aggressive_variant: str = "pa", aggressiveness: float = 1.0,
                            pull_limit: int = 0, feature_count: int = 0, range_partitioning: bool = True,
                            batch_size: int = 4096, model=None, n_shards: Optional[int] = None):
    """Device twin of ``transform_binary``: consumes ``Left((vec, bool))`` / ``Right((id, vec))``
    records in mini-batches; returns ``Left((vec_id, prediction))`` for the unlabelled ones and the
    ``Right((feature, weight))`` dump."""
    pa = DevicePA(feature_count, algorithm="binary", variant=passive_aggressive_variant,
                  aggressiveness=aggressiveness, range_partitioning=range_partitioning,
                  pull_limit=pull_limit, n_shards=n_shards)
    if model is not None:
        pa.load(model)
    return _run_stream(pa, input_source, batch_size, long_id=False, to_label=lambda p: p > 0)


def transform_multiclass_device(input_source, algorithm: str = "ova", variant: str = "pa",
                                aggressiveness: float = 1.0, pull_limit: int = 0, label_count: int = 2,
                                feature_count: int = 0, range_partitioning: bool = True,
                                batch_size: int = 4096, model=None, long_id: bool = True, cost=None,
                                n_shards: Optional[int] = None):
    pa = DevicePA(feature_count, label_count=label_count, algorithm=algorithm, variant=variant,
                  aggressiveness=aggressiveness, range_partitioning=range_partitioning,
                  pull_limit=pull_limit, cost=cost, n_shards=n_shards)
    if model is not None:
        pa.load(model)
    return _run_stream(pa, input_source, batch_size, long_id=long_id, to_label=int)


def _run_stream(pa: DevicePA, input_source, batch_size: int, long_id: bool, to_label):
    import torch
    out: list = []
    buf: list = []

    def flush():
        if not buf:
            return
        labelled = buf[0].is_left
        vecs = [r.value[0] if r.is_left else r.value[1] for r in buf]
        if labelled:
            pa.step_vectors(vecs, [r.value[1] for r in buf])
        else:
            pred = pa.step_vectors(vecs, None).cpu().tolist()
            for r, p in zip(buf, pred):
                vid = r.value[0] if long_id else r.value[1]
                out.append(Left((vid, to_label(p))))
        buf.clear()

    for rec in input_source:
        if buf and (rec.is_left != buf[0].is_left or len(buf) >= batch_size):
            flush()
        buf.append(rec)
    flush()
    torch.cuda.synchronize()
    pa.ctx.barrier()
    return out + pa.dump(), pa


def method_to_device(method) -> dict:
    """Maps a host ``PassiveAggressiveAlgorithm`` object (``models/pa/algorithm.py``) to the kernel's
    parameters: ``algorithm`` (binary / ova / pb / ml), ``variant`` (pa / pa_i / pa_ii), ``aggressiveness``,
    ``label_count`` and the class-cost matrix of the cost-based rules."""
    from . import algorithm as A
    variants = {0: "pa", 1: "pa_i", 2: "pa_ii"}
    if isinstance(method, A.PassiveAggressiveCostBased):
        L = method.label_count
        cost = np.array([[float(method.cost(y, q)) for q in range(L)] for y in range(L)], dtype=np.float32)
        return {"algorithm": "pb" if method.mode.upper() == "PB" else "ml", "variant": "pa", "aggressiveness": 1.0,
                "label_count": L, "cost": cost}
    if isinstance(method, A.PassiveAggressiveOneVersusAll):
        return {"algorithm": "ova", "variant": variants[method.VARIANT], "aggressiveness": method.aggressiveness or 1.0,
                "label_count": method.label_count, "cost": None}
    if isinstance(method, A.PassiveAggressiveBinaryAlgorithm):
        return {"algorithm": "binary", "variant": variants[method.VARIANT], "aggressiveness": method.aggressiveness or 1.0,
                "label_count": 1, "cost": None}
    raise NotImplementedError(f"no device kernel for {type(method).__name__}")


def run_pa_method_device(input_source, method, *, feature_count: int, pull_limit: int = 0,
                         range_partitioning: bool = True, batch_size: int = 4096, model=None,
                         long_id: bool = False, n_shards: Optional[int] = None):
    """The online PA job of ``transformGeneric`` (``PassiveAggressiveParameterServer.scala:239-368``) on
    the device for any of the built-in update rules; ``model`` = ``(feature, param)`` pairs to warm-start
    from (``transformWithModelLoad``).  Returns ``(output stream, DevicePA)``."""
    kw = method_to_device(method)
    pa = DevicePA(feature_count, label_count=kw["label_count"], algorithm=kw["algorithm"], variant=kw["variant"],
                  aggressiveness=kw["aggressiveness"], range_partitioning=range_partitioning,
                  pull_limit=pull_limit, cost=kw["cost"], n_shards=n_shards)
    if model is not None:
        pa.load(model)
    binary = kw["algorithm"] == "binary"
    return _run_stream(pa, input_source, batch_size, long_id=long_id,
                       to_label=(lambda p: p > 0) if binary else int)
