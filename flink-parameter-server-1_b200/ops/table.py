"""HBM-resident sharded parameter table (the device replacement of the PS shards' hash maps).

Parity: the store inside ``M/server/SimplePSLogic.scala:7-26`` / ``RangePSLogicWithClose.scala:7-63``
and the two partitioners (hash ``M/FlinkParameterServer.scala:126-134``; range
``M/passive/aggressive/PassiveAggressiveParameterServer.scala:372-384``), laid out for HBM:

  * shard ``s`` is one contiguous ``[rows_per_shard, stride]`` fp32 array (stride = dim rounded up
    to 4 floats, optionally + aux columns such as the vector length);
  * ids map to (owner, slot) with the reference's formulas, evaluated on the device;
  * ``lazy=True`` keeps the reference's life-cycle (untouched / first-pulled / first-pushed) in
    one state byte per row; ``lazy=False`` materialises ``init(id)`` eagerly (the benchmark mode);
  * with ``world_size > 1`` shard ``r`` lives on rank ``r`` and every rank maps all shards
    through VMM symmetric memory (or CUDA IPC with ``peer_mapping="ipc"``), so kernels dereference
    peer rows directly over NVLink.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import List, Optional, Tuple

from . import _native as N
from . import philox


def _symm_alloc(shape, dtype, ctx, keep: list):
    """Allocate a symmetric (same shape on every rank) VMM buffer and return
    ``(local tensor, [base pointer on rank 0, 1, ...])``."""
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(shape, dtype=dtype, device=ctx.device)
    hdl = symm.rendezvous(t, ctx.group if ctx.group is not None else dist.group.WORLD)
    keep.append((t, hdl))
    return t, [int(p) for p in hdl.buffer_ptrs]


def padded_stride(dim: int, aux_cols: int = 0) -> int:
    return ((dim + aux_cols + 3) // 4) * 4


class ShardedTable:
    def __init__(self, n_ids: int, dim: int, *, n_shards: Optional[int] = None, partition: str = "hash",
                 init_range: Tuple[float, float] = (0.0, 0.0), seed: int = 0,
                 stream_tag: int = philox.STREAM_ITEM, lazy: bool = False, aux_cols: int = 0,
                 ctx=None, local_only: bool = False, scalar: bool = False,
                 peer_mapping: Optional[str] = None, replicated: bool = False):
        """``local_only`` = a worker-resident table (MF user factors): sharded by the same formula
        but a rank only ever touches its own shard, so no peer mapping is made.
        ``scalar`` = one float per id with no padding (binary PA weights).
        ``replicated`` (several ranks only) = every GPU holds the whole table and pulls are local.
        Pushes are *deferred* by default: a push adds to the local replica and to a local delta array,
        ``flush_deltas()`` (called by the step functions after each launch) adds the accumulated
        deltas to the other replicas - a hot weight crosses NVLink once per launch instead of once
        per push.  ``FPS_REPLICA_PUSH=immediate`` applies every push to all replicas at once with one
        ``multimem.red`` on the NVLS multicast alias (or one ``red.add`` per replica without
        multicast); measured 15x slower for 4-byte rows (single-word packets cap near 5 G/s per GPU).
        For small hot tables, where sharding would send 4-byte rows across NVLink one by one."""
        import torch
        from ..parallel.device_context import get_context
        self.ctx = ctx or get_context()
        world = self.ctx.world_size
        self.n_ids, self.dim = int(n_ids), int(dim)
        if scalar and (dim != 1 or aux_cols):
            raise ValueError("a scalar table has dim 1 and no aux columns")
        self.stride = 1 if scalar else padded_stride(dim, aux_cols)
        self.aux_cols = aux_cols
        self.replicated = bool(replicated) and world > 1 and not local_only
        if self.replicated and lazy and tuple(init_range) != (0.0, 0.0):
            raise ValueError("a replicated lazy table needs a zero initialiser (state bytes are per replica)")
        self.n_shards = 1 if self.replicated else int(n_shards or world)
        if world > 1 and not self.replicated and self.n_shards != world:
            raise ValueError("with several ranks the table has exactly one shard per rank")
        if not 1 <= self.n_shards <= N.FPS_MAX_SHARDS:
            raise ValueError(f"n_shards must be in [1, {N.FPS_MAX_SHARDS}]")
        self.partition = {"hash": N.PART_HASH, "range": N.PART_RANGE}[partition]
        self.div = int(math.ceil(self.n_ids / self.n_shards))
        self.rows_per_shard = max(self.div, 1)
        self.init_range = (float(init_range[0]), float(init_range[1]))
        self.seed, self.stream_tag, self.lazy = int(seed), int(stream_tag), bool(lazy)
        self.local_only = local_only
        self.local_shards = [0] if self.replicated else ([self.ctx.rank] if world > 1 else list(range(self.n_shards)))

        self._bufs: dict = {}
        self._state_bufs: dict = {}
        self._peer_ptrs: List[int] = []
        self._symm = []               # keeps symmetric-memory handles alive
        self.rows: dict = {}
        self.state: dict = {}
        self._view = N.FpsTableView()
        v = self._view
        v.n_shards, v.dim, v.stride, v.partition = self.n_shards, self.dim, self.stride, self.partition
        v.div, v.rows_per_shard = self.div, self.rows_per_shard
        v.n_ids = self.n_ids
        v.init_lo, v.init_hi = self.init_range
        v.seed, v.stream_tag, v.lazy = self.seed, self.stream_tag, int(self.lazy)
        n_state = ((self.rows_per_shard + 3) // 4) * 4
        shared = world > 1 and not local_only and not self.replicated
        self.multicast = False
        self.deferred = False
        if self.replicated:
            self.rows[0], ptrs = _symm_alloc((self.rows_per_shard, self.stride), torch.float32, self.ctx, self._symm)
            v.rows[0] = ptrs[self.ctx.rank]
            for r in range(world):
                v.replica_rows[r] = ptrs[r]
            v.n_replicas = world
            v.replica_rank = self.ctx.rank
            self.deferred = os.environ.get("FPS_REPLICA_PUSH", "deferred") != "immediate"
            if self.deferred:
                # pushes add to the local replica and to this array; flush_deltas() forwards the sums
                self._delta = torch.zeros((self.rows_per_shard, self.stride), device="cuda", dtype=torch.float32)
                v.delta_rows = self._delta.data_ptr()
            mc = 0
            if os.environ.get("FPS_NO_MULTICAST", "0") != "1":
                try:
                    mc = int(self._symm[-1][1].multicast_ptr or 0)
                except Exception:
                    mc = 0
            v.mc_rows = mc or None
            self.multicast = bool(mc)
            if self.lazy:
                sb = N.DeviceBuffer(n_state)
                self._state_bufs[0] = sb
                self.state[0] = sb.tensor((n_state,), "|u1")
                v.state[0] = sb.ptr
        self.peer_mapping = (peer_mapping or os.environ.get("FPS_PEER_MAPPING", "vmm")) if shared else "none"
        if self.replicated:
            pass
        elif shared and self.peer_mapping == "vmm":
            # cuMem (VMM) allocations exchanged through torch symmetric memory.  Measured on 2xB200:
            # random 512-byte rows over a 2.4 GB peer shard run at the NVLink rate (2.2 ms / 2M rows)
            # through this mapping but 8x slower (17 ms) through cudaIpcOpenMemHandle mappings, which
            # collapse once the randomly accessed span exceeds ~1 GB (profiles/r1_2gpu/ipc_vs_vmm.md).
            me = self.ctx.rank
            self.rows[me], ptrs = _symm_alloc((self.rows_per_shard, self.stride), torch.float32, self.ctx, self._symm)
            for r in range(world):
                v.rows[r] = ptrs[r]
            if self.lazy:
                self.state[me], sptrs = _symm_alloc((n_state,), torch.uint8, self.ctx, self._symm)
                for r in range(world):
                    v.state[r] = sptrs[r]
        else:
            for s in self.local_shards:
                b = N.DeviceBuffer(self.rows_per_shard * self.stride * 4)
                self._bufs[s] = b
                self.rows[s] = b.tensor((self.rows_per_shard, self.stride), "<f4")
                v.rows[s] = b.ptr
                if self.lazy:
                    sb = N.DeviceBuffer(n_state)
                    self._state_bufs[s] = sb
                    self.state[s] = sb.tensor((n_state,), "|u1")
                    v.state[s] = sb.ptr
            if shared:
                self._map_peers_ipc()
        if self.lazy:
            for s in self.local_shards:
                self.rows[s].zero_()
                self.state[s].zero_()
        else:
            self.init_eager()
        torch.cuda.synchronize()
        self.ctx.barrier()

    # -- peer mapping ------------------------------------------------------------------------
    def _map_peers_ipc(self) -> None:
        """Legacy path (``peer_mapping="ipc"``): cudaMalloc + cudaIpc handles.  Correct, but slow
        for random access over large shards - kept for comparison and for small tables."""
        me = self.ctx.rank
        mine = (self._bufs[me].ipc_handle(), self._state_bufs[me].ipc_handle() if self.lazy else None)
        handles = self.ctx.all_gather_object(mine)
        for r, (hrows, hstate) in enumerate(handles):
            if r == me:
                continue
            p = N.open_ipc(hrows)
            self._peer_ptrs.append(p)
            self._view.rows[r] = p
            if hstate is not None:
                ps = N.open_ipc(hstate)
                self._peer_ptrs.append(ps)
                self._view.state[r] = ps

    # -- basic ops -----------------------------------------------------------------------------
    def view(self) -> N.FpsTableView:
        return self._view

    def view_ref(self):
        return C.byref(self._view)

    def init_eager(self) -> None:
        for s in self.local_shards:
            N.check(N.lib().fps_table_init_shard(C.byref(self._view), s, self.rows_per_shard,
                                                 N.current_stream()), "fps_table_init_shard")

    def owner_of(self, ids):
        """Vectorised torch twin of the device ``locate`` (tests / host routing)."""
        if self.partition == N.PART_HASH:
            return ids % self.n_shards, ids // self.n_shards
        owner = ids // self.div
        return owner, ids - owner * self.div

    def gather(self, ids, flow=None, out=None):
        """Pull without compute: ``[len(ids), dim]`` values (remote rows come over NVLink).  ``out``: an
        optional ``[len(ids), stride]`` fp32 device tensor to write into (e.g. a slice of an append log)."""
        import torch
        ids = ids.to(device="cuda", dtype=torch.int64).contiguous()
        if out is None:
            out = torch.empty((ids.numel(), self.stride), device="cuda", dtype=torch.float32)
        assert out.shape == (ids.numel(), self.stride) and out.is_contiguous()
        N.check(N.lib().fps_table_gather(C.byref(self._view), N.ptr(ids), ids.numel(), N.ptr(out),
                                         self.stride, N.ptr(flow), N.current_stream()), "fps_table_gather")
        return out[:, : self.dim + self.aux_cols]

    def scatter(self, ids, vals, mode: str = "add", flow=None) -> None:
        """Push: ``add`` (vectorSum), ``overwrite`` (model load / top-K users) or ``add_norm``
        (add, then refresh the length column)."""
        import torch
        ids = ids.to(device="cuda", dtype=torch.int64).contiguous()
        vals = vals.to(device="cuda", dtype=torch.float32).contiguous()
        if vals.dim() == 1:
            vals = vals[:, None]
        m = {"add": 0, "overwrite": 1, "add_norm": 2}[mode]
        if self.replicated and m == 2:
            raise ValueError("add_norm is not available on a replicated table")
        N.check(N.lib().fps_table_scatter(C.byref(self._view), N.ptr(ids), ids.numel(), N.ptr(vals),
                                          vals.shape[1], m, N.ptr(flow), N.current_stream()),
                "fps_table_scatter")
        if m == 0:
            self.flush_deltas()

    def locked_axpby(self, ids, alpha: float, deltas, *, lock_col: Optional[int] = None, use_lock: bool = True,
                     flow=None) -> None:
        """``row <- alpha * row + delta`` as a locked read-modify-write per record (the device form of
        ``LockPSLogicA``: a cycle on a key excludes every other cycle on it, on this GPU and on the peers).
        The lock word lives in an auxiliary column of the row (``aux_cols >= 1``; default: the last one)."""
        import torch
        if self.aux_cols < 1 or self.stride == 1:
            raise ValueError("locked_axpby needs a vector table with aux_cols >= 1 (the lock word's column)")
        col = self.dim + self.aux_cols - 1 if lock_col is None else int(lock_col)
        ids = ids.to(device="cuda", dtype=torch.int64).contiguous()
        deltas = deltas.to(device="cuda", dtype=torch.float32).contiguous()
        if deltas.dim() == 1:
            deltas = deltas[:, None]
        N.check(N.lib().fps_table_locked_axpby(C.byref(self._view), N.ptr(ids), ids.numel(), float(alpha), N.ptr(deltas),
                                               deltas.shape[1], col, int(bool(use_lock)), N.ptr(flow),
                                               N.current_stream()), "fps_table_locked_axpby")

    def flush_deltas(self) -> None:
        """Replicated table with deferred pushes: forward the deltas accumulated since the last
        flush to the other replicas (no-op for every other kind of table)."""
        if self.deferred:
            N.check(N.lib().fps_table_flush_delta(C.byref(self._view), N.current_stream()), "fps_table_flush_delta")

    def load(self, ids, vals) -> None:
        """Model load (``transform_with_model_load``): store the given values; rows become
        first-pushed, i.e. exactly the loaded value with no init added."""
        self.scatter(ids, vals, "overwrite")

    def dump(self):
        """``(ids, rows[:, :dim+aux])`` of the live rows of the local shard(s), sorted by id -
        the on-close model dump (``Right((id, value))`` records)."""
        import torch
        if self.replicated:
            # replicas hold the same values; which rows are live is the union of every replica's touches
            import torch.distributed as dist
            self.flush_deltas()
            torch.cuda.synchronize()
            self.ctx.barrier()
            if self.lazy:
                dist.all_reduce(self.state[0], op=dist.ReduceOp.MAX, group=self.ctx.group)
            if self.ctx.rank != 0:
                return (torch.empty(0, device="cuda", dtype=torch.int64),
                        torch.empty((0, self.dim + self.aux_cols), device="cuda", dtype=torch.float32))
        ids_out, rows_out = [], []
        for s in self.local_shards:
            cap = self.rows_per_shard
            d_ids = torch.empty(cap, device="cuda", dtype=torch.int64)
            d_rows = torch.empty((cap, self.stride), device="cuda", dtype=torch.float32)
            cnt = torch.zeros(1, device="cuda", dtype=torch.int64)
            N.check(N.lib().fps_table_dump_shard(C.byref(self._view), s, self.rows_per_shard, self.n_ids,
                                                 N.ptr(d_ids), N.ptr(d_rows), cap, N.ptr(cnt),
                                                 N.current_stream()), "fps_table_dump_shard")
            k = int(cnt.item())
            ids_out.append(d_ids[:k])
            rows_out.append(d_rows[:k, : self.dim + self.aux_cols])
        ids = torch.cat(ids_out)
        rows = torch.cat(rows_out)
        order = torch.argsort(ids)
        return ids[order], rows[order]

    # -- streaming checkpoint -----------------------------------------------------------------------
    def dump_chunks(self, chunk_rows: int = 1 << 18):
        """The live rows of the local shard(s) as a stream of ``(ids int64[k], rows fp32[k, dim + aux])``
        numpy pairs, at most ``chunk_rows`` slots at a time through two pinned staging buffers: a table of
        any size is dumped without ever being materialised on the host or turned into Python objects
        (the on-close dump of ``SimplePSLogicWithClose.scala:29-31``, streamed).  The yielded arrays are
        views of the staging buffers: consume (write out / copy) each pair before asking for the next."""
        import torch
        if self.replicated:
            self.flush_deltas()
            torch.cuda.synchronize()
            self.ctx.barrier()
            if self.ctx.rank != 0:
                return
        cols = self.dim + self.aux_cols
        d_ids = [torch.empty(chunk_rows, device="cuda", dtype=torch.int64) for _ in range(2)]
        d_rows = [torch.empty((chunk_rows, self.stride), device="cuda", dtype=torch.float32) for _ in range(2)]
        d_cnt = [torch.zeros(1, device="cuda", dtype=torch.int64) for _ in range(2)]
        h_ids = [torch.empty(chunk_rows, dtype=torch.int64).pin_memory() for _ in range(2)]
        h_rows = [torch.empty((chunk_rows, self.stride), dtype=torch.float32).pin_memory() for _ in range(2)]
        h_cnt = [torch.zeros(1, dtype=torch.int64).pin_memory() for _ in range(2)]
        done = [torch.cuda.Event() for _ in range(2)]
        work = [(s, b) for s in self.local_shards for b in range(0, self.rows_per_shard, chunk_rows)]

        def launch(k):
            s, b = work[k]
            j = k & 1
            d_cnt[j].zero_()
            N.check(N.lib().fps_table_dump_range(C.byref(self._view), s, b, min(b + chunk_rows, self.rows_per_shard),
                                                 self.n_ids, N.ptr(d_ids[j]), N.ptr(d_rows[j]), chunk_rows,
                                                 N.ptr(d_cnt[j]), N.current_stream()), "fps_table_dump_range")
            h_cnt[j].copy_(d_cnt[j], non_blocking=True)
            h_ids[j].copy_(d_ids[j], non_blocking=True)
            h_rows[j].copy_(d_rows[j], non_blocking=True)
            done[j].record()

        if work:
            launch(0)
        for k in range(len(work)):
            j = k & 1
            done[j].synchronize()
            n = int(h_cnt[j][0])
            ids_np, rows_np = h_ids[j][:n].numpy(), h_rows[j][:n, :cols].numpy()
            if k + 1 < len(work):
                launch(k + 1)            # the other buffer: its copy overlaps the consumer's work on this one
            if n:
                yield ids_np, rows_np

    def dump_to(self, path: str, chunk_rows: int = 1 << 18) -> int:
        """Streams the local live rows into ``path`` (binary: a JSON header line, then per chunk
        ``int64 k, k int64 ids, k * cols fp32 values``, closed by ``k = 0``).  Returns the row count."""
        import json
        import numpy as np
        total = 0
        cols = self.dim + self.aux_cols
        with open(path, "wb") as f:
            f.write((json.dumps({"format": "fps_b200.table.v1", "cols": cols, "dim": self.dim, "n_ids": self.n_ids,
                                 "rank": self.ctx.rank, "world": self.ctx.world_size}) + "\n").encode())
            for ids, rows in self.dump_chunks(chunk_rows):
                f.write(np.int64(len(ids)).tobytes())
                f.write(np.ascontiguousarray(ids).tobytes())
                f.write(np.ascontiguousarray(rows).tobytes())
                total += len(ids)
            f.write(np.int64(0).tobytes())
        return total

    def load_from(self, path: str) -> int:
        """Streams a ``dump_to`` file back in (``transformWithModelLoad``: stored values replace whatever
        was there; rows of a worker-resident table that belong to another rank are skipped).  Any rank
        may load any file: PS-side rows are routed to their owner shard."""
        import json
        import numpy as np
        import torch
        total = 0
        with open(path, "rb") as f:
            hdr = json.loads(f.readline().decode())
            if hdr.get("format") != "fps_b200.table.v1" or hdr["cols"] != self.dim + self.aux_cols:
                raise ValueError(f"{path}: not a table dump of {self.dim + self.aux_cols} columns")
            while True:
                k = int(np.frombuffer(f.read(8), dtype=np.int64)[0])
                if k == 0:
                    break
                ids = torch.from_numpy(np.frombuffer(f.read(8 * k), dtype=np.int64).copy())
                rows = torch.from_numpy(np.frombuffer(f.read(4 * k * hdr["cols"]), dtype=np.float32)
                                        .reshape(k, hdr["cols"]).copy())
                if self.local_only and self.ctx.world_size > 1:
                    keep = self.owner_of(ids)[0] == self.ctx.rank
                    ids, rows = ids[keep], rows[keep]
                if ids.numel():
                    self.load(ids.pin_memory().cuda(non_blocking=True), rows.pin_memory().cuda(non_blocking=True))
                    total += ids.numel()
        torch.cuda.synchronize()
        return total

    def resync(self, src: int = 0) -> None:
        """Replicated table: make every replica bit-identical to rank ``src``'s (replicas apply the
        same adds in different orders, so they drift by fp32 rounding)."""
        if self.replicated:
            import torch
            import torch.distributed as dist
            torch.cuda.synchronize()
            self.ctx.barrier()
            dist.broadcast(self.rows[0], src=src, group=self.ctx.group)
            torch.cuda.synchronize()

    def expected_init(self, ids):
        """Host oracle of ``init(id)`` (numpy, fp32)."""
        return philox.ranged_uniform(ids, self.dim, self.init_range[0], self.init_range[1], self.seed,
                                     self.stream_tag)

    def close(self) -> None:
        for p in self._peer_ptrs:
            try:
                N.lib().fps_ipc_close_handle(p)
            except Exception:
                pass
        self._peer_ptrs = []


def new_flow():
    """A zeroed device FpsFlow block (256 bytes) as a uint8 tensor."""
    import torch
    return torch.zeros(N.FLOW_BYTES, device="cuda", dtype=torch.uint8)


def read_flow(flow) -> dict:
    """Copy the flow block to the host and decode it."""
    raw = bytes(flow.cpu().numpy().tobytes())
    f = N.FpsFlow.from_buffer_copy(raw)
    d = {name: getattr(f, name) for name, _ in N.FpsFlow._fields_ if name != "reserved"}
    d["combined_rows"] = int(f.reserved[0])      # item rows really pulled/pushed by the combine variant
    return d


class DeviceAppendLog:
    """A device-resident append buffer of ``(id, row)`` records with a host-side cursor, drained by the
    host in one copy: the per-update form of the output stream (``SimplePSLogic`` emits ``(id, value)``
    on every push, ``M/server/SimplePSLogic.scala:24``; the MF workers emit ``(user, vec)`` per rating,
    ``PSOfflineMatrixFactorizationWorker.scala:144``) without a Python object per record.  ``append``
    gathers the current values of ``ids`` from the table straight into the buffer; when it would not
    fit, the buffer is drained to pinned host memory first (kept as numpy chunks until ``drain``)."""

    def __init__(self, table: ShardedTable, capacity: int = 1 << 20):
        import torch
        self.table, self.capacity = table, int(capacity)
        self.ids = torch.empty(self.capacity, device="cuda", dtype=torch.int64)
        self.rows = torch.empty((self.capacity, table.stride), device="cuda", dtype=torch.float32)
        self.cursor = 0
        self._spilled: list = []
        self.records = 0

    def append(self, ids) -> None:
        import torch
        ids = ids.to(device="cuda", dtype=torch.int64).contiguous()
        n, pos = ids.numel(), 0
        while pos < n:
            if self.cursor == self.capacity:
                self._spill()
            take = min(n - pos, self.capacity - self.cursor)
            part = ids[pos:pos + take]
            self.ids[self.cursor:self.cursor + take].copy_(part)
            self.table.gather(part, out=self.rows[self.cursor:self.cursor + take])
            self.cursor += take
            pos += take
        self.records += n

    def _spill(self) -> None:
        if self.cursor:
            cols = self.table.dim + self.table.aux_cols
            self._spilled.append((self.ids[:self.cursor].cpu().numpy(), self.rows[:self.cursor, :cols].cpu().numpy()))
            self.cursor = 0

    def drain(self):
        """``(ids int64[n], rows fp32[n, dim + aux])`` numpy arrays of everything appended since the last
        drain, in append order."""
        import numpy as np
        self._spill()
        chunks, self._spilled = self._spilled, []
        cols = self.table.dim + self.table.aux_cols
        if not chunks:
            return np.zeros(0, np.int64), np.zeros((0, cols), np.float32)
        return np.concatenate([c[0] for c in chunks]), np.concatenate([c[1] for c in chunks])
