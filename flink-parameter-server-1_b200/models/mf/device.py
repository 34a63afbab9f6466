"""Matrix factorisation on the device engine: HBM-sharded factor tables + the fused
pull-compute-push kernel (``csrc/mf.cu``).

Layout follows the reference's two-sided split (``PSOfflineMatrixFactorizationWorker.scala:43``):
user vectors are *worker-resident* (rank ``user % W`` owns a user and receives its ratings, the
input partitioning of ``PSOfflineMatrixFactorization.scala:71-74``), item vectors live on the
*parameter server* shards (``item % S``).  Worker ``r`` and shard ``r`` share GPU ``r``.

What one call to ``step`` replaces in the reference: for every rating one Pull message, one
PullAnswer carrying the item row, the SGD step, one Push carrying the delta and one output
record - here a P2P row load, registers, and a 16-byte-per-lane ``red.add``.

Semantics kept: asynchronous / stale reads (Hogwild inside a launch, no barrier between
ranks), ``param_update = add``, lazy ``param_init`` (optional), bounded in-flight pulls
(``pull_limit`` -> device credit counter), output as ``Left((user, vec))`` / ``Right((item, vec))``.
Output mode is "on close" (one record per live row, what every reference sink folds to).
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Iterable, Optional

import numpy as np

from ...entities import Left, Right
from ...ops import _native as N
from ...ops import philox
from ...ops.table import ShardedTable, new_flow, read_flow
from ...utils.observability import nvtx_range
from .vector import FactorIsNotANumberException

# direct  : one (sub-)warp per rating, rows in registers (csrc/mf.cu)
# tiles   : own counting sort by item bucket + per-CTA shared-memory sort, run-carrying SGD, rows staged
#           by the TMA engine (csrc/mf_tiles.cu) - the hot-key pre-combining path
# combine : the round-1 pre-combining path (cub radix sort + register-carried runs, csrc/mf_staged.cu);
#           kept for A/B comparisons only
VARIANTS = {"direct": 0, "combine": 2, "tiles": 3}
REPLICATE_BELOW_BYTES = 16 << 20
# "auto": see DeviceMF._variant_for.


class DeviceMF:
    def __init__(self, n_users: int, n_items: int, num_factors: int = 10, range_min: float = -0.01,
                 range_max: float = 0.01, learning_rate: float = 0.01, *, lambda_: float = 0.0,
                 pull_limit: int = 0, seed: int = 0, lazy: bool = False, variant: str = "auto",
                 ctx=None, n_shards: Optional[int] = None, replicate_items: Optional[bool] = None,
                 emit: str = "on_close", emit_capacity: int = 1 << 20):
        """``replicate_items`` (several GPUs, eager tables): every GPU keeps the whole item table, pulls
        stay local and the item deltas of a launch cross NVLink once per touched row
        (``ShardedTable(replicated=True)``).  Default: item tables up to ``REPLICATE_BELOW_BYTES``."""
        import torch
        from ...parallel.device_context import get_context
        self.ctx = ctx or get_context()
        self.n_users, self.n_items, self.num_factors = n_users, n_items, num_factors
        if replicate_items is None:
            # opt-in (FPS_MF_REPLICATE_ITEMS=1 or the argument): with deferred pushes every GPU's deltas of a
            # whole launch land on the others at once, i.e. hot rows take N steps from the same stale value.
            # SGD tolerates that at gentle learning rates but it is a different staleness profile from the
            # sharded table (whose pushes land immediately): measured stable on 2 GPUs with the direct
            # kernel (config 2), divergent with the combine kernel, whose pushes are whole-run deltas.
            replicate_items = (os.environ.get("FPS_MF_REPLICATE_ITEMS", "0") == "1" and self.ctx.distributed
                               and not lazy
                               and n_items * ((num_factors + 3) // 4 * 4) * 4 <= REPLICATE_BELOW_BYTES)
        if replicate_items and lazy:
            raise ValueError("a replicated item table must be eager (lazy=False)")
        self.items = ShardedTable(n_items, num_factors, n_shards=n_shards, partition="hash",
                                  init_range=(range_min, range_max), seed=seed,
                                  stream_tag=philox.STREAM_ITEM, lazy=lazy, ctx=self.ctx,
                                  replicated=bool(replicate_items))
        if self.items.replicated and not self.items.deferred:
            raise ValueError("replicated MF items need deferred pushes (unset FPS_REPLICA_PUSH=immediate)")
        self.users = ShardedTable(n_users, num_factors, n_shards=n_shards, partition="hash",
                                  init_range=(range_min, range_max), seed=seed,
                                  stream_tag=philox.STREAM_USER, lazy=lazy, ctx=self.ctx,
                                  local_only=True)
        self.params = N.FpsMfParams()
        self.params.lr, self.params.lambda_ = learning_rate, lambda_
        self.params.pull_limit = int(pull_limit)
        self.params.local_shard = 0 if self.items.replicated else (self.ctx.rank if self.ctx.distributed else -1)
        # "auto" picks per step (see _variant_for): the combine kernel whenever combining pays -
        #   * items sharded over several GPUs: every combined row is NVLink traffic saved;
        #   * a batch much larger than the item table (>= 4 ratings per item on average and >= 256K ratings,
        #     so the sort is amortised): otherwise the hot rows' atomics serialise (measured, rank 64,
        #     26.7K items, 1M-rating launches: 0.222 ms combine vs 0.387 ms direct);
        # never for lazy tables (the combine kernel has no life-cycle path) nor under a pull limiter (a
        # combine sub-warp holds its credits while it walks its slice serially: 109 ms vs ~10 ms per 4M
        # ratings at limit 1600).
        self._auto = variant == "auto"
        self._tiles_ok = (not lazy and not self.items.replicated and self.items.stride <= 512
                          and max(self.items.rows_per_shard, self.users.rows_per_shard) < (1 << 28))
        self._combine_ok = not lazy and pull_limit <= 0 and not self.items.replicated
        if variant == "tiles" and not self._tiles_ok:
            raise ValueError("variant='tiles' needs eager, sharded (not replicated) tables of rank <= 512")
        if variant == "combine" and not self._combine_ok:
            variant = "direct"
        self._prefer_tiles = os.environ.get("FPS_MF_PREFER_TILES", "0") == "1"
        if self._auto:
            variant = "tiles" if (self._tiles_ok and (self.ctx.distributed or self._prefer_tiles)) else "direct"
        self.variant = variant
        self.params.variant = VARIANTS[variant] if variant != "tiles" else 0
        self.flow = new_flow()
        self._copy_stream = None
        self._staging = None
        self._tile_slots = None
        self._tile_stream = None
        self._tile_slot = 0
        self.launches = 0
        self.kernel_launches = 0          # kernels of ours launched by step() so far
        # Output stream form.  "on_close": one record per live row at the end (dump(); what the ...WithClose
        # logics and every reference sink fold to).  "every_update": additionally one Right((item, vec)) per
        # push and one Left((user, vec)) per rating, appended on the device after every step
        # (SimplePSLogic.scala:24, PSOfflineMatrixFactorizationWorker.scala:144) and drained by drain_updates().
        if emit not in ("on_close", "every_update"):
            raise ValueError("emit must be 'on_close' or 'every_update'")
        self.emit = emit
        self._item_log = self._user_log = None
        if emit == "every_update":
            from ...ops.table import DeviceAppendLog
            self._item_log = DeviceAppendLog(self.items, emit_capacity)
            self._user_log = DeviceAppendLog(self.users, emit_capacity)

    def _variant_for(self, n: int) -> str:
        """``auto``: the tile path whenever combining pays -
          * items sharded over several GPUs: every combined row is NVLink traffic saved;
          * a batch much larger than the item table (>= 4 ratings per item on average and >= 256K
            ratings): otherwise the hot rows' atomics serialise;
        else the direct kernel.  Lazy and replicated tables always run the direct kernel."""
        if not self._tiles_ok:
            return "direct"
        dense = n >= 4 * self.n_items and n >= (1 << 18)
        return "tiles" if (dense or self.ctx.distributed or self._prefer_tiles) else "direct"

    # -- tile path plumbing -----------------------------------------------------------------------
    def _tile_scratch(self, n: int):
        """Two caller-owned scratch buffers (never allocated inside a step once they are large
        enough) and a side stream on which the bucketing of a step can run."""
        import torch
        need = int(N.lib().fps_mf_tiles_scratch_bytes(n))
        if self._tile_slots is None or self._tile_slots[0]["bytes"] < need:
            torch.cuda.synchronize()
            grow = need + need // 4
            self._tile_slots = [{"buf": torch.empty(grow, device="cuda", dtype=torch.uint8), "bytes": grow,
                                 "ready": torch.cuda.Event(), "free": torch.cuda.Event(), "used": False}
                                for _ in range(2)]
            if self._tile_stream is None:
                self._tile_stream = torch.cuda.Stream(priority=-1)
        s = self._tile_slots[self._tile_slot]
        self._tile_slot ^= 1
        return s

    def _bucket(self, users, items, ratings, n: int, ready_event=None, on_stream=None):
        """Phase 1 of a tile step.  ``ready_event`` (a ``torch.cuda.Event`` after which the inputs are
        valid) lets it run on the side stream while the previous step's kernel is still busy; without
        it the inputs are taken to be ordered after everything already queued on the current stream,
        previous step included."""
        import torch
        s = self._tile_scratch(n)
        main = torch.cuda.current_stream()
        side = on_stream or (self._tile_stream if ready_event is not None else main)
        if side is not main:
            if ready_event is not None:
                side.wait_event(ready_event)
            if s["used"]:
                side.wait_event(s["free"])           # the kernel that last read this scratch has finished
        N.check(N.lib().fps_mf_bucket(N.ptr(users), N.ptr(items), N.ptr(ratings), n, s["buf"].data_ptr(),
                                      s["bytes"], side.cuda_stream), "fps_mf_bucket")
        if side is not main:
            s["ready"].record(side)
            main.wait_event(s["ready"])
        self.kernel_launches += 3                 # k_bucket_count, k_bucket_scan, k_bucket_scatter
        return s

    def _run_tiles(self, s, n: int) -> None:
        import torch
        N.check(N.lib().fps_mf_sgd_tiles(self.items.view_ref(), self.users.view_ref(), s["buf"].data_ptr(), n,
                                         C.byref(self.params), N.ptr(self.flow), N.current_stream()),
                "fps_mf_sgd_tiles")
        s["free"].record(torch.cuda.current_stream())
        s["used"] = True
        self.kernel_launches += 2                 # k_tile_bounds + k_mf_tiles_tma / k_mf_tiles_reg

    # -- training ------------------------------------------------------------------------------
    def step(self, users, items, ratings, ready_event=None) -> None:
        """One fused pass over a batch already on the device (int32, int32, fp32).  With several
        ranks every user id in the batch must satisfy ``user % world == rank``.
        ``ready_event``: optional ``torch.cuda.Event`` after which the three tensors are valid; with it
        the tile path groups this batch on a side stream while the previous step is still running."""
        n = users.numel()
        if n == 0:
            return
        with nvtx_range("fps.mf.step"):
            if self._auto:
                self.variant = self._variant_for(n)
            if self.variant == "tiles":
                s = self._bucket(users, items, ratings, n, ready_event)
                self._run_tiles(s, n)
            else:
                self.params.variant = VARIANTS[self.variant]
                N.check(N.lib().fps_mf_sgd_step(self.items.view_ref(), self.users.view_ref(), N.ptr(users),
                                                N.ptr(items), N.ptr(ratings), n, C.byref(self.params),
                                                N.ptr(self.flow), N.current_stream()), "fps_mf_sgd_step")
                self.kernel_launches += 3 if self.variant == "combine" else 1
            self.items.flush_deltas()            # replicated items: this launch's summed deltas -> the other replicas
            if self._item_log is not None:       # values as of the end of this step, one record per push / rating
                keep = users != N.SKIP_ID
                self._item_log.append(items[keep] if not bool(keep.all()) else items)
                self._user_log.append(users[keep] if not bool(keep.all()) else users)
        self.launches += 1

    def _ensure_staging(self, n: int):
        import torch
        if self._staging is None or self._staging[0]["u"].numel() < n:
            self._staging = [
                {"u": torch.empty(n, device="cuda", dtype=torch.int32),
                 "i": torch.empty(n, device="cuda", dtype=torch.int32),
                 "r": torch.empty(n, device="cuda", dtype=torch.float32),
                 "copied": torch.cuda.Event(), "done": torch.cuda.Event(), "used": False,
                 "loss": torch.zeros(2, dtype=torch.float64).pin_memory()}
                for _ in range(2)]
            self._copy_stream = torch.cuda.Stream()
            self._slot = 0
            self._last_mse = float("nan")
        return self._staging

    def _read_loss(self, st) -> float:
        import torch
        loss_sum = float(st["loss"][0])
        records = int(st["loss"][1:2].view(torch.int64)[0])
        return loss_sum / max(records, 1)

    def step_from_host(self, users_h, items_h, ratings_h) -> float:
        """End-to-end step through the public API.  Every call (a) copies this step's pinned
        host inputs to the device on a copy stream, (b) runs the fused kernel, (c) copies the
        running loss (16 bytes of the flow block) back to pinned host memory.  Two staging slots
        pipeline it: the call only blocks on the step that used this slot two calls ago, so the
        H2D copy of step k+1 overlaps the kernel of step k.  Returns the mean squared error as of
        the most recent *retired* step (``flush_host_steps()`` returns the final one)."""
        import torch
        n = users_h.numel()
        stg = self._ensure_staging(n)
        st = stg[self._slot]
        self._slot ^= 1
        if st["used"]:
            st["done"].synchronize()                 # host waits for step k-2 only
            self._last_mse = self._read_loss(st)
        main = torch.cuda.current_stream()
        with torch.cuda.stream(self._copy_stream):
            st["u"][:n].copy_(users_h, non_blocking=True)
            st["i"][:n].copy_(items_h, non_blocking=True)
            st["r"][:n].copy_(ratings_h, non_blocking=True)
            st["copied"].record(self._copy_stream)
        if self._auto:
            self.variant = self._variant_for(n)
        if self.variant == "tiles":
            # the grouping of this batch runs on the copy stream right behind its H2D copies, i.e. while
            # the previous step's kernel is still busy on the main stream
            s = self._bucket(st["u"][:n], st["i"][:n], st["r"][:n], n, on_stream=self._copy_stream)
            self._run_tiles(s, n)
            self.launches += 1
        else:
            main.wait_event(st["copied"])
            self.step(st["u"][:n], st["i"][:n], st["r"][:n])
        view = self.flow[N.FLOW_LOSS_OFFSET:N.FLOW_LOSS_OFFSET + 16].view(torch.float64)
        st["loss"].copy_(view, non_blocking=True)
        st["done"].record(main)
        st["used"] = True
        return self._last_mse

    def flush_host_steps(self) -> float:
        import torch
        torch.cuda.current_stream().synchronize()
        if self._staging:
            last = self._staging[self._slot ^ 1]
            if last["used"]:
                self._last_mse = self._read_loss(last)
        return self._last_mse

    H2D_BYTES_PER_RATING = 12
    D2H_BYTES_PER_STEP = 16

    # -- evaluation / output ---------------------------------------------------------------------
    def predict(self, users, items):
        import torch
        out = torch.empty(users.numel(), device="cuda", dtype=torch.float32)
        N.check(N.lib().fps_mf_predict(self.items.view_ref(), self.users.view_ref(), N.ptr(users),
                                       N.ptr(items), users.numel(), N.ptr(out), N.current_stream()),
                "fps_mf_predict")
        return out

    def rmse(self, users, items, ratings) -> float:
        import torch
        p = self.predict(users, items)
        return float(torch.sqrt(torch.mean((p - ratings) ** 2)))

    def stats(self, check_nan: bool = True) -> dict:
        s = read_flow(self.flow)
        if check_nan and s["nan_count"]:
            raise FactorIsNotANumberException(f"{s['nan_count']} ratings produced a non-finite error")
        return s

    def reset_stats(self) -> None:
        self.flow.zero_()

    def dump(self) -> list:
        """The job's output stream in on-close mode: ``Left((user, vec))`` for every live local
        user row and ``Right((item, vec))`` for every live row of the local item shard."""
        out = []
        uid, urows = self.users.dump()
        for k, v in zip(uid.tolist(), urows.cpu().numpy().astype(np.float64)):
            out.append(Left((k, v)))
        iid, irows = self.items.dump()
        for k, v in zip(iid.tolist(), irows.cpu().numpy().astype(np.float64)):
            out.append(Right((k, v)))
        return out

    def drain_updates(self, as_stream: bool = False):
        """``emit="every_update"``: everything emitted since the last drain, in emission order, as
        ``((user_ids, user_rows), (item_ids, item_rows))`` numpy arrays - or, with ``as_stream``, as the
        reference's ``Left((user, vec))`` / ``Right((item, vec))`` records."""
        if self._item_log is None:
            raise ValueError("this model was built with emit='on_close'")
        import torch
        torch.cuda.current_stream().synchronize()
        u, i = self._user_log.drain(), self._item_log.drain()
        if not as_stream:
            return u, i
        return ([Left((int(k), v.astype(np.float64))) for k, v in zip(*u)] +
                [Right((int(k), v.astype(np.float64))) for k, v in zip(*i)])

    def save(self, prefix: str, chunk_rows: int = 1 << 18) -> dict:
        """Streaming checkpoint: every rank writes ``<prefix>.users.r<rank>`` and ``<prefix>.items.r<rank>``
        (its worker-resident user rows and its shard of the item table) chunk by chunk; no Python object per
        row, no host copy of a whole table."""
        import torch
        torch.cuda.synchronize()
        self.ctx.barrier()
        r = self.ctx.rank
        out = {"users": self.users.dump_to(f"{prefix}.users.r{r}", chunk_rows),
               "items": self.items.dump_to(f"{prefix}.items.r{r}", chunk_rows)}
        self.ctx.barrier()
        return out

    def restore(self, prefix: str, world: Optional[int] = None) -> dict:
        """Loads a ``save`` written by ``world`` ranks (default: as many as now).  Item rows are pushed to
        their owner shards whatever the sharding was; user rows are kept by the rank that owns them now."""
        import torch
        world = world or self.ctx.world_size
        cur, me = self.ctx.world_size, self.ctx.rank
        n_u = n_i = 0
        if world == cur:                        # same layout: every rank reads back its own two files
            n_u = self.users.load_from(f"{prefix}.users.r{me}")
            n_i = self.items.load_from(f"{prefix}.items.r{me}")
        else:
            for r in range(world):
                n_u += self.users.load_from(f"{prefix}.users.r{r}")    # rows of users owned elsewhere are skipped
                if r % cur == me:                                      # item files are shared out; rows go to their owners
                    n_i += self.items.load_from(f"{prefix}.items.r{r}")
        torch.cuda.synchronize()
        self.ctx.barrier()
        return {"users": n_u, "items": n_i}

    def load(self, model: Iterable) -> None:
        """Warm start from a previous output stream (``transform_with_double_model_load``):
        ``Left((user, vec))`` -> user table, ``Right((item, vec))`` -> item table.  Rows owned by
        another rank are skipped for the worker-resident half."""
        import torch
        us, uv, its, iv = [], [], [], []
        for e in model:
            k, v = e.value
            (us if e.is_left else its).append(k)
            (uv if e.is_left else iv).append(np.asarray(v, dtype=np.float32))
        W = self.ctx.world_size
        if us:
            ids = torch.tensor(us, dtype=torch.int64)
            vals = torch.from_numpy(np.stack(uv))
            if W > 1:
                keep = (ids % W) == self.ctx.rank
                ids, vals = ids[keep], vals[keep]
            if ids.numel():
                self.users.load(ids, vals)
        if its:
            self.items.load(torch.tensor(its, dtype=torch.int64), torch.from_numpy(np.stack(iv)))
        torch.cuda.synchronize()
        self.ctx.barrier()


class NegativeSampler:
    """Device-resident negative sampler of one worker (``csrc/sampler.cu``): per-user ring of the
    last ``user_memory`` items, pool of the items seen so far, Philox rejection draws.

    Parity: the ingest half of ``PSOnlineMatrixFactorizationWorker.onRecv`` (``:59-89``) and of
    ``PSOfflineMatrixFactorizationWorker.onRecv`` (``:64-95``).  ``expand`` turns a batch of positives
    into ``n * (rate + 1)`` records - first negatives of every rating, second negatives, ..., then the
    ratings themselves (records of one user are kept apart: neighbours run concurrently) - with unused
    slots marked ``SKIP_ID`` (the SGD kernels pass over them)."""

    def __init__(self, mf: "DeviceMF", negative_sample_rate: int, user_memory: int = 128, seed: int = 0):
        import torch
        if negative_sample_rate < 0 or user_memory <= 0:
            raise ValueError("negative_sample_rate >= 0 and user_memory > 0")
        self.mf, self.rate, self.user_memory = mf, int(negative_sample_rate), int(user_memory)
        users = mf.users
        distributed = mf.ctx.distributed
        n_rings = users.rows_per_shard if distributed else users.n_shards * users.rows_per_shard
        z = lambda n, dt: torch.zeros(max(int(n), 1), device="cuda", dtype=dt)   # noqa: E731
        self._seen = z(n_rings * self.user_memory, torch.int32)
        self._seen_count = z(n_rings, torch.int32)
        self._bitmap = z((mf.n_items + 31) // 32, torch.int32)
        self._pool = z(mf.n_items, torch.int32)
        self._pool_count = z(1, torch.int32)
        st = N.FpsNegSampler()
        st.seen, st.seen_count = self._seen.data_ptr(), self._seen_count.data_ptr()
        st.item_bitmap, st.item_pool = self._bitmap.data_ptr(), self._pool.data_ptr()
        st.pool_count = self._pool_count.data_ptr()
        st.n_items, st.seed = mf.n_items, (seed * 0x9E3779B97F4A7C15 + 0x1234567) & (2 ** 64 - 1)
        st.user_memory, st.rate = self.user_memory, self.rate
        st.local_shard = mf.ctx.rank if distributed else -1
        self._state = st
        self._tag = 0

    def expand(self, users, items, ratings):
        """``(users, items, ratings)`` int32 / int32 / fp32 on the device -> the expanded batch."""
        import torch
        n = users.numel()
        m = n * (self.rate + 1)
        ou = torch.empty(m, device="cuda", dtype=torch.int32)
        oi = torch.empty(m, device="cuda", dtype=torch.int32)
        orr = torch.empty(m, device="cuda", dtype=torch.float32)
        self._tag += 1
        N.check(N.lib().fps_mf_sample_negatives(self.mf.users.view_ref(), C.byref(self._state), N.ptr(users),
                                                N.ptr(items), N.ptr(ratings), n, self._tag, N.ptr(ou), N.ptr(oi),
                                                N.ptr(orr), N.current_stream()), "fps_mf_sample_negatives")
        return ou, oi, orr

    def items_seen(self) -> int:
        return int(self._pool_count.item())


def _ratings_to_tensors(ratings, rank: int, world: int):
    import torch
    u = np.fromiter((r.user for r in ratings), dtype=np.int32)
    i = np.fromiter((r.item for r in ratings), dtype=np.int32)
    v = np.fromiter((r.rating for r in ratings), dtype=np.float32)
    if world > 1:
        keep = (u % world) == rank        # input partitioning by user (X4)
        u, i, v = u[keep], i[keep], v[keep]
    return (torch.from_numpy(u).cuda(), torch.from_numpy(i).cuda(), torch.from_numpy(v).cuda())


def ps_offline_mf_device(src, num_factors: int = 10, range_min: float = -0.01, range_max: float = 0.01,
                         learning_rate: float = 0.01, negative_sample_rate: int = 0, user_memory: int = 128,
                         iterations: int = 1, pull_limit: int = 1600,
                         *, n_users: Optional[int] = None, n_items: Optional[int] = None, seed: int = 0,
                         shuffle: bool = True, lambda_: float = 0.0, lazy: bool = True,
                         conflict_free: bool = False, n_shards: Optional[int] = None,
                         variant: str = "auto", return_model: bool = False):
    """Device twin of ``ps_offline_mf`` (same arguments where they apply): multi-epoch SGD MF,
    returns the ``Left((user, vec))`` / ``Right((item, vec))`` stream.

    ``conflict_free=True`` splits every epoch into waves in which no user and no item repeats,
    which makes the asynchronous result equal to a serial schedule (used to check the kernel
    against the fp64 oracle); the default is one Hogwild launch per epoch."""
    import torch
    from ...parallel.device_context import get_context
    ctx = get_context()
    ratings = list(src)
    n_users = n_users or (max(r.user for r in ratings) + 1)
    n_items = n_items or (max(r.item for r in ratings) + 1)
    mf = DeviceMF(n_users, n_items, num_factors, range_min, range_max, learning_rate, lambda_=lambda_,
                  pull_limit=pull_limit, seed=seed, lazy=lazy, variant=variant, ctx=ctx, n_shards=n_shards)
    u, i, v = _ratings_to_tensors(ratings, ctx.rank, ctx.world_size)
    if negative_sample_rate > 0 and u.numel():
        # ingest phase of the reference worker (":64-95"): every rating becomes a block of its negatives +
        # itself, drawn once, in arrival order (micro-batches keep "items seen so far" meaningful); the
        # epochs then replay the blocks
        sampler = NegativeSampler(mf, negative_sample_rate, user_memory, seed)
        parts = [sampler.expand(u[s:s + 4096], i[s:s + 4096], v[s:s + 4096]) for s in range(0, u.numel(), 4096)]
        u, i, v = (torch.cat([p[k] for p in parts]) for k in range(3))
        keep = u != N.SKIP_ID
        u, i, v = u[keep].contiguous(), i[keep].contiguous(), v[keep].contiguous()
    gen = torch.Generator(device="cpu").manual_seed(seed)
    for _ in range(iterations):
        if shuffle and u.numel():
            perm = torch.randperm(u.numel(), generator=gen).cuda()
            u, i, v = u[perm], i[perm], v[perm]
        if conflict_free:
            for wu, wi, wv in conflict_free_waves(u, i, v):
                mf.step(wu, wi, wv)
        else:
            mf.step(u, i, v)
    torch.cuda.synchronize()
    ctx.barrier()
    mf.stats()
    out = mf.dump()
    return (out, mf) if return_model else out


def conflict_free_waves(u, i, v):
    """Greedy split of a batch into waves with pairwise distinct users and items, preserving the
    batch order across waves (wave k holds, for every record, the k-th occurrence rank)."""
    import torch
    uu, ii = u.cpu().numpy(), i.cpu().numpy()
    wave_of = np.zeros(len(uu), dtype=np.int64)
    next_u, next_i = {}, {}
    for k in range(len(uu)):
        w = max(next_u.get(int(uu[k]), 0), next_i.get(int(ii[k]), 0))
        wave_of[k] = w
        next_u[int(uu[k])] = w + 1
        next_i[int(ii[k])] = w + 1
    for w in range(int(wave_of.max()) + 1 if len(uu) else 0):
        idx = torch.from_numpy(np.nonzero(wave_of == w)[0]).cuda()
        yield u[idx], i[idx], v[idx]


class _HostSeenRing:
    """Host twin of the device sampler's per-user ring, for the streaming producer (which is the host):
    ``[n_users, user_memory]`` item ids, vectorised membership test."""

    def __init__(self, n_users: int, user_memory: int):
        self.ring = np.full((n_users, user_memory), -1, dtype=np.int32)
        self.count = np.zeros(n_users, dtype=np.int64)
        self.mem = user_memory

    def add(self, u: np.ndarray, i: np.ndarray) -> None:
        for a, b in zip(u.tolist(), i.tolist()):          # arrival order matters for the FIFO
            self.ring[a, self.count[a] % self.mem] = b
            self.count[a] += 1

    def contains(self, u: np.ndarray, cand: np.ndarray) -> np.ndarray:
        """``cand`` ``[n, k]`` -> bool ``[n, k]``: candidate is in its user's memory."""
        return (self.ring[u][:, None, :] == cand[:, :, None]).any(axis=2)


def ps_online_mf_device(src, num_factors: int = 10, range_min: float = -0.01, range_max: float = 0.01,
                        learning_rate: float = 0.01, negative_sample_rate: int = 9, user_memory: int = 128,
                        pull_limit: int = 1600, *, n_users: Optional[int] = None, n_items: Optional[int] = None,
                        batch_size: int = 65536, seed: int = 0, lazy: bool = True, streaming: bool = False,
                        iteration_wait_time: int = 1000, return_model: bool = False):
    """Device twin of ``ps_online_mf`` (single pass over a rating stream with negative sampling).

    Records are consumed in arrival order in micro-batches.  Per rating, up to ``negative_sample_rate``
    negatives (rating 0) are drawn from the items seen so far in the stream and rejected while they are
    among the user's last ``user_memory`` items (``PSOnlineMatrixFactorizationWorker.scala:59-89``) - on
    the device, by ``NegativeSampler`` (per-user rings + Philox rejection sampler in ``csrc/sampler.cu``);
    within a micro-batch all negatives come before the positives.
    ``streaming=True`` feeds the persistent ring-consumer kernel (``OnlineStream``) instead of one
    launch per micro-batch; the job then ends by the device-side idle timeout.  There the producer is
    the host, so the same sampling runs on the host (``_HostSeenRing``)."""
    import torch
    from ...parallel.device_context import get_context
    ctx = get_context()
    ratings = list(src) if not isinstance(src, list) else src
    n_users = n_users or (max(r.user for r in ratings) + 1)
    n_items = n_items or (max(r.item for r in ratings) + 1)
    mf = DeviceMF(n_users, n_items, num_factors, range_min, range_max, learning_rate, pull_limit=pull_limit,
                  seed=seed, lazy=lazy, variant="direct", ctx=ctx)
    u_all, i_all, r_all = _ratings_to_tensors(ratings, ctx.rank, ctx.world_size)
    if streaming:
        # The producer of a stream is the host: negatives are drawn with numpy and the records go straight
        # into the ring.  (Besides being the natural place, it keeps the GPU free of producer work - a
        # torch.randint on a CUDA generator was measured to wait for the resident consumer kernel to leave,
        # scripts/stream_concurrency_diag.py: 1.5 s = one idle timeout per call.)
        from .stream import OnlineStream
        stream = OnlineStream(mf, capacity=max(1 << 16, 2 * batch_size * (1 + negative_sample_rate)),
                              iteration_wait_ms=iteration_wait_time, max_lifetime_ms=600_000)
        rng = np.random.default_rng(seed * 31 + ctx.rank)
        u_np, i_np, r_np = u_all.cpu().numpy(), i_all.cpu().numpy(), r_all.cpu().numpy()
        seen_np = np.zeros(n_items, dtype=bool)
        memory = _HostSeenRing(n_users, user_memory) if negative_sample_rate > 0 else None
        for s in range(0, len(u_np), batch_size):
            u, i, r = u_np[s:s + batch_size], i_np[s:s + batch_size], r_np[s:s + batch_size]
            seen_np[i] = True
            pool = np.flatnonzero(seen_np).astype(np.int32)
            if memory is not None and pool.size > 1:
                memory.add(u, i)
                k = negative_sample_rate
                neg_i = pool[rng.integers(0, pool.size, size=(len(u), k))]
                bad = memory.contains(u, neg_i)
                for _ in range(4):                       # re-draw the rejected candidates a few times
                    if not bad.any():
                        break
                    neg_i[bad] = pool[rng.integers(0, pool.size, size=int(bad.sum()))]
                    bad = memory.contains(u, neg_i)
                keep = ~bad
                # draw-major order (all first negatives, ..., then the positives): the records of one user
                # must not sit next to each other, neighbours run concurrently on the device
                keep_t = keep.T.ravel()
                nu = np.repeat(u[None, :], k, 0).ravel()[keep_t]
                ni = neg_i.T.ravel()[keep_t]
                u = np.concatenate([nu, u])
                i = np.concatenate([ni, i])
                r = np.concatenate([np.zeros(len(nu), np.float32), r])
            stream.feed(u, i, r)
        stream.join()
    else:
        sampler = NegativeSampler(mf, negative_sample_rate, user_memory, seed) if negative_sample_rate > 0 else None
        for s in range(0, u_all.numel(), batch_size):
            u, i, r = u_all[s:s + batch_size], i_all[s:s + batch_size], r_all[s:s + batch_size]
            if sampler is not None:
                u, i, r = sampler.expand(u.contiguous(), i.contiguous(), r.contiguous())
            mf.step(u.contiguous(), i.contiguous(), r.contiguous())
    torch.cuda.synchronize()
    ctx.barrier()
    mf.stats()
    out = mf.dump()
    return (out, mf) if return_model else out
