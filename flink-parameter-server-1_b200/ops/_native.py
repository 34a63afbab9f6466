"""ctypes binding of ``libfps_b200.so`` (the C ABI declared in ``include/fps_b200.h``).

The library is built in-tree by ``build_native.py`` (``nvcc -gencode arch=compute_100a,code=sm_100a``)
and lives next to this file under ``_lib/`` so it travels with a repo snapshot.  There is no
PyTorch fallback for the device ops: on a CUDA box a missing library is an error, not a slow path.
"""
from __future__ import annotations

import ctypes as C
import os
import pathlib
from typing import Optional

FPS_MAX_SHARDS = 16
PART_HASH, PART_RANGE = 0, 1
ROW_UNTOUCHED, ROW_PULLED, ROW_PUSHED = 0, 1, 2

_LIB_DIR = pathlib.Path(__file__).resolve().parent / "_lib"
LIB_PATH = _LIB_DIR / "libfps_b200.so"


class FpsTableView(C.Structure):
    _fields_ = [
        ("rows", C.c_void_p * FPS_MAX_SHARDS),
        ("state", C.c_void_p * FPS_MAX_SHARDS),
        ("n_shards", C.c_int32),
        ("dim", C.c_int32),
        ("stride", C.c_int32),
        ("partition", C.c_int32),
        ("div", C.c_int64),
        ("rows_per_shard", C.c_int64),
        ("init_lo", C.c_float),
        ("init_hi", C.c_float),
        ("seed", C.c_uint64),
        ("stream_tag", C.c_int32),
        ("lazy", C.c_int32),
        ("mc_rows", C.c_void_p),
        ("delta_rows", C.c_void_p),
        ("replica_rows", C.c_void_p * FPS_MAX_SHARDS),
        ("n_replicas", C.c_int32),
        ("replica_rank", C.c_int32),
        ("n_ids", C.c_int64),
    ]


class FpsFlow(C.Structure):
    _fields_ = [
        ("in_flight", C.c_int32),
        ("max_in_flight", C.c_int32),
        ("limiter_stalls", C.c_uint64),
        ("pulls", C.c_uint64),
        ("pushes", C.c_uint64),
        ("remote_pulls", C.c_uint64),
        ("remote_pushes", C.c_uint64),
        ("nan_count", C.c_uint64),
        ("bad_ids", C.c_uint64),
        ("loss_sum", C.c_double),
        ("records", C.c_uint64),
        ("last_activity_ns", C.c_uint64),
        ("ring_head", C.c_uint64),
        ("ring_done", C.c_uint64),
        ("exit_reason", C.c_int32),
        ("combined_groups", C.c_int32),
        ("reserved", C.c_uint64 * 18),
    ]


assert C.sizeof(FpsFlow) == 256, C.sizeof(FpsFlow)
FLOW_BYTES = 256
FLOW_LOSS_OFFSET = FpsFlow.loss_sum.offset
FLOW_RECORDS_OFFSET = FpsFlow.records.offset


class FpsMfParams(C.Structure):
    _fields_ = [
        ("lr", C.c_float),
        ("lambda_", C.c_float),
        ("pull_limit", C.c_int32),
        ("local_shard", C.c_int32),
        ("variant", C.c_int32),
        ("unroll", C.c_int32),
        ("blocks_per_sm", C.c_int32),
        ("reserved", C.c_int32),
    ]


class FpsNegSampler(C.Structure):
    _fields_ = [
        ("seen", C.c_void_p),
        ("seen_count", C.c_void_p),
        ("item_bitmap", C.c_void_p),
        ("item_pool", C.c_void_p),
        ("pool_count", C.c_void_p),
        ("n_items", C.c_int64),
        ("seed", C.c_uint64),
        ("user_memory", C.c_int32),
        ("rate", C.c_int32),
        ("local_shard", C.c_int32),
        ("reserved", C.c_int32),
        ("n_rings", C.c_int64),
    ]


SKIP_ID = -2147483648


class FpsPaParams(C.Structure):
    _fields_ = [
        ("variant", C.c_int32),
        ("C", C.c_float),
        ("multiclass", C.c_int32),
        ("n_classes", C.c_int32),
        ("pull_limit", C.c_int32),
        ("local_shard", C.c_int32),
        ("predict_only", C.c_int32),
        ("reserved", C.c_int32),
    ]


class NativeError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None


def available() -> bool:
    return LIB_PATH.exists()


def lib() -> C.CDLL:
    """Load (once) and return the native library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise NativeError(
            f"{LIB_PATH} is missing - run `python build_native.py` (or __graft_entry__.build()) first; "
            "the device ops have no PyTorch fallback")
    L = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
    vp, i32, i64, u64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_float
    P = C.POINTER
    sig = {
        "fps_last_error": (C.c_char_p, []),
        "fps_version": (C.c_int, []),
        "fps_malloc": (C.c_int, [P(vp), C.c_size_t]),
        "fps_free": (C.c_int, [vp]),
        "fps_memset": (C.c_int, [vp, C.c_int, C.c_size_t, vp]),
        "fps_ipc_get_handle": (C.c_int, [vp, C.c_char_p]),
        "fps_ipc_open_handle": (C.c_int, [C.c_char_p, P(vp)]),
        "fps_ipc_close_handle": (C.c_int, [vp]),
        "fps_enable_peer_access": (C.c_int, [C.c_int]),
        "fps_host_alloc_mapped": (C.c_int, [P(vp), P(vp), C.c_size_t]),
        "fps_host_free": (C.c_int, [vp]),
        "fps_p2p_probe": (C.c_int, [vp, vp, C.c_size_t, C.c_int, C.c_int, P(f32), vp]),
        "fps_p2p_probe_random": (C.c_int, [vp, i64, C.c_int, i64, C.c_int, C.c_int, C.c_int, C.c_int, P(f32), vp]),
        "fps_p2p_probe_mixed": (C.c_int, [vp, vp, i64, i64, C.c_int, C.c_int, P(f32), vp]),
        "fps_table_init_shard": (C.c_int, [P(FpsTableView), C.c_int, i64, vp]),
        "fps_table_gather": (C.c_int, [P(FpsTableView), vp, i64, vp, C.c_int, vp, vp]),
        "fps_table_scatter": (C.c_int, [P(FpsTableView), vp, i64, vp, C.c_int, C.c_int, vp, vp]),
        "fps_table_locked_axpby": (C.c_int, [P(FpsTableView), vp, i64, C.c_float, vp, i32, i32, i32, vp, vp]),
        "fps_table_dump_shard": (C.c_int, [P(FpsTableView), C.c_int, i64, i64, vp, vp, i64, vp, vp]),
        "fps_table_dump_range": (C.c_int, [P(FpsTableView), C.c_int, i64, i64, i64, vp, vp, i64, vp, vp]),
        "fps_mf_sgd_step": (C.c_int, [P(FpsTableView), P(FpsTableView), vp, vp, vp, i64, P(FpsMfParams), vp, vp]),
        "fps_mf_predict": (C.c_int, [P(FpsTableView), P(FpsTableView), vp, vp, i64, vp, vp]),
        "fps_mf_sample_negatives": (C.c_int, [P(FpsTableView), P(FpsNegSampler), vp, vp, vp, i64, u64, vp, vp, vp, vp]),
        "fps_mf_sample_negatives_dtag": (C.c_int, [P(FpsTableView), P(FpsNegSampler), vp, vp, vp, i64, vp, vp, vp, vp, vp]),
        "fps_seen_ring_append": (C.c_int, [vp, vp, i32, i64, vp, vp, i64, vp]),
        "fps_mark_live_items": (C.c_int, [P(FpsTableView), i32, vp, vp, i64, vp, vp]),
        "fps_mf_tiles_scratch_bytes": (C.c_size_t, [i64]),
        "fps_mf_bucket": (C.c_int, [vp, vp, vp, i64, vp, C.c_size_t, vp]),
        "fps_mf_sgd_tiles": (C.c_int, [P(FpsTableView), P(FpsTableView), vp, i64, P(FpsMfParams), vp, vp]),
        "fps_mf_online_stream": (C.c_int, [P(FpsTableView), P(FpsTableView), vp, vp, vp, u64, vp, vp,
                                           P(FpsMfParams), C.c_uint32, C.c_uint32, vp, vp]),
        "fps_pa_step": (C.c_int, [P(FpsTableView), vp, vp, vp, vp, i64, P(FpsPaParams), vp, vp, vp]),
        "fps_pa_step_cost": (C.c_int, [P(FpsTableView), vp, vp, vp, vp, i64, P(FpsPaParams), vp, vp, vp, vp]),
        "fps_table_flush_delta": (C.c_int, [vp, vp]),
        "fps_topk_scratch_bytes": (C.c_size_t, [i64, i32, i64]),
        "fps_topk_scratch_bytes_full": (C.c_size_t, [i64, i32, i64]),
        "fps_lemp_topk": (C.c_int, [vp, i32, i64, vp, i32, vp, vp, i64, i32, i32, i32, vp, C.c_size_t,
                                    vp, vp, vp, vp]),
        "fps_topk_packed_bytes": (C.c_size_t, [i64, i32]),
        "fps_topk_packed_scratch_bytes": (C.c_size_t, [i64, i32, i64]),
        "fps_topk_pack_items": (C.c_int, [vp, i32, i64, i32, vp, vp]),
        "fps_lemp_topk_packed": (C.c_int, [vp, i32, i64, vp, vp, vp, i64, i32, i32, i32, vp, C.c_size_t,
                                           vp, vp, vp, vp]),
        "fps_topk_wide_packed_bytes": (C.c_size_t, [i64, i32]),
        "fps_topk_wide_scratch_bytes": (C.c_size_t, [i64, i32, i64]),
        "fps_topk_pack_items_wide": (C.c_int, [vp, i32, i64, i32, vp, vp]),
        "fps_topk_pack_items_wide_live": (C.c_int, [vp, i32, i64, i32, vp, vp, vp]),
        "fps_lemp_topk_wide": (C.c_int, [vp, i32, i64, vp, vp, vp, i64, i32, i32, i32, vp, C.c_size_t,
                                         vp, vp, vp, vp]),
        "fps_lemp_topk_static": (C.c_int, [vp, i32, i64, vp, i32, vp, vp, i64, i32, i32, i32, vp, vp, vp, vp]),
        "fps_topk_simt": (C.c_int, [vp, i32, i64, vp, i32, vp, i64, i32, i32, vp, C.c_size_t, vp, vp, vp]),
        "fps_topk_merge": (C.c_int, [vp, vp, i32, i64, i32, i32, vp, i32, vp, vp, vp]),
        "fps_umma_probe": (C.c_int, [i32, i32, i32, vp, vp]),
        "fps_lemp_keep_mask": (C.c_int, [vp, i32, i64, vp, i32, vp, i64, i32, vp, i32, i32, i32, C.c_float, vp, vp]),
        "fps_lemp_topk_filtered": (C.c_int, [vp, i32, i64, vp, i32, vp, vp, i64, i32, i32, i32, i32, i32, C.c_float,
                                             vp, vp, vp, vp, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name, None)
        if fn is None:
            continue          # optional entry points are checked at call time
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = lib().fps_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'native call'} failed (code {code}): {msg}")


def ptr(t) -> int:
    """Device pointer of a torch tensor (or 0 for None)."""
    return 0 if t is None else t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream


class DeviceBuffer:
    """A cudaMalloc'd buffer owned by this process and exportable through CUDA IPC.  Exposed to
    torch zero-copy via ``__cuda_array_interface__``."""

    def __init__(self, nbytes: int):
        p = C.c_void_p()
        check(lib().fps_malloc(C.byref(p), max(int(nbytes), 16)), f"fps_malloc({nbytes})")
        self.ptr = p.value
        self.nbytes = int(nbytes)
        self._freed = False

    def tensor(self, shape, dtype_str: str = "<f4"):
        import torch

        class _Iface:
            pass

        holder = _Iface()
        holder.__cuda_array_interface__ = {
            "shape": tuple(int(s) for s in shape), "typestr": dtype_str,
            "data": (self.ptr, False), "version": 2, "strides": None,
        }
        t = torch.as_tensor(holder, device="cuda")
        t._fps_owner = self            # keep the allocation alive as long as the view is
        return t

    def ipc_handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        check(lib().fps_ipc_get_handle(self.ptr, buf), "fps_ipc_get_handle")
        return buf.raw

    def free(self) -> None:
        if not self._freed and self.ptr:
            self._freed = True
            try:
                lib().fps_free(self.ptr)
            except Exception:
                pass

    def __del__(self):
        self.free()


def open_ipc(handle: bytes) -> int:
    p = C.c_void_p()
    check(lib().fps_ipc_open_handle(handle, C.byref(p)), "fps_ipc_open_handle")
    return p.value
