#!/usr/bin/env python
"""Random-row NVLink probe (torchrun, >= 2 ranks): GB/s of the PS access pattern - one warp per
random 512-byte row - against local HBM and a peer shard, for two ways of mapping the peer:
cudaMalloc + CUDA IPC, and torch symmetric memory (cuMem VMM).  Writes gpurun_out/p2p_random.json.
"""
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fps_b200.ops import _native as N  # noqa: E402
from fps_b200.parallel.device_context import init_context  # noqa: E402


def probe(ptr, span_rows, row_floats, n_rows, mode, unroll=2, bps=8, iters=3):
    L = N.lib()
    L.fps_p2p_probe_random.restype = C.c_int
    L.fps_p2p_probe_random.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.POINTER(C.c_float), C.c_void_p]
    ms = C.c_float()
    N.check(L.fps_p2p_probe_random(ptr, span_rows, row_floats, n_rows, mode, unroll, bps, iters, C.byref(ms),
                                   N.current_stream()), "probe")
    passes = 2 if mode == 2 else 1
    return n_rows * row_floats * 4 * passes * iters / (ms.value * 1e-3) / 1e9


def main():
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    row_floats = 128
    total_rows = 5 * 1024 * 1024          # 2.5 GiB per shard, like the 10M x 128 item table on 2 GPUs
    nbytes = total_rows * row_floats * 4
    res = {}

    # (a) cudaMalloc + IPC
    buf = N.DeviceBuffer(nbytes)
    t = buf.tensor((total_rows, row_floats))
    t.zero_()
    handles = ctx.all_gather_object(buf.ipc_handle())
    peer_ipc = N.open_ipc(handles[(rank + 1) % world]) if world > 1 else buf.ptr
    torch.cuda.synchronize()
    ctx.barrier()

    # (b) symmetric memory (VMM)
    peer_symm = local_symm = None
    try:
        import torch.distributed._symmetric_memory as symm
        st = symm.empty((total_rows, row_floats), dtype=torch.float32, device=ctx.device)
        hdl = symm.rendezvous(st, dist.group.WORLD)
        st.zero_()
        local_symm = int(hdl.buffer_ptrs[rank])
        peer_symm = int(hdl.buffer_ptrs[(rank + 1) % world])
    except Exception as e:  # noqa: BLE001
        res["symm_error"] = repr(e)[:300]
    torch.cuda.synchronize()
    ctx.barrier()

    if rank == 0:
        n = 2 * 1024 * 1024
        targets = [("local", buf.ptr), ("peer_ipc", peer_ipc)]
        if peer_symm:
            targets += [("local_symm", local_symm), ("peer_symm", peer_symm)]
        for name, ptr in targets:
            for span_mb in (64, 1024, 2560):
                span_rows = span_mb * 1024 * 1024 // (row_floats * 4)
                for mode, mname in ((0, "pull"), (1, "push"), (2, "pull+push")):
                    key = f"{name}/span{span_mb}MB/{mname}"
                    res[key] = round(probe(ptr, span_rows, row_floats, n, mode), 1)
                    print(key, res[key], "GB/s", flush=True)
        # MLP sweep on the peer, widest span
        for name, ptr in targets[1:2] + targets[3:4]:
            for unroll in (1, 2, 4):
                for bps in (4, 8, 16):
                    key = f"{name}/span2560MB/pull/unroll{unroll}/bps{bps}"
                    res[key] = round(probe(ptr, total_rows, row_floats, n, 0, unroll, bps), 1)
                    print(key, res[key], "GB/s", flush=True)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", "p2p_random.json"), "w"), indent=1)
    ctx.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
