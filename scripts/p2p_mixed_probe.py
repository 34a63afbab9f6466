#!/usr/bin/env python
"""Which mix of local / peer accesses inside one warp is slow?  (torchrun, 2 ranks; rank 0 drives)"""
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


def main():
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    rows = 4 * 1024 * 1024                  # 2 GiB of 512-byte rows
    buf = N.DeviceBuffer(rows * 512)
    buf.tensor((rows, 128)).zero_()
    loc = N.DeviceBuffer(rows * 512)
    loc.tensor((rows, 128)).zero_()
    handles = ctx.all_gather_object(buf.ipc_handle())
    peer = N.open_ipc(handles[(rank + 1) % world])
    torch.cuda.synchronize()
    ctx.barrier()
    L = N.lib()
    L.fps_p2p_probe_mixed.restype = C.c_int
    L.fps_p2p_probe_mixed.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                      C.POINTER(C.c_float), C.c_void_p]
    res = {}
    if rank == 0:
        n = 1 << 21
        names = {1: "rLD", 2: "rRED", 4: "lLD", 8: "lST", 16: "lRED", 32: "weak", 64: "redfirst"}
        combos = [1, 2, 3, 4 | 8, 1 | 4, 1 | 8, 1 | 4 | 8, 2 | 8, 2 | 4, 3 | 4, 3 | 8, 3 | 4 | 8, 3 | 4 | 8 | 32,
                  3 | 4 | 8 | 64, 3 | 16, 1 | 16, 2 | 16, 3 | 4 | 16]
        for target, tname in ((peer, "peer"), (buf.ptr, "self")):
            for f in combos:
                ms = C.c_float()
                N.check(L.fps_p2p_probe_mixed(target, loc.ptr, rows, n, f, 3, C.byref(ms), N.current_stream()))
                label = "+".join(v for k, v in names.items() if f & k)
                res[f"{tname}:{label}"] = round(ms.value / 3, 3)
                print(f"{tname}: {label:28s} {ms.value / 3:8.3f} ms per 2M rows", flush=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", "p2p_mixed.json"), "w"), indent=1)
    ctx.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
