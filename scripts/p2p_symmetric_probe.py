#!/usr/bin/env python
"""The NVLink roofline of the PS access pattern when EVERY GPU does it at once (torchrun, N ranks):
each rank pulls random 512-byte rows from the next rank's shard (128-bit loads) and pushes a 512-byte
delta to the same rows (red.add.v4) - so each GPU's ingress carries the rows it pulls plus the deltas
pushed to it, and its egress the mirror image, like the fused MF kernels.  The streaming one-way probes
(load 759, red.add 687 GB/s) say nothing about this mix: requests and acknowledgements of one
direction share the wire with the payload of the other.  Prints GB/s per direction and writes
gpurun_out/p2p_symmetric_<N>gpu.json."""
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fps_b200.ops import _native as N  # noqa: E402
from fps_b200.ops.table import ShardedTable  # noqa: E402
from fps_b200.parallel.device_context import init_context  # noqa: E402


def main():
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    rows = 4 * 1024 * 1024                                   # 2 GiB of 512-byte rows per rank
    t = ShardedTable(rows * world, 128, init_range=(0.0, 0.0), ctx=ctx)          # VMM symmetric memory
    loc = torch.zeros((rows, 128), device="cuda")
    peer = t._view.rows[(rank + 1) % world]
    n = 1 << 22
    res = {}
    for label, flags in (("pull_only", 1), ("push_only", 2), ("pull+push", 3), ("pull+push+local_rw", 3 | 4 | 8),
                         ("push_only_bulk_reduce", 2 | 256), ("pull+push_bulk_reduce", 3 | 256)):
        torch.cuda.synchronize()
        ctx.barrier()
        ms = C.c_float()
        N.check(N.lib().fps_p2p_probe_mixed(peer, loc.data_ptr(), rows, n, flags | 128, 5, C.byref(ms), N.current_stream()))
        tms = torch.tensor([ms.value / 5], device="cuda", dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        per_dir = (bool(flags & 1) + bool(flags & 2)) * 512 * n / (float(tms.item()) * 1e-3) / 1e9
        res[label] = {"ms": float(tms.item()), "gbs_per_direction_per_gpu": per_dir,
                      "rows_per_sec_per_gpu": n / (float(tms.item()) * 1e-3)}
        if rank == 0:
            print(f"{label:22s} {float(tms.item()):7.3f} ms per {n} rows  -> {per_dir:7.1f} GB/s per direction", flush=True)
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump({"world": world, "rows": n, "row_bytes": 512, **res},
                  open(os.path.join(ROOT, "gpurun_out", f"p2p_symmetric_{world}gpu.json"), "w"), indent=1)
    ctx.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
