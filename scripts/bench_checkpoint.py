#!/usr/bin/env python
"""Streaming checkpoint round trip of a 10M x 128 fp32 table (5.1 GB): dump_to() then load_from() into a
fresh table with another initialisation, rows compared bit for bit.  Prints one JSON line."""
import json
import os
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fps_b200.ops.table import ShardedTable  # noqa: E402
from fps_b200.parallel.device_context import init_context  # noqa: E402


def main():
    ctx = init_context()
    n, dim = int(os.environ.get("FPS_CKPT_ROWS", 10_000_000)), 128
    a = ShardedTable(n, dim, init_range=(-0.01, 0.01), seed=1, ctx=ctx)
    d = tempfile.mkdtemp(dir=os.environ.get("FPS_CKPT_DIR", "/dev/shm" if os.path.isdir("/dev/shm") else None))
    path = os.path.join(d, "table.r0")
    torch.cuda.synchronize()
    t0 = time.time()
    rows = a.dump_to(path, chunk_rows=1 << 18)
    t1 = time.time()
    b = ShardedTable(n, dim, init_range=(5.0, 6.0), seed=2, ctx=ctx)
    torch.cuda.synchronize()
    t2 = time.time()
    loaded = b.load_from(path)
    torch.cuda.synchronize()
    t3 = time.time()
    equal = all(torch.equal(a.rows[s], b.rows[s]) for s in a.local_shards)
    size = os.path.getsize(path)
    os.remove(path)
    os.rmdir(d)
    print(json.dumps({"what": "streaming checkpoint round trip", "rows": rows, "dim": dim, "file_gb": size / 1e9,
                      "dump_s": t1 - t0, "load_s": t3 - t2, "round_trip_s": (t1 - t0) + (t3 - t2),
                      "dump_gbs": size / 1e9 / (t1 - t0), "load_gbs": size / 1e9 / (t3 - t2),
                      "rows_loaded": loaded, "bit_equal": bool(equal), "dir": d}))


if __name__ == "__main__":
    main()
