#!/usr/bin/env python
"""Why is the MF kernel slow on IPC-mapped peers under torchrun?  Cross-tests (2 ranks):
  A. mixed probe kernel on the ShardedTable's own peer pointer
  B. MF kernel on hand-made IPC buffers (no ShardedTable)
  C. MF kernel on the ShardedTable with local_shard = -1
"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fps_b200.models.mf.device import DeviceMF  # noqa: E402
from fps_b200.ops import _native as N  # noqa: E402
from fps_b200.parallel.device_context import init_context  # noqa: E402


def time_steps(fn, n=3):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    ctx = init_context()
    rank, world = ctx.rank, ctx.world_size
    L = N.lib()
    L.fps_p2p_probe_mixed.restype = C.c_int
    L.fps_p2p_probe_mixed.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                      C.POINTER(C.c_float), C.c_void_p]
    n_users, n_items, B = 10_000_000, 10_000_000, 1 << 21
    mf = DeviceMF(n_users, n_items, 128, -0.01, 0.01, 0.01, ctx=ctx)
    g = torch.Generator().manual_seed(5 + rank)
    users = (torch.randint(0, n_users // world, (B,), generator=g) * world + rank).int().cuda()
    base = torch.randint(0, n_items // world, (B,), generator=g)
    remote_items = (base * world + (rank + 1) % world).int().cuda()
    local_items = (base * world + rank).int().cuda()
    ratings = torch.rand(B, generator=g).cuda()
    torch.cuda.synchronize()
    dist.barrier()

    def report(name, ms):
        if rank == 0:
            print(f"{name}: {ms:.3f} ms per {B} rows", flush=True)

    if rank == 0:
        peer_ptr = mf.items._view.rows[(rank + 1) % world]
        loc_ptr = mf.users._view.rows[rank]
        ms = C.c_float()
        N.check(L.fps_p2p_probe_mixed(peer_ptr, loc_ptr, min(mf.items.rows_per_shard, mf.users.rows_per_shard), B, 1 | 2 | 4 | 8, 3, C.byref(ms),
                                      N.current_stream()))
        report("A probe(rLD+rRED+lLD+lST) on table pointers", ms.value / 3)
        for fl, nm in ((1 | 128, "rLD128"), (1, "rLD32"), (1 | 2 | 128, "rLD128+rRED"), (1 | 2 | 4 | 8 | 128, "rLD128+rRED+lLD128+lST"), (4 | 8 | 128, "lLD128+lST")):
            N.check(L.fps_p2p_probe_mixed(peer_ptr, loc_ptr, min(mf.items.rows_per_shard, mf.users.rows_per_shard), B, fl, 3, C.byref(ms), N.current_stream()))
            report("A probe " + nm, ms.value / 3)
        report("C0 MF all_local", time_steps(lambda: mf.step(users, local_items, ratings)))
        report("C1 MF all_remote (table, local_shard=rank)", time_steps(lambda: mf.step(users, remote_items, ratings)))
        for dbg, nm in () if not hasattr(L, "fps_debug_set") else ((1, "no item load"), (2, "no item red"), (3, "no item load/red"), (4 | 8, "no user load/store"),
                        (1 | 4 | 8, "only item red"), (2 | 4 | 8, "only item load"), (8, "no user store"), (4, "no user load")):
            N.check(L.fps_debug_set(dbg))
            report(f"C MF all_remote dbg[{nm}]", time_steps(lambda: mf.step(users, remote_items, ratings)))
        if hasattr(L, "fps_debug_set"):
            N.check(L.fps_debug_set(0))
        for span in (1 << 17, 1 << 19, 1 << 21, 5_000_000):
            it = ((base % span) * world + (rank + 1) % world).int().cuda()
            report(f"D MF all_remote span {span * 512 >> 20} MB", time_steps(lambda: mf.step(users, it, ratings)))
            N.check(L.fps_p2p_probe_mixed(peer_ptr, loc_ptr, span, B, 1 | 2 | 4 | 8 | 128, 3, C.byref(ms), N.current_stream()))
            report(f"D probe rLD128+rRED+lLD128+lST span {span * 512 >> 20} MB", ms.value / 3)
    torch.cuda.synchronize()
    dist.barrier()

    # B: hand-made buffers
    rows = n_items // world
    buf = N.DeviceBuffer(rows * 512)
    buf.tensor((rows, 128)).zero_()
    handles = ctx.all_gather_object(buf.ipc_handle())
    peer = N.open_ipc(handles[(rank + 1) % world])
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        saved = (mf.items._view.rows[0], mf.items._view.rows[1])
        mf.items._view.rows[rank] = buf.ptr
        mf.items._view.rows[(rank + 1) % world] = peer
        report("B MF all_remote (hand-made IPC buffers)", time_steps(lambda: mf.step(users, remote_items, ratings)))
        mf.items._view.rows[0], mf.items._view.rows[1] = saved
        report("B2 MF all_remote (table again)", time_steps(lambda: mf.step(users, remote_items, ratings)))
        ms = C.c_float()
        N.check(L.fps_p2p_probe_mixed(peer, mf.users._view.rows[rank], min(rows, mf.users.rows_per_shard), B, 1 | 2 | 4 | 8, 3, C.byref(ms),
                                      N.current_stream()))
        report("A2 probe on hand-made buffers", ms.value / 3)
    torch.cuda.synchronize()
    dist.barrier()
    # E: the same kernel on torch symmetric memory (cuMem VMM + fd/fabric handles) instead of cudaIpc
    import torch.distributed._symmetric_memory as symm
    st = symm.empty((rows, 128), dtype=torch.float32, device=ctx.device)
    hdl = symm.rendezvous(st, dist.group.WORLD)
    st.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        for r in range(world):
            mf.items._view.rows[r] = int(hdl.buffer_ptrs[r])
        report("E MF all_remote full span (symmetric memory)", time_steps(lambda: mf.step(users, remote_items, ratings)))
        ms = C.c_float()
        N.check(L.fps_p2p_probe_mixed(int(hdl.buffer_ptrs[1]), mf.users._view.rows[rank], min(rows, mf.users.rows_per_shard), B,
                                      1 | 2 | 4 | 8 | 128, 3, C.byref(ms), N.current_stream()))
        report("E probe full span (symmetric memory)", ms.value / 3)
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
