"""Streaming (online) MF on the device: a persistent kernel consumes a ring of ratings.

Parity: ``PSOnlineMatrixFactorizationWorker.scala:22-90`` driven by
``FlinkParameterServer.transform(..., iterationWaitTime)`` (``M/FlinkParameterServer.scala:331-335``):
records arrive continuously, the job ends after the feedback edge has been idle for
``iteration_wait_time`` ms (0 = run until stopped).  Here the ring, its tail and the stop flag
are in mapped pinned host memory (the producer is a host thread - "the data source"), the
consumer is ``fps_mf_online_stream``: CTAs claim chunks with an atomic on the device-side
head, idle time is measured with ``%globaltimer`` on the device, and the kernel leaves by
itself - no host round trip decides termination.  ``max_lifetime_ms`` is a hard cap so a lost
producer can never wedge the GPU.
"""
from __future__ import annotations

import ctypes as C
import time
from typing import Optional

import numpy as np

from ...ops import _native as N
from ...ops.table import read_flow


class _Mapped:
    def __init__(self, nbytes: int):
        h, d = C.c_void_p(), C.c_void_p()
        N.check(N.lib().fps_host_alloc_mapped(C.byref(h), C.byref(d), nbytes), "fps_host_alloc_mapped")
        self.h, self.d, self.nbytes = h.value, d.value, nbytes

    def array(self, dtype, count: int, offset: int = 0) -> np.ndarray:
        ct = np.ctypeslib.as_ctypes_type(np.dtype(dtype))
        return np.ctypeslib.as_array((ct * count).from_address(self.h + offset))

    def free(self) -> None:
        if self.h:
            N.lib().fps_host_free(self.h)
            self.h = 0


class OnlineStream:
    def __init__(self, mf, capacity: int = 1 << 16, iteration_wait_ms: int = 1000,
                 max_lifetime_ms: int = 60_000):
        import torch
        self.mf = mf
        self.capacity = int(capacity)
        self.iteration_wait_ms = int(iteration_wait_ms)
        self.max_lifetime_ms = int(max_lifetime_ms)
        self._ring = _Mapped(self.capacity * 12)
        self._ctl = _Mapped(64)
        self.ring_u = self._ring.array(np.int32, self.capacity, 0)
        self.ring_i = self._ring.array(np.int32, self.capacity, self.capacity * 4)
        self.ring_r = self._ring.array(np.float32, self.capacity, self.capacity * 8)
        self.tail = self._ctl.array(np.uint64, 1, 0)
        self.stop = self._ctl.array(np.int32, 1, 32)
        self.tail[0] = 0
        self.stop[0] = 0
        self._kernel_stream = torch.cuda.Stream()
        self._copy_stream = torch.cuda.Stream()
        self._done_host = torch.zeros(N.FLOW_BYTES, dtype=torch.uint8).pin_memory()
        self._fed = 0
        self._started = False

    def start(self) -> None:
        import torch
        mf = self.mf
        mf.flow.zero_()
        torch.cuda.synchronize()
        with torch.cuda.stream(self._kernel_stream):
            N.check(N.lib().fps_mf_online_stream(
                mf.items.view_ref(), mf.users.view_ref(), self._ring.d, self._ring.d + self.capacity * 4,
                self._ring.d + self.capacity * 8, self.capacity, self._ctl.d, self._ctl.d + 32,
                C.byref(mf.params), self.iteration_wait_ms, self.max_lifetime_ms, N.ptr(mf.flow),
                self._kernel_stream.cuda_stream), "fps_mf_online_stream")
        self._started = True

    def _retired(self) -> int:
        import torch
        with torch.cuda.stream(self._copy_stream):
            self._done_host.copy_(self.mf.flow, non_blocking=True)
        self._copy_stream.synchronize()
        f = N.FpsFlow.from_buffer_copy(bytes(self._done_host.numpy().tobytes()))
        return int(f.ring_done)

    def feed(self, users, items, ratings, timeout_s: float = 30.0) -> None:
        """Append records (CPU tensors / arrays).  Blocks while the ring is full (back-pressure)."""
        u = np.asarray(users, dtype=np.int32)
        i = np.asarray(items, dtype=np.int32)
        r = np.asarray(ratings, dtype=np.float32)
        n, pos = len(u), 0
        deadline = time.time() + timeout_s
        while pos < n:
            free = self.capacity - (self._fed - self._retired())
            if free <= 0:
                if time.time() > deadline:
                    raise TimeoutError("ring stayed full: the consumer kernel is not running")
                time.sleep(0.0005)
                continue
            take = min(free, n - pos)
            start = self._fed % self.capacity
            first = min(take, self.capacity - start)
            for ring, src in ((self.ring_u, u), (self.ring_i, i), (self.ring_r, r)):
                ring[start:start + first] = src[pos:pos + first]
                if take > first:
                    ring[0:take - first] = src[pos + first:pos + take]
            self._fed += take
            pos += take
            self.tail[0] = self._fed          # publish after the payload (x86 store order)

    def request_stop(self) -> None:
        self.stop[0] = 1

    def join(self) -> int:
        """Wait for the kernel to leave; returns its exit reason (1 idle, 2 stop, 3 lifetime)."""
        self._kernel_stream.synchronize()
        reason = read_flow(self.mf.flow)["exit_reason"]
        self._ring.free()
        self._ctl.free()
        return reason
