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
        self._exit_seen = 0
        self.relaunches = 0

    def start(self, resume: bool = False) -> None:
        """Launch the consumer.  ``resume=True`` re-arms a consumer that left on its idle timeout while
        records were still to come: ring_head / ring_done / counters are kept, only exit_reason is reset."""
        import torch
        mf = self.mf
        if resume:
            self._kernel_stream.synchronize()
            off = N.FpsFlow.exit_reason.offset
            mf.flow.view(torch.uint8)[off:off + 4].zero_()
            self.relaunches += 1
        else:
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
        self._exit_seen = int(f.exit_reason)
        return int(f.ring_done)

    def feed(self, users, items, ratings, timeout_s: float = 30.0) -> None:
        """Append records (CPU tensors / arrays).  Blocks while the ring is full (back-pressure).
        The consumer kernel is launched by the first call: the device-side idle clock must not run
        while the producer is still getting ready (a consumer started earlier could time out before
        the first record arrives and leave the ring unread)."""
        if not self._started:
            self.start()
        u = np.asarray(users, dtype=np.int32)
        i = np.asarray(items, dtype=np.int32)
        r = np.asarray(ratings, dtype=np.float32)
        n, pos = len(u), 0
        deadline = time.time() + timeout_s
        while pos < n:
            free = self.capacity - (self._fed - self._retired())
            if self._exit_seen == 1:
                # the consumer timed out during a producer gap longer than iteration_wait_ms (first-use
                # module loads of the producer's own GPU work can serialise behind a running kernel);
                # it left with head == done, so it can simply be re-armed
                self.start(resume=True)
                self._exit_seen = 0
            elif self._exit_seen in (2, 3):
                raise RuntimeError("the consumer kernel has left (stop flag / max_lifetime_ms) while records "
                                   "were still being fed")
            if free <= 0:
                if time.time() > deadline:
                    raise TimeoutError("ring stayed full: the consumer kernel is not retiring records")
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
        if not self._started:
            return 0
        self._kernel_stream.synchronize()
        flow = read_flow(self.mf.flow)
        if flow["exit_reason"] == 1 and flow["ring_done"] < self._fed:      # left just before the last feed
            self.start(resume=True)
            self._kernel_stream.synchronize()
        reason = read_flow(self.mf.flow)["exit_reason"]
        self._ring.free()
        self._ctl.free()
        return reason
