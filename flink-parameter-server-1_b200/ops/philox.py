"""Counter-based parameter initialisation shared bit-for-bit by host and device.

The reference draws initial factors from a process-global unseeded RNG and ignores the id
(``M/matrix/factorization/factors/RangedRandomFactorInitializer.scala:7-9``), which is
irreproducible and cannot be evaluated independently on several GPUs.  Here the initial value
of coordinate ``j`` of parameter ``id`` is a pure function ``f(seed, stream, id, j)``:
Philox4x32-10 with key = seed (64 bit), counter = (id_lo, id_hi, j // 4, stream); word ``j % 4``
of the output block becomes ``u = (word >> 8) * 2**-24`` in ``[0, 1)`` and the value is
``fmaf(u, hi - lo, lo)`` evaluated in fp32.  ``csrc/philox.cuh`` is the device twin.

Because initialisation is a pure function of the id, "initialise on first pull" is idempotent
on the device: any GPU may evaluate it without coordination.
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)

STREAM_ITEM = 0   # PS-side table (MF item factors, PA weights)
STREAM_USER = 1   # worker-side table (MF user factors)


def philox4x32(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32-10.  Counters are uint32 arrays (broadcastable); returns 4 uint32 arrays."""
    c0 = np.asarray(c0, dtype=np.uint64) & _MASK
    c1 = np.asarray(c1, dtype=np.uint64) & _MASK
    c2 = np.asarray(c2, dtype=np.uint64) & _MASK
    c3 = np.asarray(c3, dtype=np.uint64) & _MASK
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 &= 0xFFFFFFFF
    k1 &= 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def uniform01(ids, dim: int, seed: int, stream: int = STREAM_ITEM) -> np.ndarray:
    """``[len(ids), dim]`` fp32 uniforms in [0, 1)."""
    ids = np.atleast_1d(np.asarray(ids, dtype=np.int64)).astype(np.uint64)
    nblk = (dim + 3) // 4
    blk = np.arange(nblk, dtype=np.uint64)[None, :]
    lo = (ids & _MASK)[:, None]
    hi = (ids >> np.uint64(32))[:, None]
    w = philox4x32(lo, hi, blk, np.uint64(stream), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    words = np.stack(w, axis=-1).reshape(len(ids), nblk * 4)[:, :dim]
    return ((words >> np.uint32(8)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)


def ranged_uniform(ids, dim: int, lo: float, hi: float, seed: int, stream: int = STREAM_ITEM) -> np.ndarray:
    """``fmaf(u, hi - lo, lo)`` in fp32 - identical rounding to the device kernel."""
    u = uniform01(ids, dim, seed, stream).astype(np.float64)
    scale = np.float64(np.float32(np.float32(hi) - np.float32(lo)))
    # one rounding at the end == fused multiply-add semantics
    return (u * scale + np.float64(np.float32(lo))).astype(np.float32)
