"""``fps_lemp_topk_static`` - top-K retrieval for callers that are not allowed to allocate.

Some users of this library answer recommendation queries from a process in which every byte of
device memory already belongs to someone else (a trainer, another framework's caching allocator),
or from a request path that cannot tolerate the device-wide stall of ``cudaMalloc`` / ``cudaFree``.
For them the library carries its own fixed scratch array; this test drives that entry point
through the C ABI exactly as such a caller would (inputs and outputs only, no scratch argument)
and checks it against the regular entry point, twice in a row with different data, so that
anything the first call leaves behind in the shared array would show up in the second.

The file name sorts last on purpose: a failure here cannot hide other tests behind ``-x``.
"""
import ctypes as C

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

STATIC_BYTES = 16 << 20
FPS_E_TOO_LARGE = 10002


def _lib():
    """The shared ctypes binding (``ops/_native.py`` declares the entry point's argument types)."""
    from fps_b200.ops import _native as N
    L = N.lib()
    assert L.fps_lemp_topk_static.argtypes is not None
    return N, L


def _device_free_bytes() -> int:
    """Free device memory as the driver sees it (``cudaMemGetInfo``).  torch's own counters only see
    torch's caching allocator: a ``cudaMalloc`` inside the library would not move them."""
    rt = C.CDLL("libcudart.so.12")
    free, total = C.c_size_t(0), C.c_size_t(0)
    rc = rt.cudaMemGetInfo(C.byref(free), C.byref(total))
    assert rc == 0, rc
    return free.value


def _problem(seed, n_q, n_items, dim):
    g = torch.Generator(device="cuda").manual_seed(seed)
    q = torch.randn(n_q, dim, device="cuda", generator=g).contiguous()
    items = torch.randn(n_items, dim, device="cuda", generator=g)
    items = items * (0.2 + 2.0 * torch.rand(n_items, 1, device="cuda", generator=g))
    lens = torch.linalg.vector_norm(items, dim=1)
    order = torch.argsort(lens, descending=True)
    return q, items[order].contiguous(), lens[order].contiguous(), order.int().contiguous()


def _call_regular(N, L, q, items, lens, ids, k):
    n_q, dim = q.shape
    need = int(L.fps_topk_scratch_bytes(n_q, k, items.shape[0]))
    scratch = torch.empty(need, device="cuda", dtype=torch.uint8)
    out_s = torch.empty((n_q, k), device="cuda")
    out_i = torch.empty((n_q, k), device="cuda", dtype=torch.int32)
    N.check(L.fps_lemp_topk(q.data_ptr(), dim, n_q, items.data_ptr(), dim, lens.data_ptr(), ids.data_ptr(),
                            items.shape[0], dim, k, 1, scratch.data_ptr(), need, out_s.data_ptr(),
                            out_i.data_ptr(), None, N.current_stream()), "fps_lemp_topk")
    return out_s, out_i


def _call_static(N, L, q, items, lens, ids, k, out_s, out_i):
    n_q, dim = q.shape
    return L.fps_lemp_topk_static(q.data_ptr(), dim, n_q, items.data_ptr(), dim, lens.data_ptr(), ids.data_ptr(),
                                  items.shape[0], dim, k, 1, out_s.data_ptr(), out_i.data_ptr(), None,
                                  N.current_stream())


def test_static_scratch_entry_matches_regular_entry_twice():
    N, L = _lib()
    # outputs are allocated up front: between the two static calls nothing is allocated at all
    shapes = [(1, 200, 30_000, 64, 50), (2, 97, 8_000, 128, 100)]
    outs = [(torch.empty((n_q, k), device="cuda"), torch.empty((n_q, k), device="cuda", dtype=torch.int32))
            for _seed, n_q, _n, _d, k in shapes]
    problems = [_problem(seed, n_q, n_items, dim) for seed, n_q, n_items, dim, _k in shapes]
    # a first call per device does the one-time address look-up; measure around the calls after it
    rc = _call_static(N, L, *problems[0], shapes[0][4], *outs[0])
    assert rc == 0, N.lib().fps_last_error()
    torch.cuda.synchronize()
    before, before_torch = _device_free_bytes(), torch.cuda.memory_allocated()
    for (seed, n_q, n_items, dim, k), (q, items, lens, ids), (out_s, out_i) in zip(shapes, problems, outs):
        rc = _call_static(N, L, q, items, lens, ids, k, out_s, out_i)
        assert rc == 0, N.lib().fps_last_error()
    torch.cuda.synchronize()
    # neither the driver's free-memory figure nor torch's allocator moved: nothing was allocated or freed
    assert _device_free_bytes() == before
    assert torch.cuda.memory_allocated() == before_torch
    for (seed, n_q, n_items, dim, k), (q, items, lens, ids), (out_s, out_i) in zip(shapes, problems, outs):
        ref_s, ref_i = _call_regular(N, L, q, items, lens, ids, k)
        torch.cuda.synchronize()
        torch.testing.assert_close(out_s, ref_s, rtol=2e-5, atol=2e-5)
        # same items wherever scores are not tied to within the tolerance
        same = (out_i == ref_i).float().mean().item()
        assert same > 0.99, same
        # and the scores are right in absolute terms (fp32 reference of the same op)
        truth = (q.double() @ items.double().T).topk(k, dim=1).values
        torch.testing.assert_close(out_s.double(), truth, rtol=2e-5, atol=2e-5)


def test_static_scratch_entry_refuses_oversized_problems():
    N, L = _lib()
    q, items, lens, ids = _problem(3, 2000, 4_000, 32)          # 16 query tiles: far beyond 16 MiB
    assert int(L.fps_topk_scratch_bytes(2000, 100, 4_000)) > STATIC_BYTES
    out_s = torch.empty((2000, 100), device="cuda")
    out_i = torch.empty((2000, 100), device="cuda", dtype=torch.int32)
    rc = _call_static(N, L, q, items, lens, ids, 100, out_s, out_i)
    assert rc == FPS_E_TOO_LARGE
    msg = L.fps_last_error().decode()
    assert "scratch" in msg and "split the query batch" in msg
