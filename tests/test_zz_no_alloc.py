"""hb_kmeans_step_static: the K-means pass for callers that cannot supply scratch memory.

Some users of the C ABI have no way to allocate device memory for this library: a JVM or
other foreign runtime calling in through FFI owns no CUDA allocator, and a process whose GPU
memory was reserved wholesale by another pool cannot afford a cudaMalloc (it may fail, and it
synchronises the device).  For them the library carries its own fixed scratch array.  This
test drives that entry point exactly as such a caller would — through ctypes, raw pointers
only — and checks it against hb_kmeans_step (caller-provided workspace) twice in a row, the
second time with a smaller problem so that leftovers of the first call sit in the shared
array.  It lives in a file that sorts last so a `-x` run has already exercised everything else.
"""
import ctypes
from ctypes import c_int, c_longlong, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from harp_b200 import _native
    lib = _native.lib()
    lib.hb_kmeans_step_static.argtypes = [c_void_p, c_longlong, c_int, c_void_p, c_int, c_void_p,
                                          c_void_p, c_void_p, c_void_p, c_void_p]
    lib.hb_kmeans_step_static.restype = c_int
    return lib, _native


def _run_both(lib, nat, n, d, k, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    X = torch.rand(n, d, device="cuda", generator=g)
    C = torch.rand(k, d, device="cuda", generator=g)
    stream = nat.stream_ptr()
    outs = []
    for static in (False, True):
        sums = torch.full((k, d), -1.0, device="cuda")
        counts = torch.full((k,), -1.0, device="cuda")
        assign = torch.full((n,), -1, dtype=torch.int32, device="cuda")
        obj = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
        if static:
            rc = lib.hb_kmeans_step_static(nat.ptr(X), n, d, nat.ptr(C), k, nat.ptr(sums), nat.ptr(counts),
                                           nat.ptr(assign), nat.ptr(obj), stream)
        else:
            # same kernel as the built-in-scratch entry runs (flag 4 = single-CTA tcgen05 kernel), so the
            # two outputs are comparable element by element
            nbytes = int(lib.hb_kmeans_workspace_bytes(k, d))
            ws = torch.empty((nbytes + 255) // 256 * 256, dtype=torch.uint8, device="cuda")
            rc = lib.hb_kmeans_step(nat.ptr(X), n, d, nat.ptr(C), k, nat.ptr(sums), nat.ptr(counts),
                                    nat.ptr(assign), nat.ptr(obj), nat.ptr(ws), ws.numel(), 4, stream)
        assert rc == 0, lib.hb_last_error().decode()
        torch.cuda.synchronize()
        outs.append((sums, counts, assign, obj))
    (s0, c0, a0, o0), (s1, c1, a1, o1) = outs
    assert torch.equal(a0, a1)
    assert torch.equal(c0, c1)
    assert float(c1.sum()) == n
    torch.testing.assert_close(s1, s0, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(o1, o0, rtol=1e-6, atol=1e-6)


def test_static_scratch_matches_workspace_entry_point_twice(native_built):
    lib, nat = _lib()
    _run_both(lib, nat, n=20000, d=128, k=1000, seed=11)   # first call: resolves the scratch address
    _run_both(lib, nat, n=5000, d=64, k=300, seed=12)      # second call: smaller, stale tail in the array


def test_static_scratch_rejects_what_it_cannot_hold(native_built):
    lib, nat = _lib()
    X = torch.rand(256, 128, device="cuda")
    kmax = 16384
    C = torch.rand(kmax + 1, 128, device="cuda")
    sums = torch.zeros(kmax + 1, 128, device="cuda"); counts = torch.zeros(kmax + 1, device="cuda")
    rc = lib.hb_kmeans_step_static(nat.ptr(X), 256, 128, nat.ptr(C), kmax + 1, nat.ptr(sums), nat.ptr(counts),
                                   None, None, nat.stream_ptr())
    assert rc != 0 and b"16384" in lib.hb_last_error()
    X2 = torch.rand(256, 130, device="cuda"); C2 = torch.rand(8, 130, device="cuda")
    rc = lib.hb_kmeans_step_static(nat.ptr(X2), 256, 130, nat.ptr(C2), 8, nat.ptr(sums), nat.ptr(counts),
                                   None, None, nat.stream_ptr())
    assert rc != 0 and b"hb_kmeans_step" in lib.hb_last_error()
    torch.cuda.synchronize()


def test_static_scratch_alignment_and_largest_k(native_built):
    """A misaligned input is refused with a message about alignment (not about a workspace this caller never gave),
    before anything is queued; k = 16384, the documented limit, still runs."""
    lib, nat = _lib()
    n, d, k = 4096, 64, 16384
    X = torch.rand(n, d, device="cuda"); C = torch.rand(k, d, device="cuda")
    sums = torch.full((k, d), -1.0, device="cuda"); counts = torch.full((k,), -1.0, device="cuda")
    flat = torch.rand(n * d + 1, device="cuda")
    Xmis = flat[1:]                                         # 4 bytes past a 16-byte boundary
    rc = lib.hb_kmeans_step_static(c_void_p(Xmis.data_ptr()), n, d, nat.ptr(C), k, nat.ptr(sums), nat.ptr(counts),
                                   None, None, nat.stream_ptr())
    assert rc != 0 and b"16-byte aligned" in lib.hb_last_error()
    torch.cuda.synchronize()
    assert float(sums.min()) == -1.0 and float(counts.min()) == -1.0        # nothing was touched
    rc = lib.hb_kmeans_step_static(nat.ptr(X), n, d, nat.ptr(C), k, nat.ptr(sums), nat.ptr(counts), None, None,
                                   nat.stream_ptr())
    assert rc == 0, lib.hb_last_error().decode()
    torch.cuda.synchronize()
    assert float(counts.sum()) == n
