"""dlti_sumsq_noalloc, called straight through the C ABI, against dlti_sumsq (GPU tier).

The no-alloc entry point exists for hosts that cannot give the library a device workspace: a
server whose KV-cache pool already owns all free HBM, a loop that must avoid the device-wide
synchronisation of cudaMalloc/cudaFree, or a bare C caller with no allocator to carve scratch
from. It keeps its partial sums in an array compiled into the library, so the checks here are the
ones that design can get wrong: the answer must match the caller-scratch version, a second call
must not be polluted by what the first left in the shared array (a long vector followed by a short
one), accumulation must work, and an input beyond the array's capacity must be refused with a
readable message instead of overrunning it.

This file sorts last on purpose: with `-x`, every other test has already run if this one fails.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

ERR_TOO_LARGE = -4


def _lib():
    from dlti_b200.ops import _lib as L

    lib = C.CDLL(str(L._GPU_LIB_PATH))
    lib.dlti_sumsq.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.dlti_sumsq.restype = C.c_int
    lib.dlti_sumsq_noalloc.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_void_p]
    lib.dlti_sumsq_noalloc.restype = C.c_int
    lib.dlti_sumsq_noalloc_max_elems.restype = C.c_longlong
    lib.dlti_sumsq_scratch_floats.restype = C.c_int
    lib.dlti_last_error.restype = C.c_char_p
    return lib


def test_noalloc_sumsq_matches_scratch_version_twice_in_a_row():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    scratch = torch.empty(lib.dlti_sumsq_scratch_floats(), dtype=torch.float32, device=dev)
    gen = torch.Generator(device=dev).manual_seed(0)
    # two calls back to back: a long vector, then a short unaligned-length one
    for n in (5_000_003, 1_000):
        g = torch.randn(n, device=dev, generator=gen) * 3.0
        ref = torch.zeros(1, device=dev)
        out = torch.full((1,), -7.0, device=dev)
        assert lib.dlti_sumsq(g.data_ptr(), n, scratch.data_ptr(), ref.data_ptr(), 0, st) == 0
        assert lib.dlti_sumsq_noalloc(g.data_ptr(), n, out.data_ptr(), 0, st) == 0
        torch.cuda.synchronize()
        torch.testing.assert_close(out, ref, rtol=1e-4, atol=0)
        torch.testing.assert_close(out.cpu(), g.double().pow(2).sum().float().reshape(1).cpu(), rtol=1e-4, atol=0)
    # accumulate into an existing value, from a view that is not 16-byte aligned
    base = torch.randn(4097, device=dev, generator=gen)
    g = base[1:]
    acc = torch.full((1,), 10.0, device=dev)
    assert lib.dlti_sumsq_noalloc(g.data_ptr(), g.numel(), acc.data_ptr(), 1, st) == 0
    torch.cuda.synchronize()
    torch.testing.assert_close(acc.cpu(), (10.0 + g.double().pow(2).sum()).float().reshape(1).cpu(), rtol=1e-4, atol=0)
    # a non-finite element must surface (this is how fp16 overflow is detected)
    g2 = torch.ones(70_000, device=dev)
    g2[65_432] = float("inf")
    o2 = torch.zeros(1, device=dev)
    assert lib.dlti_sumsq_noalloc(g2.data_ptr(), g2.numel(), o2.data_ptr(), 0, st) == 0
    torch.cuda.synchronize()
    assert not torch.isfinite(o2).item()


def test_noalloc_sumsq_rejects_oversize_input():
    lib = _lib()
    torch.cuda.set_device(0)
    cap = lib.dlti_sumsq_noalloc_max_elems()
    assert cap == 16384 * 8192
    small = torch.zeros(8, device="cuda")
    out = torch.zeros(1, device="cuda")
    rc = lib.dlti_sumsq_noalloc(small.data_ptr(), cap + 1, out.data_ptr(), 0, None)   # refused before any launch
    assert rc == ERR_TOO_LARGE
    msg = lib.dlti_last_error().decode()
    assert "exceeds" in msg and str(cap) in msg
    torch.cuda.synchronize()
