"""smmc_reduce_mean_device_noalloc against smmc_reduce_mean_device, through the C ABI.

The no-alloc entry point exists for processes in which libsmmc must not allocate device memory: a host whose own arena
or caching allocator already owns the whole GPU (a cudaMalloc from a plugin then fails), a host that has work in flight on
other streams (cudaMalloc / cudaFree would stall all of it), or a deployment whose memory budget is fixed when the library
is loaded.  Its scratch is a fixed 262 400-byte __device__ array inside the library, shared by every call on a device and
never cleared between calls except for the words that must start at zero -- which is why each case below is run twice in a
row, with a different vector in between, and must return the same answer both times.

This file is named to be collected LAST: under `pytest -x` a failure here must not hide the parity tests.  It makes no
assumption about what ran before it: it passes alone in a fresh process (then its first call is also the per-device set-up
of the entry point) and after the rest of the suite (then the library's allocating context already exists beside it).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-6  # the tolerance tests/test_gpu_parity.py::test_reduce_mean_and_vector_add allows on a mean


def _both(lib, ptr, n, stream=None):
    """(no-alloc mean, allocating entry point's mean) of the n floats at device address ptr, straight through the C ABI"""
    from stock_market_monte_carlo_b200 import _native as N
    a, b, ms = C.c_float(), C.c_float(), C.c_float()
    N.check(lib.smmc_reduce_mean_device_noalloc(C.c_void_p(ptr), C.c_int64(n), 0, stream, C.byref(a)))
    N.check(lib.smmc_reduce_mean_device(C.c_void_p(ptr), C.c_int64(n), 0, stream, C.byref(b), C.byref(ms)))
    return float(a.value), float(b.value)


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(S):
    import torch
    lib = S.load()
    rng = np.random.default_rng(11)
    h = (rng.standard_normal(19_400_003) * 1000.0 + 5000.0).astype(np.float32)
    d = torch.from_numpy(h).cuda()
    other = torch.full((3_000_000,), -7.25, dtype=torch.float32, device="cuda")  # leaves different partials behind
    # (offset in floats, n): every head length 0..3, fewer floats than one 16-byte quad, exactly one 32768-float chunk,
    # one float more, ragged ends, many chunks per CTA
    cases = ((0, h.size), (1, h.size - 1), (2, 1), (3, 2), (1, 3), (0, 4), (0, 32768), (0, 32769), (3, 32768 + 5),
             (2, 4_000_001), (0, 148 * 4 * 32768 + 17))
    for off, n in cases:
        ptr = d.data_ptr() + 4 * off
        want = h[off:off + n].astype(np.float64).mean()
        first, ref = _both(lib, ptr, n)
        assert abs(first - ref) <= RTOL * abs(ref), (off, n, first, ref)
        assert abs(first - want) <= RTOL * abs(want), (off, n, first, want)
        # a different, shorter vector in between: the shared scratch now holds ITS partials, ticket and sum
        mid, _ = _both(lib, other.data_ptr(), other.numel())
        assert mid == -7.25
        second, ref2 = _both(lib, ptr, n)
        assert second == first, (off, n, first, second)  # deterministic bits, nothing inherited from the call before
        assert ref2 == ref
    # on the caller's stream, input larger than L2
    big = torch.full((1 << 28,), 1.5, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        a, b = _both(lib, big.data_ptr(), big.numel(), stream=C.c_void_p(st))
        assert a == 1.5 and b == 1.5
    # the python wrapper is the same call
    assert S.reduce_mean_device_noalloc(d.data_ptr(), h.size) == _both(lib, d.data_ptr(), h.size)[0]


def test_noalloc_rejects_what_its_fixed_scratch_cannot_hold(S):
    import torch
    from stock_market_monte_carlo_b200 import _native as N
    lib = S.load()
    d = torch.zeros(16, dtype=torch.float32, device="cuda")
    out = C.c_float(123.0)
    # the size check comes before anything touches the device: the pointer is never dereferenced for an oversized n
    rc = lib.smmc_reduce_mean_device_noalloc(C.c_void_p(d.data_ptr()), C.c_int64(N.SMMC_NOALLOC_MAX_N + 1), 0, None, C.byref(out))
    assert rc == 1 and out.value == 123.0
    msg = lib.smmc_last_error().decode()
    assert "exceeds the fixed scratch" in msg and str(N.SMMC_NOALLOC_MAX_N) in msg
    assert lib.smmc_reduce_mean_device_noalloc(C.c_void_p(d.data_ptr()), C.c_int64(0), 0, None, C.byref(out)) == 1
    assert lib.smmc_reduce_mean_device_noalloc(None, C.c_int64(4), 0, None, C.byref(out)) == 1
    assert lib.smmc_reduce_mean_device_noalloc(C.c_void_p(d.data_ptr() + 2), C.c_int64(4), 0, None, C.byref(out)) == 1
    assert lib.smmc_reduce_mean_device_noalloc(C.c_void_p(d.data_ptr()), C.c_int64(4), 99, None, C.byref(out)) == 1
    # and a valid call still works afterwards
    assert lib.smmc_reduce_mean_device_noalloc(C.c_void_p(d.data_ptr()), C.c_int64(16), 0, None, C.byref(out)) == 0
    assert out.value == 0.0
