"""pcp_exclusive_sum_u32_noalloc (include/pcp_cuda.h): the prefix sum whose scratch is a fixed array in device
memory, for callers that cannot allocate.  Checked through the C ABI against the existing entry point
(pcp_exclusive_sum_u32) -- bit-exact, integer work -- on two calls in a row that share the scratch array.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _excl(a):
    """Exclusive prefix sums mod 2^32, as the device computes them in uint32."""
    out = np.zeros(a.shape[0], dtype=np.uint64)
    np.cumsum(a[:-1], dtype=np.uint64, out=out[1:])
    return (out & np.uint64(0xFFFFFFFF)).astype(np.uint32)


def _run_pair(ctx, a):
    """-> (existing entry point's result, no-alloc entry point's result) for the uint32 array a, both on device buffers."""
    L = ctx.L
    n = a.shape[0]
    d_in, d_ref, d_out = ctx.alloc(n * 4), ctx.alloc(n * 4), ctx.alloc(n * 4)   # the CALLER's buffers
    try:
        d_in.from_host(a)
        ctx._check(L.pcp_exclusive_sum_u32(ctx.h, d_in.ptr, d_ref.ptr, n))
        ctx._check(L.pcp_exclusive_sum_u32_noalloc(ctx.h, d_in.ptr, d_out.ptr, n))
        ctx.synchronize()
        return d_ref.to_host(np.uint32, n), d_out.to_host(np.uint32, n)
    finally:
        for d in (d_in, d_ref, d_out):
            d.free()


def test_noalloc_scan_matches_existing_entry_point_twice(ctx):
    rng = np.random.default_rng(17)
    # first call: many tiles (fills a long stretch of the shared scratch); second call: fewer, other values, so
    # anything the first call left behind would show
    for n in (5_000_003, 1_234_567, 4096, 1):
        a = rng.integers(0, 1 << 12, n, dtype=np.uint32)
        ref, got = _run_pair(ctx, a)
        assert np.array_equal(ref, _excl(a))
        assert np.array_equal(got, ref), n


def test_noalloc_scan_in_place_and_back_to_back(ctx):
    """Two calls queued on the stream with no host synchronisation between them, the second in place."""
    L = ctx.L
    rng = np.random.default_rng(23)
    n = 2_000_001
    a = rng.integers(0, 7, n, dtype=np.uint32)
    d0, d1 = ctx.alloc(n * 4), ctx.alloc(n * 4)
    try:
        d0.from_host(a)
        ctx._check(L.pcp_exclusive_sum_u32_noalloc(ctx.h, d0.ptr, d1.ptr, n))
        ctx._check(L.pcp_exclusive_sum_u32_noalloc(ctx.h, d1.ptr, d1.ptr, n))  # scan of the scan, out == in
        ctx.synchronize()
        got = d1.to_host(np.uint32, n)
    finally:
        d0.free()
        d1.free()
    assert np.array_equal(got, _excl(_excl(a)))  # wraps mod 2^32 like the device


def test_noalloc_scan_rejects_what_it_cannot_do(ctx):
    L = ctx.L
    d = ctx.alloc(64)
    try:
        # beyond the fixed scratch: refused before anything is touched
        st = L.pcp_exclusive_sum_u32_noalloc(ctx.h, d.ptr, d.ptr, 16384 * 4096 + 1)
        assert st == -4 and b"PCP_NOALLOC_SCAN_MAX_N" in L.pcp_last_error(ctx.h)
        # host arrays would need a staging allocation
        h = np.zeros(16, dtype=np.uint32)
        st = L.pcp_exclusive_sum_u32_noalloc(ctx.h, h.ctypes.data, d.ptr, 16)
        assert st == -1 and b"device pointers" in L.pcp_last_error(ctx.h)
        assert L.pcp_exclusive_sum_u32_noalloc(ctx.h, None, None, 0) == 0
    finally:
        d.free()


def test_noalloc_scan_at_its_maximum(ctx):
    """Exactly PCP_NOALLOC_SCAN_MAX_N items: every word of the fixed scratch is used."""
    n = 16384 * 4096
    a = (np.arange(n, dtype=np.uint32) * np.uint32(2654435761)) >> np.uint32(27)
    ref, got = _run_pair(ctx, a)
    assert np.array_equal(got, ref)
    assert int(got[-1]) == int(np.sum(a[:-1], dtype=np.uint64) & 0xFFFFFFFF)
