"""The hand-written radix sort and prefix sums (csrc/sort.cu) against numpy, and against CUB as a second opinion
(PCP_SORT_CUB=1, a subprocess: the switch is read once per process).  Bit-exact: a stable sort by the low bits."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _expect(keys, bits):
    kb = 8 * keys.dtype.itemsize
    masked = keys if bits >= kb else (keys & keys.dtype.type((1 << bits) - 1))
    order = np.argsort(masked, kind="stable")
    return keys[order], order.astype(np.uint32)


def _cases():
    rng = np.random.default_rng(3)
    yield "random u64 full", rng.integers(0, 1 << 63, 300_001, dtype=np.uint64) * np.uint64(2) + np.uint64(1), 64
    yield "random u64 42 bits", rng.integers(0, 1 << 42, 1_000_003, dtype=np.uint64), 42
    yield "random u32 full", rng.integers(0, 1 << 32, 500_000, dtype=np.uint32), 32
    yield "random u32 19 bits", rng.integers(0, 1 << 19, 77_777, dtype=np.uint32), 19
    yield "few distinct", rng.integers(0, 3, 200_000, dtype=np.uint64) << np.uint64(17), 40
    yield "all equal", np.full(100_000, 0x1234567, dtype=np.uint64), 33
    yield "sorted", np.arange(250_000, dtype=np.uint64) * np.uint64(7), 30
    yield "reversed", (np.arange(250_000, dtype=np.uint64) * np.uint64(7))[::-1].copy(), 30
    yield "constant high bits", (np.uint64(0xABC) << np.uint64(40)) | rng.integers(0, 1 << 20, 400_000, dtype=np.uint64), 52
    yield "dropped keys last", np.concatenate([rng.integers(0, 1 << 30, 50_000, dtype=np.uint64),
                                               np.full(777, 0xFFFFFFFFFFFFFFFF, dtype=np.uint64)]), 64
    yield "one", np.array([5], dtype=np.uint64), 8
    yield "tile edge", rng.integers(0, 1 << 24, 4096, dtype=np.uint32), 24
    yield "tile edge + 1", rng.integers(0, 1 << 24, 4097, dtype=np.uint32), 24
    yield "low bits only", rng.integers(0, 1 << 32, 123_456, dtype=np.uint32), 5


@pytest.mark.parametrize("name,keys,bits", list(_cases()), ids=[c[0] for c in _cases()])
def test_sort_pairs_matches_numpy(ctx, name, keys, bits):
    ek, ev = _expect(keys, bits)
    k = keys.copy()
    v = np.arange(len(keys), dtype=np.uint32)
    ctx.sort_pairs(k, v, bits)
    assert np.array_equal(v, ev), name  # stable: the permutation itself is pinned
    assert np.array_equal(k, ek), name
    k2 = keys.copy()
    ctx.sort_pairs(k2, None, bits)       # keys only (iota values generated and dropped)
    assert np.array_equal(k2, ek), name
    v3 = (np.arange(len(keys), dtype=np.uint32) * np.uint32(2654435761)) ^ np.uint32(0x5bd1e995)
    k3 = keys.copy()
    e3 = v3[ev]
    ctx.sort_pairs(k3, v3, bits)         # arbitrary values travel with their keys
    assert np.array_equal(v3, e3), name


def test_sort_large_voxel_like_keys(ctx):
    """25 M 42-bit cell keys of a city-like cloud: nearly sorted runs, constant top bits, 6 passes of 7 bits."""
    rng = np.random.default_rng(11)
    n = 25_000_000
    i = rng.integers(0, 30000, n, dtype=np.uint64)
    j = rng.integers(0, 30000, n, dtype=np.uint64)
    k = (rng.standard_normal(n) * 3 + 20).clip(0, 2999).astype(np.uint64)
    keys = i + j * np.uint64(30000) + k * np.uint64(30000 * 30000)
    ek, ev = _expect(keys, 42)
    v = np.arange(n, dtype=np.uint32)
    ctx.sort_pairs(keys, v, 42)
    assert np.array_equal(keys, ek) and np.array_equal(v, ev)


@pytest.mark.parametrize("n", [1, 31, 4096, 4097, 65_537, 3_000_001])
def test_exclusive_sum(ctx, n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 1000, n, dtype=np.uint32)
    got = ctx.exclusive_sum(a)
    exp = np.concatenate([[0], np.cumsum(a[:-1], dtype=np.uint64)]).astype(np.uint32)
    assert np.array_equal(got, exp)


def test_cub_switch_gives_the_same_order():
    """Second opinion: the same sort through CUB (library code, kept only behind this switch)."""
    code = ("import numpy as np, pointcloudprocess_b200 as pcp\n"
            "ctx = pcp.Context(0)\n"
            "rng = np.random.default_rng(5)\n"
            "k = rng.integers(0, 1 << 41, 2_000_003, dtype=np.uint64); v = np.arange(len(k), dtype=np.uint32)\n"
            "ctx.sort_pairs(k, v, 41)\n"
            "import hashlib; print(hashlib.sha1(k.tobytes() + v.tobytes()).hexdigest())\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for env in ({}, {"PCP_SORT_CUB": "1"}):
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root, env=e, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs.append(r.stdout.strip().splitlines()[-1])
    assert outs[0] == outs[1]
