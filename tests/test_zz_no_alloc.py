"""rd_sort_pairs_noalloc (include/radiance_b200.h): the build's pair sort with its scratch in a file-scope static device
array, for callers that cannot allocate. No counterpart in the reference, so
the checker is the existing entry point rd_debug_sort_pairs (itself pinned to numpy in tests/test_gpu_parity.py).
Integer work: the bar is bit-exact. Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest


def _sort_through_static_scratch(rd, plt, keys, vals, bits):
    n = keys.size
    kin, vin = rd.CreateBuffer(plt, max(8 * n, 8)), rd.CreateBuffer(plt, max(4 * n, 4))
    kout, vout = rd.CreateBuffer(plt, max(8 * n, 8)), rd.CreateBuffer(plt, max(4 * n, 4))
    try:
        if n:
            rd.WriteBuffer(plt, kin, 8 * n, keys)
            rd.WriteBuffer(plt, vin, 4 * n, vals)
            # poison the outputs so a pass that never ran cannot look sorted
            rd.WriteBuffer(plt, kout, 8 * n, np.full(n, 0xDEADBEEFDEADBEEF, np.uint64))
            rd.WriteBuffer(plt, vout, 4 * n, np.full(n, 0xFFFFFFFF, np.uint32))
        rd.sort_pairs_noalloc(plt, kin, vin, kout, vout, n, bits)
        k, v = np.empty(n, np.uint64), np.empty(n, np.uint32)
        kin_after = np.empty(n, np.uint64)
        if n:
            rd.ReadBuffer(plt, kout, 8 * n, k)      # blocking read on the platform stream: waits for the sort
            rd.ReadBuffer(plt, vout, 4 * n, v)
            rd.ReadBuffer(plt, kin, 8 * n, kin_after)
            assert np.array_equal(kin_after, keys), "inputs must be left unchanged"
        return k, v
    finally:
        for h in (kin, vin, kout, vout):
            h.free()


@pytest.mark.gpu
def test_noalloc_sort_matches_existing_entry_point_twice(pkg, plt):
    rd = pkg.rd
    rng = np.random.default_rng(20261105)
    # two calls in a row on the shared array: a large input first, then a smaller one with other key widths, so stale
    # tables / ping-pong contents of call 1 would show in call 2; then the maximum size and the small sizes
    cases = [(200_003, 64), (50_021, 30), (rd.SORT_NOALLOC_MAX_PAIRS, 40), (4097, 8), (1, 64), (2, 17), (0, 64)]
    for n, bits in cases:
        keys = rng.integers(0, 1 << 64, n, dtype=np.uint64)
        if n > 10:
            keys[: n // 3] = keys[n // 3: 2 * (n // 3)]          # duplicates: stability is visible in the values
        vals = np.arange(n, dtype=np.uint32)
        exp_k, exp_v = rd.debug_sort_pairs(plt, keys, vals, bits)   # the existing entry point
        got_k, got_v = _sort_through_static_scratch(rd, plt, keys, vals, bits)
        assert np.array_equal(got_k, exp_k), "keys differ at n=%d bits=%d" % (n, bits)
        assert np.array_equal(got_v, exp_v), "values differ at n=%d bits=%d" % (n, bits)
        # and against numpy directly: stable sort by the bits the passes cover
        covered = 8 * ((bits + 7) // 8)
        mask = np.uint64((1 << covered) - 1) if covered < 64 else np.uint64(0xFFFFFFFFFFFFFFFF)
        order = np.argsort(keys & mask, kind="stable")
        assert np.array_equal(got_v, vals[order])
    # same input twice in a row: identical output
    keys = rng.integers(0, 1 << 62, 123_457, dtype=np.uint64)
    vals = np.arange(keys.size, dtype=np.uint32)
    a = _sort_through_static_scratch(rd, plt, keys, vals, 64)
    b = _sort_through_static_scratch(rd, plt, keys, vals, 64)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.all(a[0][1:] >= a[0][:-1])


@pytest.mark.gpu
def test_noalloc_sort_rejects_what_does_not_fit(pkg, plt):
    rd = pkg.rd
    L = rd.lib()
    buf = rd.CreateBuffer(plt, 64)
    other = rd.CreateBuffer(plt, 64)
    try:
        d, o = buf.device_ptr(), other.device_ptr()
        rc = L.rd_sort_pairs_noalloc(plt.handle, d, d, o, o, rd.SORT_NOALLOC_MAX_PAIRS + 1, 64)
        assert rc == -2 and b"at most 262144 pairs" in L.rd_last_error()
        assert L.rd_sort_pairs_noalloc(plt.handle, d, d, d, o, 4, 64) == -2          # output aliases input
        assert L.rd_sort_pairs_noalloc(plt.handle, d, d, o, o, 4, 0) == -2           # bits out of range
        assert L.rd_sort_pairs_noalloc(plt.handle, None, None, None, None, 0, 64) == 0   # empty input: nothing to do
    finally:
        buf.free()
        other.free()


def test_noalloc_symbol_exported(pkg):
    L = pkg.rd.lib()
    assert hasattr(L, "rd_sort_pairs_noalloc")
    assert L.rd_sort_pairs_noalloc.argtypes[-1] is C.c_int
