"""grtSortPairsNoAlloc (allocation-free (key, value) radix sort, temporaries in a static device array) against the existing
grtDebugRadixSort entry point, through the C ABI.  Collected last on purpose (file name).
Integer work: the comparison is bit-exact."""
import importlib

import numpy as np
import pytest

grt = importlib.import_module("gaussian-ray-tracing_b200")

pytestmark = pytest.mark.gpu


def _pairs(n, seed, key_bits=63):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << key_bits, size=n, dtype=np.uint64)
    if n > 16:
        keys[rng.integers(0, n, size=n // 8)] = keys[rng.integers(0, n, size=n // 8)]  # duplicates: the sort must be stable
    vals = np.arange(n, dtype=np.uint32)
    return keys, vals


def _sort_noalloc(ctx, d_keys, d_vals, keys, vals):
    n = keys.shape[0]
    ctx.to_device(d_keys, keys)
    ctx.to_device(d_vals, vals)
    ctx.sort_pairs_noalloc(d_keys, d_vals, n)
    ctx.sync()
    return ctx.to_host(d_keys, (n,), np.uint64).copy(), ctx.to_host(d_vals, (n,), np.uint32).copy()


def test_no_alloc_sort_matches_existing_entry_point_twice():
    ctx = grt.Context(0)
    nmax = 1 << 20
    d_keys = ctx.malloc(8 * nmax)
    d_vals = ctx.malloc(4 * nmax)
    try:
        # two calls in a row on the same static scratch array: a large one, then a smaller one that sees the first one's leftovers,
        # then the maximum size and a tiny one
        for n, seed, bits in ((777_777, 1, 63), (100_003, 2, 40), (nmax, 3, 63), (5, 4, 8)):
            keys, vals = _pairs(n, seed, bits)
            k_ref, v_ref = ctx.radix_sort(keys, vals)                 # existing entry point (allocates its temporaries)
            k_new, v_new = _sort_noalloc(ctx, d_keys, d_vals, keys, vals)
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(k_ref, keys[order]) and np.array_equal(v_ref, vals[order])
            assert np.array_equal(k_new, k_ref), f"keys differ at n={n}"
            assert np.array_equal(v_new, v_ref), f"values differ at n={n}"
        # the same input twice in a row gives the same output (the array keeps the last call's contents; nothing may leak)
        keys, vals = _pairs(300_000, 9)
        a = _sort_noalloc(ctx, d_keys, d_vals, keys, vals)
        b = _sort_noalloc(ctx, d_keys, d_vals, keys, vals)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    finally:
        ctx.free(d_keys)
        ctx.free(d_vals)
        ctx.close()


def test_no_alloc_sort_rejects_sizes_beyond_the_static_array():
    ctx = grt.Context(0)
    d_keys = ctx.malloc(64)
    d_vals = ctx.malloc(64)
    try:
        with pytest.raises(Exception) as ei:
            ctx.sort_pairs_noalloc(d_keys, d_vals, (1 << 20) + 1)
        assert "exceeds" in str(ei.value) and "1048576" in str(ei.value)
        ctx.sort_pairs_noalloc(d_keys, d_vals, 0)  # empty input is a no-op
    finally:
        ctx.free(d_keys)
        ctx.free(d_vals)
        ctx.close()
