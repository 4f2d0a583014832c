"""splat_sort_pairs_u64_u32_static: the sort entry point that allocates nothing (its scratch is a fixed device array inside the
library; no counterpart in the reference). Through the C ABI, against the
existing entry point splat_sort_pairs_u64_u32 on the same inputs: bit-exact, the bar of every sort test here (integer work).
Two calls in a row on one stream: the second finds the first one's histograms, tickets and lookback words in the shared
array and must not be disturbed by them. Collected last (file name) so that a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sort_both(ctx, keys_np, begin_bit, end_bit):
    """Returns (keys, vals) of the existing entry point and of the static one for the same input."""
    import torch
    from gaussian_splat_b200 import capi
    L = capi.lib()
    n = keys_np.size
    s = torch.cuda.current_stream().cuda_stream
    out = []
    for static in (False, True):
        k = torch.from_numpy(keys_np.view(np.int64).copy()).cuda()
        v = torch.arange(n, dtype=torch.int32, device="cuda")
        ka, va = torch.empty_like(k), torch.empty_like(v)        # the caller's buffers; the library allocates none in the static call
        torch.cuda.synchronize()
        if static:
            alt = C.c_int(-1)
            rc = L.splat_sort_pairs_u64_u32_static(C.c_void_p(s), C.c_void_p(k.data_ptr()), C.c_void_p(v.data_ptr()),
                                                   C.c_void_p(ka.data_ptr()), C.c_void_p(va.data_ptr()), n, begin_bit, end_bit,
                                                   C.byref(alt))
            assert rc == 0, L.splat_last_error_string()
            alt = alt.value
        else:
            alt = ctx.sort_pairs(s, k.data_ptr(), v.data_ptr(), ka.data_ptr(), va.data_ptr(), n, begin_bit, end_bit)
        torch.cuda.synchronize()
        rk, rv = (ka, va) if alt else (k, v)
        out.append((rk.cpu().numpy().view(np.uint64), rv.cpu().numpy().view(np.uint32)))
    return out


def test_static_sort_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gaussian_splat_b200 import capi
    rng = np.random.default_rng(21)
    ctx = capi.Context(0)
    try:
        # first call: close to the maximum, all 64 bits; second call: a smaller, ragged size with few distinct keys (long runs of
        # equal keys test stability) on a bit sub-range - it reuses the part of the array the first call left dirty
        cases = [(capi.STATIC_SORT_MAX_PAIRS - 4097, 0, 64, False), (300_001, 0, 40, True), (1, 0, 64, False)]
        for n, b0, b1, dup in cases:
            keys = rng.integers(0, 2 ** 63, n, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, n, dtype=np.uint64)
            if dup:
                keys = (keys & np.uint64(0x3FF)) << np.uint64(20)
            (ek, ev), (gk, gv) = _sort_both(ctx, keys, b0, b1)
            assert np.array_equal(gk, ek), n
            assert np.array_equal(gv, ev), n
            m = gk & np.uint64(((1 << b1) - 1) & ~((1 << b0) - 1))        # the bits that were sorted
            assert np.all(m[1:] >= m[:-1])
            assert np.array_equal(keys[gv], gk)                  # values still name their keys: a permutation of the input
    finally:
        ctx.close()


def test_static_sort_refuses_sizes_beyond_its_scratch():
    from gaussian_splat_b200 import capi
    L = capi.lib()
    alt = C.c_int(0)
    rc = L.splat_sort_pairs_u64_u32_static(None, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), C.c_void_p(8),
                                           capi.STATIC_SORT_MAX_PAIRS + 1, 0, 64, C.byref(alt))
    assert rc == -1                                              # SPLAT_ERR_ARG, before any pointer is touched
    msg = L.splat_last_error_string().decode()
    assert "SPLAT_STATIC_SORT_MAX_PAIRS" in msg and "2097152" in msg
    assert L.splat_sort_pairs_u64_u32_static(None, None, None, None, None, 0, 0, 64, C.byref(alt)) == 0      # nothing to do
    assert L.splat_sort_pairs_u64_u32_static(None, C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), C.c_void_p(8), 16, 8, 4, None) == -1
