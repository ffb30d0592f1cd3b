"""lzg_distance_matrix_noalloc (no counterpart in the reference; added for callers that cannot call cudaMalloc): through the
C ABI, against lzg_distance_matrix_device_async / lzg_distance_matrix on the same profiles, bit-exact, on calls in a row that
share the library's static scratch array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from lazac_copy_number_b200 import synth
from lazac_copy_number_b200._capi import LazacGpuError
from lazac_copy_number_b200.api import Profiles

pytestmark = pytest.mark.gpu


def _both(ctx, bp):
    n = bp.shape[0]
    prof = Profiles(ctx, bp)
    want_host = prof.distance_matrix()
    d_ref = torch.full((n, n), -1, dtype=torch.int32, device="cuda:0")
    d_new = torch.full((n, n), -2, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    prof.distance_device_async(d_ref.data_ptr())
    prof.distance_noalloc(d_new.data_ptr())
    ctx.sync()
    got_ref, got_new = d_ref.cpu().numpy(), d_new.cpu().numpy()
    prof.free()
    assert np.array_equal(got_ref, want_host)
    assert np.array_equal(got_new, got_ref)
    return got_new


def test_noalloc_equals_the_allocating_entry_point_on_calls_in_a_row(ctx, port):
    # a large plan (split tiles, 148 CTAs), then a small one that leaves most of the last plan in the array, then the
    # single-tile 64 x 64 form, then the large one again; int8 and the 16-bit format
    big = synth.random_bp(1500, 900, seed=11, lo=-9, hi=9, density=0.3)
    small = synth.random_bp(200, 300, seed=12, lo=-5, hi=5, density=0.5)
    tiny = synth.random_bp(40, 77, seed=13)
    wide = synth.random_bp(300, 500, seed=14, lo=-3000, hi=3000, density=0.2)
    first = _both(ctx, big)
    got_small = _both(ctx, small)
    assert np.array_equal(got_small, port.distance_matrix(small))
    _both(ctx, tiny)
    _both(ctx, wide)
    again = _both(ctx, big.astype(np.int8))
    assert np.array_equal(first, again)


def test_noalloc_same_buffer_twice(ctx):
    """Two calls in a row into the same output: tiles accumulated with atomics must be zeroed by the call itself."""
    bp = synth.random_bp(700, 4000, seed=21, lo=-9, hi=9, density=0.3)
    prof = Profiles(ctx, bp)
    want = prof.distance_matrix()
    d = torch.empty((700, 700), dtype=torch.int32, device="cuda:0")
    for _ in range(2):
        prof.distance_noalloc(d.data_ptr())
        ctx.sync()
        assert np.array_equal(d.cpu().numpy(), want)
    prof.free()


def test_noalloc_rejects_a_plan_beyond_the_static_array(ctx):
    lib = ctx.lib
    cap = lib.lzg_distance_matrix_noalloc_max_bytes()
    assert cap == 1 << 20
    n, L = 48000, 16
    info = (C.c_int32 * 4)()
    lib.lzg_distance_plan(n, L, 148, 0, 1, 0, 2 ** 31 - 1, None, 0, None, 0, info)  # sizes only
    assert 16 * info[1] > cap  # the segments alone do not fit
    prof = Profiles(ctx, np.zeros((n, L), np.int8))
    d = torch.zeros(16, dtype=torch.int32, device="cuda:0")  # never written: the call is refused before any launch
    with pytest.raises(LazacGpuError, match="noalloc.*scratch"):
        prof.distance_noalloc(d.data_ptr())
    prof.free()
    # the array still serves the next call
    bp = synth.random_bp(130, 200, seed=31)
    _both(ctx, bp)
