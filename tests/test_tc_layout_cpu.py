"""CPU: the thermometer code behind the tensor-core distance (lzg_distance_tc.cu) -- the host-only column layout
(lzg_distance_tc_layout), the kernel's gather (two words + PRMT selector) and SWAR compare restated in numpy, and the identity
D = s 1^t + 1 s^t - 2 T T^t against the oracle's distance matrix.  No GPU: the library's host code only."""
import ctypes as C

import numpy as np
import pytest

from lazac_copy_number_b200 import _capi, synth


def layout(bp):
    lib = _capi.load()
    lo, hi = bp.min(0).astype(np.int32), bp.max(0).astype(np.int32)
    K = C.c_int64(0)
    n_words = lib.lzg_distance_tc_layout(lo.ctypes.data_as(C.c_void_p), hi.ctypes.data_as(C.c_void_p), bp.shape[1], None, 0, C.byref(K))
    words = np.zeros((max(1, n_words), 4), np.uint32)
    assert lib.lzg_distance_tc_layout(lo.ctypes.data_as(C.c_void_p), hi.ctypes.data_as(C.c_void_p), bp.shape[1], words.ctypes.data_as(C.c_void_p),
                                      n_words, C.byref(K)) == n_words
    return words[:n_words], int(K.value)


def expand(bp, words):
    """k_tc_expand in numpy: per word the eight bytes at its base, one byte picked per column by the selector nibble,
    compared with the threshold byte the way the kernel's SWAR expression does, masked by the valid bits."""
    n, L = bp.shape
    l_pad = (L + 127) // 128 * 128 + 16
    rows = np.full((n, l_pad), 128, np.uint8)           # the device format: value + 128, padding = the bias
    rows[:, :L] = (bp + 128).astype(np.uint8)
    T = np.zeros((n, 4 * len(words)), np.uint8)
    for w, (base, sel, thr4, mask) in enumerate(words.tolist()):
        eight = rows[:, 4 * base:4 * base + 8]
        for e in range(4):
            if not (mask >> (8 * e)) & 1:
                continue
            x = eight[:, (sel >> (4 * e)) & 0xF].astype(np.uint32)
            y = np.uint32((thr4 >> (8 * e)) & 0xFF)
            d = (x | 0x80) - (y & 0x7F)
            ge = ((x & ~y) | (~(x ^ y) & d)) >> 7 & 1    # bit 7 of the byte lane: x >= y
            T[:, 4 * w + e] = ge
    return T


@pytest.mark.parametrize("n,L,lo,hi,density,seed", [(40, 300, -3, 3, 0.05, 1), (25, 129, -128, 127, 1.0, 2), (60, 64, 0, 1, 0.02, 3),
                                                    (33, 1000, -9, 9, 0.004, 4), (12, 50, 5, 5, 1.0, 5)])
def test_code_and_identity(port, n, L, lo, hi, density, seed):
    bp = synth.random_bp(n, L, seed=seed, lo=lo, hi=hi, density=density)
    words, K = layout(bp)
    assert K == int((bp.max(0).astype(np.int64) - bp.min(0)).sum())
    base, sel, mask = words[:, 0].astype(np.int64), words[:, 1], words[:, 3]
    valid = np.stack([(mask >> (8 * e)) & 1 for e in range(4)], 1).astype(bool)
    assert int(valid.sum()) == K                          # every threshold of every position exactly once
    off = np.stack([(sel >> (4 * e)) & 0xF for e in range(4)], 1).astype(np.int64)
    assert (off[valid] <= 7).all()                        # within the two words the kernel loads
    pos = (4 * base[:, None] + off)[valid]
    assert (np.diff(pos) >= 0).all() and pos.max(initial=0) < L
    thr = np.stack([(words[:, 2] >> (8 * e)) & 0xFF for e in range(4)], 1)[valid].astype(np.int64) - 128
    for q in np.unique(pos):                              # position q: thresholds lo+1 .. hi, in order
        assert thr[pos == q].tolist() == list(range(int(bp[:, q].min()) + 1, int(bp[:, q].max()) + 1))
    T = expand(bp, words).astype(np.int64)
    s = T.sum(1)
    assert np.array_equal(s, (bp - bp.min(0)).sum(1))     # a row's code holds v - lo ones per position
    D = s[:, None] + s[None, :] - 2 * (T @ T.T)
    assert np.array_equal(D, port.distance_matrix(bp))


def test_layout_rejects_bad_ranges():
    lib = _capi.load()
    lo, hi = np.array([3], np.int32), np.array([2], np.int32)
    assert lib.lzg_distance_tc_layout(lo.ctypes.data_as(C.c_void_p), hi.ctypes.data_as(C.c_void_p), 1, None, 0, None) == -1
    lo, hi = np.array([-200], np.int32), np.array([2], np.int32)
    assert lib.lzg_distance_tc_layout(lo.ctypes.data_as(C.c_void_p), hi.ctypes.data_as(C.c_void_p), 1, None, 0, None) == -1


def test_identity_on_random_shapes_and_ranges(port):
    """Seeded sweep over shapes, value ranges and densities (isolated wide positions far apart, constant columns, full-range
    columns next to narrow ones): the code's identity must reproduce the oracle's matrix every time."""
    rng = np.random.default_rng(2024)
    for case in range(40):
        n, L = int(rng.integers(2, 24)), int(rng.integers(1, 400))
        bp = np.zeros((n, L), np.int32)
        for q in rng.choice(L, size=int(rng.integers(0, L + 1)), replace=False):
            lo = int(rng.integers(-128, 128))
            hi = int(rng.integers(lo, min(127, lo + int(rng.choice([0, 1, 2, 7, 255]))) + 1))
            bp[:, q] = rng.integers(lo, hi + 1, n)
        words, K = layout(bp)
        assert K == int((bp.max(0).astype(np.int64) - bp.min(0)).sum())
        T = expand(bp, words).astype(np.int64)
        s = T.sum(1)
        assert np.array_equal(s[:, None] + s[None, :] - 2 * (T @ T.T), port.distance_matrix(bp)), case
