"""GPU parity of the 16-bit device format (row a1 of SURVEY.md 8: the reference's profiles are plain int, so a
breakpoint jump above 127 -- a focal amplification -- just works there, src/copy_number.cxx:100-113).  Uploads
whose values leave [-127, 127] are stored as 16-bit words (|v| <= 32767); every kernel family has a 16-bit
instantiation: the distance kernel (one VABSDIFF per position), full scoring, the neighbourhood pass, move
application, labeling, the device-side climb, the multi-GPU distance.  LZG_ERR_RANGE only beyond int16."""
import numpy as np
import pytest

from lazac_copy_number_b200 import synth
from lazac_copy_number_b200._capi import LazacGpuError
from lazac_copy_number_b200.api import MultiContext, Profiles, Tree
from oracle import PortTree
from test_gpu_climb import check_climb

pytestmark = pytest.mark.gpu


def wide_bp(n, L, seed, big=30000, density=0.2, big_fraction=0.05):
    """Sparse breakpoint vectors, mostly small values, a few of them up to +-big."""
    rng = np.random.default_rng(seed)
    v = rng.integers(-9, 10, size=(n, L))
    wide = rng.integers(-big, big + 1, size=(n, L))
    v = np.where(rng.random((n, L)) < big_fraction, wide, v)
    return np.where(rng.random((n, L)) < density, v, 0).astype(np.int32)


@pytest.mark.parametrize("n,L", [(1, 1), (2, 3), (10, 255), (63, 129), (65, 127), (130, 1000), (300, 2048), (1000, 700)])
def test_distance_wide_values(ctx, port, n, L):
    bp = wide_bp(n, L, seed=n + L)
    bp[0, 0] = 32767
    bp[n - 1, L - 1] = -32767
    prof = Profiles(ctx, bp)
    D = prof.distance_matrix()
    prof.free()
    assert np.array_equal(D, port.distance_matrix(bp))


def test_one_value_of_128_takes_the_wide_format(ctx, port):
    bp = synth.random_bp(70, 333, seed=5)
    bp[2, 3] = 128                                  # LZG_ERR_RANGE in round 1
    prof = Profiles(ctx, bp)
    assert np.array_equal(prof.distance_matrix(), port.distance_matrix(bp))
    prof.free()
    i8 = synth.random_bp(40, 100, seed=6).astype(np.int8)
    i8[1, 39] = -128                                # the one int8 value the 8-bit format lacks
    prof = Profiles(ctx, i8)
    assert np.array_equal(prof.distance_matrix(), port.distance_matrix(i8.astype(np.int32)))
    prof.free()


def test_range_error_only_beyond_int16(ctx):
    for bad in (32768, -32768, 40000, -(1 << 20)):
        bp = np.zeros((4, 10), np.int32)
        bp[2, 3] = bad
        with pytest.raises(LazacGpuError) as e:
            Profiles(ctx, bp)
        assert e.value.status == 3
    ok = np.zeros((4, 10), np.int32)
    ok[2, 3] = -32767
    Profiles(ctx, ok).free()


def test_forced_wide_format_equals_narrow(ctx, port, monkeypatch):
    """The 16-bit kernels on data that also fits 8 bits: same matrix, same scores."""
    bp = synth.random_bp(300, 900, seed=9, density=0.3)
    prof = Profiles(ctx, bp)
    narrow = prof.distance_matrix()
    prof.free()
    monkeypatch.setenv("LZG_FORCE_WIDTH", "2")
    prof = Profiles(ctx, bp)
    assert np.array_equal(prof.distance_matrix(), narrow)
    prof.free()


def tree_case(n_leaves, L, seed):
    bp = wide_bp(n_leaves, L, seed)
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(seed + 100))
    n = 2 * n_leaves - 1
    leaf_row = np.full(n, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    leaf_bp = np.zeros((n, L), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    return bp, c0, c1, leaf_row, leaf_bp


@pytest.mark.parametrize("n_leaves,L,greedy", [(3, 40, False), (10, 255, False), (60, 1000, False), (60, 1000, True), (150, 2100, False)])
def test_tree_kernels_wide_values(ctx, port, n_leaves, L, greedy):
    bp, c0, c1, leaf_row, leaf_bp = tree_case(n_leaves, L, seed=3 * n_leaves)
    prof = Profiles(ctx, bp)
    gt = Tree(ctx, prof, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    S, E, sc = gt.intervals()
    assert np.array_equal(sc, pt.score) and np.array_equal(S, pt.start) and np.array_equal(E, pt.end)
    got, want = gt.ancestral_labeling(), pt.ancestral_labeling()
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    eu, ev = pt.edges()
    order = np.random.default_rng(1).permutation(len(eu)).astype(np.int32)
    full = pt.greedy_nni(eu, ev, order, greedy=False)
    want = pt.greedy_nni(eu, ev, order, greedy=greedy, want_all=False)
    moves = gt.enumerate_moves(eu, ev, order)
    idx, score, allsc = gt.nni_best(moves, greedy=greedy, want_all=True)
    assert np.array_equal(allsc, full["scores"])
    assert idx == want["index"]
    if want["found"]:
        assert gt.apply_nni(*moves[idx]) == pt.apply_nni(*want["move"]) == score
        S, E, sc = gt.intervals()
        assert np.array_equal(sc, pt.score) and np.array_equal(S, pt.start) and np.array_equal(E, pt.end)
    gt.free()
    prof.free()


@pytest.mark.parametrize("n_leaves,L,greedy", [(4, 60, False), (50, 900, False), (50, 900, True), (120, 1500, False)])
def test_climb_wide_values(ctx, port, n_leaves, L, greedy):
    bp, c0, c1, leaf_row, _ = tree_case(n_leaves, L, seed=7 + n_leaves)
    check_climb(ctx, port, bp, c0, c1, leaf_row, greedy=greedy, rounds=2)


def test_multi_distance_wide_values(port):
    bp = wide_bp(400, 600, seed=4)
    with MultiContext([0]) as m:
        D, pairs = m.distance_matrix(bp)
        assert pairs == 400 * 401 // 2 and np.array_equal(D, port.distance_matrix(bp))
        bad = bp.copy()
        bad[5, 5] = 70000
        with pytest.raises(LazacGpuError) as e:
            m.distance_matrix(bad)
        assert e.value.status == 3
        assert np.array_equal(m.distance_matrix(bp)[0], D)
