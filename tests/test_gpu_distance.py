"""GPU parity: all-pairs ZCNT distance through the C ABI vs the oracle (bit-exact)."""
import numpy as np
import pytest

from lazac_copy_number_b200 import synth
from lazac_copy_number_b200._capi import LazacGpuError
from lazac_copy_number_b200.api import Profiles

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,L", [(1, 1), (2, 3), (10, 255), (63, 129), (64, 128), (65, 127), (130, 1000),
                                 (257, 513), (300, 2048), (1000, 700)])
def test_distance_matches_oracle(ctx, port, n, L):
    bp = synth.random_bp(n, L, seed=n * 7 + L, lo=-9, hi=9, density=0.3)
    prof = Profiles(ctx, bp)
    D = prof.distance_matrix()
    prof.free()
    assert D.dtype == np.int32 and D.shape == (n, n)
    assert np.array_equal(D, port.distance_matrix(bp))


def test_distance_extreme_values(ctx, port):
    rng = np.random.default_rng(5)
    bp = rng.choice(np.array([-127, 127, 0, 1, -1], np.int32), size=(70, 333))
    prof = Profiles(ctx, bp)
    assert np.array_equal(prof.distance_matrix(), port.distance_matrix(bp))
    prof.free()


def test_distance_int8_upload(ctx, port):
    bp = synth.random_bp(90, 777, seed=3).astype(np.int8)
    prof = Profiles(ctx, bp)
    assert np.array_equal(prof.distance_matrix(), port.distance_matrix(bp.astype(np.int32)))
    prof.free()


def test_range_error(ctx):
    """Values beyond the 16-bit device format (tests/test_gpu_int16.py covers everything up to it)."""
    bp = np.zeros((4, 10), np.int32)
    bp[2, 3] = 32768
    with pytest.raises(LazacGpuError) as e:
        Profiles(ctx, bp)
    assert e.value.status == 3


def test_empty(ctx):
    prof = Profiles(ctx, np.zeros((0, 5), np.int32))
    assert prof.distance_matrix().shape == (0, 0)
    prof.free()
    prof = Profiles(ctx, np.zeros((3, 0), np.int32))
    assert np.array_equal(prof.distance_matrix(), np.zeros((3, 3), np.int32))
    prof.free()


def test_large_tile_path_vs_properties(ctx, port):
    """128x128-tile path (many cells): rows checked against the oracle, whole matrix by properties."""
    n, L = 2500, 640
    bp = synth.random_bp(n, L, seed=11, density=0.2)
    prof = Profiles(ctx, bp)
    D = prof.distance_matrix()
    prof.free()
    assert np.array_equal(D, D.T) and not D.diagonal().any()
    rows = port.distance_rows(bp, 1234, 1250)
    for k in range(16):
        assert np.array_equal(D[1234 + k, 1234 + k:], rows[k, 1234 + k:])
    assert np.array_equal(D[0], np.abs(bp - bp[0]).sum(1))
    assert np.array_equal(D[n - 1], np.abs(bp - bp[n - 1]).sum(1))


@pytest.mark.parametrize("n_shards", [2, 3, 8])
def test_shards_union_is_full_matrix(ctx, port, n_shards):
    n, L = 700, 300
    bp = synth.random_bp(n, L, seed=2)
    prof = Profiles(ctx, bp)
    D = np.full((n, n), -1, np.int32)
    pairs = sum(prof.distance_shard(s, n_shards, D) for s in range(n_shards))
    prof.free()
    assert pairs == n * (n + 1) // 2
    assert np.array_equal(D, port.distance_matrix(bp))


def test_row_group_overlap_path(ctx):
    """>= 64 MiB of output: computed in tile-row groups with the D2H of finished rows overlapped."""
    n, L = 4200, 200
    bp = synth.random_bp(n, L, seed=13, density=0.3)
    prof = Profiles(ctx, bp.astype(np.int8))
    D = ctx.pinned((n, n), np.int32)
    D[...] = -1
    prof.distance_matrix(D)
    prof.free()
    assert np.array_equal(D, D.T) and not D.diagonal().any()
    for i in (0, 1, 127, 128, 2047, 2100, 4095, 4096, 4199):
        assert np.array_equal(D[i], np.abs(bp - bp[i]).sum(1))


def test_repeated_upload_free_cycles(ctx, port):
    """The pooled allocations are reused across upload/distance/free cycles of different shapes."""
    for k, (n, L) in enumerate([(300, 1000), (50, 77), (300, 1000), (129, 4097), (300, 1000)]):
        bp = synth.random_bp(n, L, seed=20 + k, density=0.2)
        prof = Profiles(ctx, bp.astype(np.int8) if k % 2 else bp)
        assert np.array_equal(prof.distance_matrix(), port.distance_matrix(bp))
        prof.free()


def test_int8_minus_128_uses_the_wide_format(ctx):
    bp = np.zeros((3, 40), np.int8)
    bp[1, 39] = -128
    prof = Profiles(ctx, bp)
    D = prof.distance_matrix()
    prof.free()
    assert D[1, 0] == 128 and D[0, 2] == 0


def test_tile_orders_give_the_same_matrix(port, monkeypatch):
    """The plan's tile order (row-major list, super-blocks, phase-aligned whole tiles)
    only decides which CTA computes what and when: 26 tile rows = 325 off-diagonal tiles is enough
    to switch the re-ordering on, and every order must give the oracle's matrix bit for bit."""
    from lazac_copy_number_b200.api import Context
    n, L = 3300, 200
    bp = synth.random_bp(n, L, seed=21, density=0.3)
    want = port.distance_matrix(bp)
    for sb, aligned in (("0", "0"), ("12", "0"), ("12", "1"), ("5", "1")):
        monkeypatch.setenv("LZG_DIST_SB", sb)
        monkeypatch.setenv("LZG_DIST_ALIGN", aligned)
        with Context(0) as own:          # schedules are cached per context
            prof = Profiles(own, bp)
            got = prof.distance_matrix()
            prof.free()
        assert np.array_equal(got, want), (sb, aligned)
