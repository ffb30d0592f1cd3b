"""GPU parity of the tensor-core distance (lzg_distance_tc.cu: thermometer code + tcgen05 kind::i8 product), through the C ABI,
bit-exact against the oracle and against k_zcnt_distance (LZG_TC=0) on the same profiles."""
import numpy as np
import pytest
import torch

from lazac_copy_number_b200 import synth
from lazac_copy_number_b200.api import Profiles

pytestmark = pytest.mark.gpu


def _alu(ctx, monkeypatch, bp):
    monkeypatch.setenv("LZG_TC", "0")
    p = Profiles(ctx, bp)
    D = p.distance_matrix()
    assert not p.distance_path()["tensor_core"]
    p.free()
    monkeypatch.delenv("LZG_TC")
    return D


@pytest.mark.parametrize("n,L,lo,hi,density", [(256, 300, -3, 3, 0.05), (257, 129, -1, 1, 0.5), (700, 1000, -9, 9, 0.3), (1000, 64, 0, 1, 0.01),
                                               (1531, 517, -2, 2, 0.02), (300, 100, -127, 127, 1.0)])
def test_forced_tensor_core_path_matches_oracle(ctx, port, monkeypatch, n, L, lo, hi, density):
    """LZG_TC=1: the path at shapes below its dispatch threshold, odd sizes, the full 8-bit range (K = 254 L)."""
    bp = synth.random_bp(n, L, seed=n * 3 + L, lo=lo, hi=hi, density=density)
    want = port.distance_matrix(bp)
    monkeypatch.setenv("LZG_TC", "1")
    p = Profiles(ctx, bp)
    got = p.distance_matrix()
    path = p.distance_path()
    p.free()
    assert path["tensor_core"] and path["launches"] >= 1
    assert path["K"] == int((bp.max(0).astype(np.int64) - bp.min(0)).sum())
    assert np.array_equal(got, want)


def test_constant_columns_and_identical_rows(ctx, port, monkeypatch):
    monkeypatch.setenv("LZG_TC", "1")
    bp = np.tile(np.arange(-3, 4, dtype=np.int32).repeat(20), (400, 1))  # K = 0: every distance is zero
    p = Profiles(ctx, bp)
    assert not p.distance_matrix().any()
    assert p.distance_path() == {"tensor_core": True, "K": 0, "k_pad": 512, "launches": 1}
    p.free()
    bp = bp.copy()
    bp[7, 5] = 100   # one cell differs at one position
    bp[9, 5] = -100
    p = Profiles(ctx, bp)
    assert np.array_equal(p.distance_matrix(), port.distance_matrix(bp))
    p.free()


def test_default_dispatch_by_size_and_value_range(ctx, port, monkeypatch):
    """Narrow values and enough tiles: tensor cores; wide values (K > 8 L), few tiles or the 16-bit format: k_zcnt_distance."""
    narrow = synth.random_bp(3000, 400, seed=1, lo=-2, hi=2, density=0.1)
    p = Profiles(ctx, narrow)
    D = p.distance_matrix()
    assert p.distance_path()["tensor_core"]
    p.free()
    assert np.array_equal(D, _alu(ctx, monkeypatch, narrow))
    for bp in (synth.random_bp(3000, 300, seed=2, lo=-60, hi=60, density=1.0),   # K = 120 L
               synth.random_bp(900, 400, seed=3, lo=-2, hi=2, density=0.1),      # 16 tiles
               synth.random_bp(3000, 200, seed=4, lo=-3000, hi=3000, density=0.3)):
        p = Profiles(ctx, bp)
        p.distance_matrix()
        assert not p.distance_path()["tensor_core"]
        p.free()


def test_row_groups_shards_and_device_entry_points(ctx, port, monkeypatch):
    """n = 4,500: lzg_distance_matrix works in groups of row blocks (81 MB of output); shards deal the row blocks; the device
    entry points write the same matrix."""
    n, L = 4500, 700
    bp = synth.random_bp(n, L, seed=9, lo=-2, hi=3, density=0.04).astype(np.int8)
    want = _alu(ctx, monkeypatch, bp)
    p = Profiles(ctx, bp)
    got = p.distance_matrix()
    path = p.distance_path()
    assert path["tensor_core"] and path["launches"] > 1
    assert np.array_equal(got, want)
    d = torch.full((n, n), -7, dtype=torch.int32, device="cuda:0")
    p.distance_device_async(d.data_ptr())
    ctx.sync()
    assert np.array_equal(d.cpu().numpy(), want)
    # three shards into one host matrix: together they write every entry
    out = np.full((n, n), -1, np.int32)
    pairs = sum(p.distance_shard(k, 3, out) for k in range(3))
    assert pairs == n * (n + 1) // 2
    assert np.array_equal(out, want)
    p.free()
