"""GPU: build_distance_matrix on several GPUs from one process (lzg_distance_matrix_multi) -- every GPU
uploads a row block and pushes it to its peers, computes its tile rows, writes the mirror of every
tile into the owner's memory over NVLink, and copies the rows it owns into the ONE host matrix.
The result must be the oracle's matrix and, byte for byte, what one GPU computes."""
import numpy as np
import pytest

from lazac_copy_number_b200 import synth
from lazac_copy_number_b200._capi import LazacGpuError
from lazac_copy_number_b200.api import MultiContext, Profiles

pytestmark = pytest.mark.gpu


def n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("n,L,dtype", [(1, 1, np.int32), (10, 255, np.int32), (64, 128, np.int8), (65, 127, np.int32),
                                       (300, 1000, np.int8), (700, 300, np.int32), (1000, 700, np.int8)])
def test_one_device_row_table_path(port, n, L, dtype):
    """G = 1: the same code path (row-block table, owner copies) on the one GPU every box has."""
    bp = synth.random_bp(n, L, seed=n + L, density=0.3)
    with MultiContext([0]) as m:
        D, pairs = m.distance_matrix(bp.astype(dtype))
    assert pairs == n * (n + 1) // 2
    assert np.array_equal(D, port.distance_matrix(bp))


def test_range_and_empty(port):
    with MultiContext([0]) as m:
        bad = np.zeros((5, 40), np.int32)
        bad[3, 7] = 40000
        with pytest.raises(LazacGpuError) as e:
            m.distance_matrix(bad)
        assert e.value.status == 3
        D, pairs = m.distance_matrix(np.zeros((0, 9), np.int32))
        assert D.shape == (0, 0) and pairs == 0
        D, _ = m.distance_matrix(np.zeros((4, 0), np.int32))
        assert not D.any()
        bp = synth.random_bp(200, 333, seed=1)          # the context is reusable after an error
        assert np.array_equal(m.distance_matrix(bp)[0], port.distance_matrix(bp))


@pytest.mark.parametrize("n,L", [(130, 500), (1000, 700), (2500, 640), (4200, 200)])
def test_multi_gpu_equals_one_gpu(ctx, port, n, L):
    g = n_gpus()
    if g < 2:
        pytest.skip("needs at least two GPUs")
    bp = synth.random_bp(n, L, seed=3 * n + L, density=0.25).astype(np.int8)
    prof = Profiles(ctx, bp)
    one = prof.distance_matrix()
    prof.free()
    for devices in ([0, 1], list(range(g))):
        with MultiContext(devices) as m:
            D = np.full((n, n), -1, np.int32)
            _, pairs = m.distance_matrix(bp, out=D)
            assert pairs == n * (n + 1) // 2
            assert D.tobytes() == one.tobytes(), devices
            _, _ = m.distance_matrix(bp.astype(np.int32), out=D)      # the int32 upload path, buffers reused
            assert D.tobytes() == one.tobytes(), devices
    if n <= 1000:
        assert np.array_equal(one, port.distance_matrix(bp.astype(np.int32)))


def test_tensor_core_path_through_the_row_table(ctx, monkeypatch):
    """Narrow value ranges and enough tiles: every GPU of lzg_distance_matrix_multi runs k_tc_distance on its row blocks (direct
    rows and mirrors through the row-block table, peers included); same bytes as k_zcnt_distance on one GPU."""
    n, L = 3300, 500
    bp = synth.random_bp(n, L, seed=17, lo=-2, hi=2, density=0.06).astype(np.int8)
    monkeypatch.setenv("LZG_TC", "0")
    prof = Profiles(ctx, bp)
    want = prof.distance_matrix()
    prof.free()
    monkeypatch.delenv("LZG_TC")
    prof = Profiles(ctx, bp)
    one = prof.distance_matrix()
    assert prof.distance_path()["tensor_core"]
    prof.free()
    assert one.tobytes() == want.tobytes()
    g = n_gpus()
    for devices in ([0], [0, 1], list(range(g))):
        if len(devices) > g:
            continue
        with MultiContext(devices) as m:
            for phases in ("1", "3"):
                monkeypatch.setenv("LZG_MULTI_PHASES", phases)
                D = np.full((n, n), -1, np.int32)
                _, pairs = m.distance_matrix(bp, out=D)
                assert pairs == n * (n + 1) // 2
                assert D.tobytes() == want.tobytes(), (devices, phases)
