"""GPU: lzg_nj_seed_tree (Clearcut-exact neighbour joining on the device) against the host
restatement of the reference's NJ (`lazac nj`, itself byte-identical to the reference CLI's
PREFIX_nj_tree.newick on the golden matrices -- tests/test_host_cpu.py).  Integer matrices with
many ties and with row sums beyond 2^24, so both the tie rule and the float32 summation order
are exercised."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
LAZAC = os.path.join(ROOT, "host", "lazac")
if not os.path.exists(LAZAC):
    subprocess.run(["make", "-C", os.path.join(ROOT, "host")], check=False, stdout=subprocess.DEVNULL)


def write_matrix(path, D):
    """PREFIX_dist_matrix.txt as write_matrix_txt / the reference writes it (src/lazac.cxx:118-131)."""
    n = len(D)
    with open(path, "w") as f:
        f.write("%d\n" % n)
        for i in range(n):
            f.write("c%05d " % i + " ".join(str(int(x)) for x in D[i]) + "\n")


def nj_both(tmp_path, D, seed):
    src = tmp_path / "m.txt"
    write_matrix(src, D)
    outs = []
    for mode in ("host", "gpu"):
        out = tmp_path / ("nj_%s.newick" % mode)
        subprocess.run([LAZAC, "nj", str(src), "-s", str(seed), "-o", str(out)], check=True, env=dict(os.environ, LAZAC_NJ=mode))
        outs.append(out.read_bytes())
    return outs


def l1_matrix(n, L, lo, hi, seed):
    rng = np.random.default_rng(seed)
    x = rng.integers(lo, hi + 1, size=(n, L))
    return np.abs(x[:, None, :] - x[None, :, :]).sum(axis=2)


@pytest.mark.parametrize("n,L,hi,seed", [(2, 5, 3, 0), (3, 5, 3, 1), (4, 6, 1, 2), (17, 12, 1, 3), (150, 40, 2, 4), (600, 30, 1, 5)])
def test_nj_gpu_equals_host_on_tie_heavy_matrices(tmp_path, n, L, hi, seed):
    host, gpu = nj_both(tmp_path, l1_matrix(n, L, 0, hi, seed), seed)
    assert host == gpu


def test_nj_gpu_equals_host_with_large_row_sums(tmp_path):
    """Row sums far beyond 2^24: every float32 addition rounds, so the summation order matters."""
    rng = np.random.default_rng(9)
    n = 700
    D = rng.integers(1, 1 << 20, size=(n, n))
    D = np.triu(D, 1)
    D = D + D.T
    host, gpu = nj_both(tmp_path, D, 11)
    assert host == gpu


@pytest.mark.parametrize("matrix,seed,tree", [
    ("ptx008_cli/PTX008_dist_matrix.txt", 0, "ptx008_cli/PTX008_nj_tree.newick"),
    ("synth_cli/full_dist_matrix.txt", 3, "synth_cli/full_nj_tree.newick"),
    ("synth_cli/greedy_dist_matrix.txt", 5, "synth_cli/greedy_nj_tree.newick"),
])
def test_nj_gpu_reproduces_the_reference_cli_files(tmp_path, matrix, seed, tree):
    out = tmp_path / "nj.newick"
    subprocess.run([LAZAC, "nj", os.path.join(GOLD, matrix), "-s", str(seed), "-o", str(out)], check=True,
                   env=dict(os.environ, LAZAC_NJ="gpu"))
    with open(os.path.join(GOLD, tree), "rb") as f:
        assert out.read_bytes() == f.read()


def test_nj_api_shapes_and_errors(ctx):
    from lazac_copy_number_b200._capi import LazacGpuError
    D = l1_matrix(9, 10, 0, 2, 1)
    left, right = ctx.nj_seed_tree(D, np.arange(9))
    assert (left[:9] == -1).all() and (right[:9] == -1).all()
    kids = np.concatenate([left[9:], right[9:]])
    assert sorted(kids.tolist()) == list(range(16))          # every node but the root is a child exactly once
    with pytest.raises(LazacGpuError):
        ctx.nj_seed_tree(D, np.zeros(9, np.int32))              # not a permutation
    big = D.copy()
    big[0, 1] = big[1, 0] = 1 << 24
    with pytest.raises(LazacGpuError):
        ctx.nj_seed_tree(big, np.arange(9))                     # not exact in float32


def test_nj_from_device_matrix_and_from_profiles(ctx):
    """lzg_nj_seed_tree_device / lzg_nj_seed_tree_profiles (the matrix never leaves HBM) against lzg_nj_seed_tree on the host copy;
    the profiles form both after lzg_distance_matrix (device copy reused) and on fresh profiles (computed now)."""
    import torch
    from lazac_copy_number_b200 import search, synth
    from lazac_copy_number_b200.api import Profiles
    from lazac_copy_number_b200._capi import LazacGpuError
    for n, L, seed in ((5, 40, 1), (200, 333, 2), (1100, 600, 3)):
        bp = synth.random_bp(n, L, seed=seed, lo=-4, hi=4, density=0.3)
        perm = search.clearcut_shuffle(n, seed)
        fresh = Profiles(ctx, bp)
        l0, r0 = fresh.nj_seed_tree(perm)            # no matrix yet: computed on the device now
        D = fresh.distance_matrix()
        l1, r1 = fresh.nj_seed_tree(perm)            # the copy lzg_distance_matrix left in HBM
        fresh.free()
        want_l, want_r = ctx.nj_seed_tree(D, perm)
        d = torch.from_numpy(D).to("cuda:0")
        l2, r2 = ctx.nj_seed_tree_device(d.data_ptr(), n, perm)
        assert np.array_equal(d.cpu().numpy(), D)    # left untouched
        for left, right in ((l0, r0), (l1, r1), (l2, r2)):
            assert np.array_equal(left, want_l) and np.array_equal(right, want_r), n
    big = torch.full((9, 9), 1 << 24, dtype=torch.int32, device="cuda:0")
    with pytest.raises(LazacGpuError):
        ctx.nj_seed_tree_device(big.data_ptr(), 9, np.arange(9))   # not exact in float32: flagged by the gather


def test_nj_gpu_equals_host_library(ctx):
    """lzg_nj_seed_tree against liblazac_host.so's CPU neighbour joining (the code `lazac` links) with the same taxon order."""
    from lazac_copy_number_b200 import search
    for n, seed in ((2, 1), (3, 2), (17, 3), (300, 4)):
        rng = np.random.default_rng(seed)
        D = rng.integers(1, 400, (n, n)).astype(np.int32)
        D = np.triu(D, 1) + np.triu(D, 1).T
        perm = search.clearcut_shuffle(n, seed)
        left, right = ctx.nj_seed_tree(D, perm)
        cl, cr = search.nj_seed_tree_cpu(D, perm)
        assert np.array_equal(left, cl) and np.array_equal(right, cr), n


@pytest.mark.parametrize("n,seed,hi", [(1500, 21, 400), (2500, 22, 3), (1026, 23, 50), (13001, 24, 9)])
def test_nj_persistent_kernel_on_larger_matrices(ctx, monkeypatch, n, seed, hi):
    """All joins in one cooperative launch (k_nj_fused: one grid barrier per join, the update of the last join virtual
    during the scan; k_nj_persistent: two barriers; the sequential row sum overlapped with the next scan) against the CPU restatement and against the two-launches-per-join form; hi = 3 makes almost every scan a tie."""
    from lazac_copy_number_b200 import search
    rng = np.random.default_rng(seed)
    D = rng.integers(1, hi + 1, (n, n)).astype(np.int32)
    D = np.triu(D, 1) + np.triu(D, 1).T
    perm = search.clearcut_shuffle(n, seed)
    monkeypatch.setenv("LZG_NJ_PERSISTENT", "2")  # k_nj_fused at every size it has shared memory for
    left, right = ctx.nj_seed_tree(D, perm)
    monkeypatch.delenv("LZG_NJ_PERSISTENT")
    if n <= 3000:
        cl, cr = search.nj_seed_tree_cpu(D, perm)
    else:  # minutes on one core: the per-join kernels (pinned to the CPU restatement above) stand in
        monkeypatch.setenv("LZG_NJ_PERSISTENT", "0")
        cl, cr = ctx.nj_seed_tree(D, perm)
        monkeypatch.delenv("LZG_NJ_PERSISTENT")
    assert np.array_equal(left, cl) and np.array_equal(right, cr)
    monkeypatch.setenv("LZG_NJ_PERSISTENT", "1")  # two grid barriers per join (k_nj_persistent): the form for n > 14,000
    l4, r4 = ctx.nj_seed_tree(D, perm)
    assert np.array_equal(l4, cl) and np.array_equal(r4, cr)
    monkeypatch.setenv("LZG_NJ_EARLY", "0")  # ... for n > 14,000: the joined row's terms through a device vector
    l5, r5 = ctx.nj_seed_tree(D, perm)
    assert np.array_equal(l5, cl) and np.array_equal(r5, cr)
    monkeypatch.delenv("LZG_NJ_EARLY")
    monkeypatch.setenv("LZG_NJ_SMEM", "0")  # ... and for n > 28,000: r2 and the joined row's terms straight from L2
    l3, r3 = ctx.nj_seed_tree(D, perm)
    assert np.array_equal(l3, cl) and np.array_equal(r3, cr)
    monkeypatch.delenv("LZG_NJ_SMEM")
    monkeypatch.setenv("LZG_NJ_PERSISTENT", "0")
    l2, r2 = ctx.nj_seed_tree(D, perm)
    assert np.array_equal(l2, cl) and np.array_equal(r2, cr)
