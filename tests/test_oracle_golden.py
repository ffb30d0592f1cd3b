"""CPU: the C restatement (oracle/zcnt_oracle.c) against the committed golden vectors, which were
produced by the unmodified reference (oracle/make_golden.py).  This is what pins the oracle on a
box where /root/reference is absent."""
import glob
import os

import numpy as np
import pytest

from oracle import PortTree

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TREE_CASES = sorted(glob.glob(os.path.join(GOLD, "tree_*.npz")))


def seg_from_bins(bins):
    """Segment offsets of a position list already in reference order (one allele)."""
    chrom = bins[:, 0]
    cut = np.flatnonzero(chrom[1:] != chrom[:-1]) + 1
    return np.concatenate([[0], cut, [len(chrom)]]).astype(np.int32)


def test_ptx008_bp_and_distance(port):
    g = np.load(os.path.join(GOLD, "ptx008.npz"))
    seg = seg_from_bins(g["bins"])
    assert len(seg) - 1 == 22
    cn = g["cn_file_order"][:, g["perm"]]
    for k in range(cn.shape[0]):
        bp = port.bp_from_cn(seg, cn[k])
        assert np.array_equal(bp, g["bp"][k])
        assert np.array_equal(port.cn_from_bp(seg, bp), cn[k])
    D = port.distance_matrix(g["bp"])
    assert np.array_equal(D, g["D"])
    assert list(D[0]) == [0, 32, 43, 47, 42, 41, 47, 40, 33, 36]


def test_ptx008_cli_matrix_matches(port):
    g = np.load(os.path.join(GOLD, "ptx008.npz"))
    with open(os.path.join(GOLD, "ptx008_cli", "PTX008_dist_matrix.txt")) as f:
        lines = f.read().split("\n")
    assert int(lines[0]) == 10
    D = port.distance_matrix(g["bp"])
    for i in range(10):
        tok = lines[1 + i].split()
        assert tok[0] == str(g["names"][i])
        assert [int(x) for x in tok[1:]] == list(D[i])


def load_tree(port, g):
    n = len(g["child0"])
    n_leaves = (n + 1) // 2
    leaf_bp = np.zeros((n, g["bp"].shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = g["bp"]
    return PortTree(port, g["child0"], g["child1"], leaf_bp)


@pytest.mark.parametrize("path", TREE_CASES, ids=[os.path.basename(p)[:-4] for p in TREE_CASES])
def test_tree_case(port, path):
    g = np.load(path)
    from lazac_copy_number_b200 import synth
    assert np.array_equal(synth.bp_from_cn(g["cn"], g["seg"]), g["bp"])
    for k in range(min(5, g["cn"].shape[0])):
        assert np.array_equal(port.bp_from_cn(g["seg"], g["cn"][k].astype(np.int32)), g["bp"][k])
    assert np.array_equal(port.distance_matrix(g["bp"].astype(np.int32)), g["D"])
    t = load_tree(port, g)
    assert t.small_rectilinear() == int(g["root_score"])
    internal = g["child0"] >= 0
    assert np.array_equal(t.score, g["score"])
    assert np.array_equal(t.start, g["start"]) and np.array_equal(t.end, g["end"])
    eu, ev = t.edges()
    assert np.array_equal(eu, g["eu"]) and np.array_equal(ev, g["ev"])
    full = t.greedy_nni(eu, ev, g["order"], greedy=False)
    assert np.array_equal(full["moves"], g["moves"])
    assert np.array_equal(full["scores"], g["cand_scores"])
    want = tuple(int(x) for x in g["move_best"])
    assert (full["move"] or (-1, -1, -1, -1)) == want
    first = t.greedy_nni(eu, ev, g["order"], greedy=True, want_all=False)
    assert (first["move"] or (-1, -1, -1, -1)) == tuple(int(x) for x in g["move_first"])
    old, par, lab, br = t.ancestral_labeling()
    assert np.array_equal(old, g["lab_old"]) and np.array_equal(par, g["lab_parent"])
    assert np.array_equal(lab, g["lab_labels"]) and np.array_equal(br, g["lab_branch"])
    assert int(br[1:].sum()) == int(g["root_score"])
    if want[0] >= 0:
        assert t.apply_nni(*want) == int(g["applied_root_score"])
        assert np.array_equal(t.score[internal], g["applied_score"][internal])
        assert np.array_equal(t.start, g["applied_start"]) and np.array_equal(t.end, g["applied_end"])


def test_simulation_distance(port):
    g = np.load(os.path.join(GOLD, "sim_n100_l1000_s0.npz"))
    assert np.array_equal(port.distance_matrix(g["bp"].astype(np.int32)), g["D"])
    t = load_tree(port, g)
    assert t.small_rectilinear() == int(g["score_before"])
    t2 = PortTree(port, g["hc_child0"], g["hc_child1"], t.start * (g["child0"] < 0)[:, None])
    assert t2.small_rectilinear() == int(g["hc_score"])
