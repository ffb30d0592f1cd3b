"""GPU parity against the committed golden vectors (outputs of the unmodified reference,
oracle/make_golden.py) through the C ABI: bit-exact distances, intervals, scores, candidate
scores, chosen moves, labels."""
import glob
import os

import numpy as np
import pytest

from lazac_copy_number_b200.api import Profiles, Tree

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TREE_CASES = sorted(glob.glob(os.path.join(GOLD, "tree_*.npz")))


def test_ptx008_distance(ctx):
    g = np.load(os.path.join(GOLD, "ptx008.npz"))
    names, D = ctx.build_distance_matrix({str(nm): g["bp"][k] for k, nm in enumerate(g["names"])})
    assert names == [str(x) for x in g["names"]]
    assert np.array_equal(D, g["D"])


def gpu_tree(ctx, g):
    n = len(g["child0"])
    n_leaves = (n + 1) // 2
    prof = Profiles(ctx, g["bp"])  # int8 upload path
    leaf_row = np.full(n, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    return prof, Tree(ctx, prof, g["child0"], g["child1"], leaf_row)


@pytest.mark.parametrize("path", TREE_CASES, ids=[os.path.basename(p)[:-4] for p in TREE_CASES])
def test_tree_case(ctx, path):
    g = np.load(path)
    prof, t = gpu_tree(ctx, g)
    assert np.array_equal(prof.distance_matrix(), g["D"])
    assert t.small_rectilinear() == int(g["root_score"])
    S, E, sc = t.intervals()
    assert np.array_equal(sc, g["score"]) and np.array_equal(S, g["start"]) and np.array_equal(E, g["end"])
    eu, ev = t.edges()
    assert np.array_equal(eu, g["eu"]) and np.array_equal(ev, g["ev"])
    moves = t.enumerate_moves(eu, ev, g["order"])
    assert np.array_equal(moves, g["moves"])
    idx, score, allsc = t.nni_best(moves, greedy=False, want_all=True)
    assert np.array_equal(allsc, g["cand_scores"])
    want = tuple(int(x) for x in g["move_best"])
    assert (tuple(int(x) for x in moves[idx]) if idx >= 0 else (-1,) * 4) == want
    idx_g, _, _ = t.nni_best(moves, greedy=True)
    assert (tuple(int(x) for x in moves[idx_g]) if idx_g >= 0 else (-1,) * 4) == tuple(int(x) for x in g["move_first"])
    old, par, lab, br = t.ancestral_labeling()
    assert np.array_equal(old, g["lab_old"]) and np.array_equal(par, g["lab_parent"])
    assert np.array_equal(lab, g["lab_labels"]) and np.array_equal(br, g["lab_branch"])
    if want[0] >= 0:
        assert score == int(g["applied_root_score"])
        assert t.apply_nni(*want) == int(g["applied_root_score"])
        S, E, sc = t.intervals()
        internal = g["child0"] >= 0
        assert np.array_equal(sc[internal], g["applied_score"][internal])
        assert np.array_equal(S, g["applied_start"]) and np.array_equal(E, g["applied_end"])
    t.free()
    prof.free()


def test_simulation(ctx):
    g = np.load(os.path.join(GOLD, "sim_n100_l1000_s0.npz"))
    prof, t = gpu_tree(ctx, g)
    assert np.array_equal(prof.distance_matrix(), g["D"])
    assert t.small_rectilinear() == int(g["score_before"])
    t.set_topology(g["hc_child0"], g["hc_child1"])
    assert t.small_rectilinear() == int(g["hc_score"])
    t.free()
    prof.free()
