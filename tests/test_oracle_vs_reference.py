"""CPU: the C restatement against the unmodified reference functions (oracle/_ref/libref_shim.so,
compiled from /root/reference/src/copy_number.cxx) on seeded random inputs.  Skipped on a box
where the shim was not built (the golden vectors cover that case)."""
import numpy as np
import pytest

from lazac_copy_number_b200 import synth
from oracle import PortTree


def case(n_leaves, bins, alleles, seed):
    prof = synth.make_profiles(n_leaves, bins, alleles, seed=seed)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(seed + 100))
    n = 2 * n_leaves - 1
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    return prof, bp, c0, c1, leaf_bp


def test_bp_transform_two_alleles_and_single_bin_segment(port, ref):
    # chrom "1" x {cn_a, cn_b} (4 bins) and a single-bin chrom "10"
    chrom = np.array([1, 1, 1, 1, 1, 1, 1, 1, 10], np.int32)
    allele = np.array([0, 0, 0, 0, 1, 1, 1, 1, 0], np.int32)
    start = np.array([0, 1, 2, 3, 0, 1, 2, 3, 0], np.int32)
    cn = np.array([2, 3, 3, 1, 1, 1, 0, 2, 5], np.int32)
    bp, perm = ref.bp_from_cn(chrom, allele, start, start, cn)
    assert list(bp) == [0, 1, 0, 1, -1, 0, -1, 0, 3]
    seg = np.array([0, 4, 8, 9], np.int32)
    assert np.array_equal(port.bp_from_cn(seg, cn[perm]), bp)
    back, _ = ref.cn_from_bp(chrom[perm], allele[perm], start[perm], start[perm], bp)
    assert np.array_equal(port.cn_from_bp(seg, bp), back)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_bp_transform_random_bin_order(port, ref, seed):
    rng = np.random.default_rng(seed)
    prof = synth.make_profiles(3, 97, 2, seed=seed)
    segs, off = prof["segs"], prof["seg"]
    chrom = np.concatenate([np.full(nb, int(ch)) for ch, a, nb in segs]).astype(np.int32)
    allele = np.concatenate([np.full(nb, a) for ch, a, nb in segs]).astype(np.int32)
    start = np.concatenate([np.arange(nb) for ch, a, nb in segs]).astype(np.int32)
    shuffle = rng.permutation(len(chrom))
    for k in range(3):
        cn = prof["cn"][k].astype(np.int32)
        bp, perm = ref.bp_from_cn(chrom[shuffle], allele[shuffle], start[shuffle], start[shuffle], cn[shuffle])
        assert np.array_equal(shuffle[perm], np.arange(len(chrom)))  # reference order == synth order
        assert np.array_equal(bp, port.bp_from_cn(off, cn))


@pytest.mark.parametrize("n,L,seed", [(1, 1, 0), (7, 33, 1), (40, 513, 2)])
def test_distance(port, ref, n, L, seed):
    bp = synth.random_bp(n, L, seed=seed, lo=-127, hi=127, density=0.5)
    assert np.array_equal(port.distance_matrix(bp), ref.distance_matrix(bp))


@pytest.mark.parametrize("n_leaves,bins,alleles,seed", [(2, 9, 1, 0), (5, 60, 2, 1), (30, 250, 1, 2), (64, 400, 2, 3)])
def test_tree_paths(port, ref, n_leaves, bins, alleles, seed):
    _, bp, c0, c1, leaf_bp = case(n_leaves, bins, alleles, seed)
    rt, pt = ref.tree(c0, c1, leaf_bp), PortTree(port, c0, c1, leaf_bp)
    assert rt.small_rectilinear() == pt.small_rectilinear()
    S, E, sc = rt.state()
    assert np.array_equal(S, pt.start) and np.array_equal(E, pt.end) and np.array_equal(sc, pt.score)
    eu, ev = rt.edges()
    peu, pev = pt.edges()
    assert np.array_equal(eu, peu) and np.array_equal(ev, pev)
    order = np.random.default_rng(seed).permutation(len(eu)).astype(np.int32)
    moves, scores = rt.candidate_scores(eu, ev, order)
    full = pt.greedy_nni(eu, ev, order)
    assert np.array_equal(moves, full["moves"]) and np.array_equal(scores, full["scores"])
    for greedy in (False, True):
        want = rt.clone().greedy_nni(eu, ev, order, greedy)
        got = pt.greedy_nni(eu, ev, order, greedy=greedy, want_all=False)["move"]
        assert got == want
    for a, b in zip(rt.ancestral_labeling(), pt.ancestral_labeling()):
        assert np.array_equal(a, b)
    # a short climb: same moves, same state after each
    for _ in range(5):
        mv = rt.greedy_nni(eu, ev, order, False)
        assert pt.greedy_nni(eu, ev, order, want_all=False)["move"] == mv
        if mv is None:
            break
        rt.apply_nni(*mv)
        assert rt.small_rectilinear() == pt.apply_nni(*mv)
        eu, ev = rt.edges()
    rt.small_rectilinear()  # a failed greedy_nni leaves stale un-visited path intervals behind
    S, E, sc = rt.state()
    assert np.array_equal(S, pt.start) and np.array_equal(E, pt.end) and np.array_equal(sc, pt.score)


def test_non_binary_is_an_error(port, ref):
    c0 = np.array([1, 3, -1, -1], np.int32)
    c1 = np.array([2, -1, -1, -1], np.int32)
    leaf_bp = np.zeros((4, 5), np.int32)
    with pytest.raises(ValueError):
        ref.tree(c0, c1, leaf_bp).small_rectilinear()
    with pytest.raises(ValueError):
        PortTree(port, c0, c1, leaf_bp).small_rectilinear()
