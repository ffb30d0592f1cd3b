"""GPU parity at BASELINE.json's full sizes.  cfg4 (1,000 cells x 5,000 positions) is small enough for
the oracle to score a whole neighbourhood in seconds, so it is compared candidate by candidate; at
cfg5's and cfg3's shapes the oracle would take minutes to hours, so the checks are the
size-independent properties of the path:
  * a candidate's score (three merges around the edge, UP intervals) equals the root score after
    the move is really applied (path re-scoring by another kernel) and after a full bottom-up
    rescoring from the leaves (a third kernel);
  * a move followed by its inverse restores score and neighbourhood scores;
  * the scores of column slices add up to the scores of the whole profiles (position sharding);
  * the distance matrix is symmetric with a zero diagonal and its rows equal the row-wise L1 sums.
"""
import numpy as np
import pytest

from _inputs import config_bp
from lazac_copy_number_b200 import sharding, synth
from lazac_copy_number_b200.api import Profiles, Tree
from oracle import PortTree

pytestmark = pytest.mark.gpu


def tree_arrays(n_leaves, seed):
    c0, c1 = synth.random_tree(n_leaves, np.random.default_rng(seed))
    n = 2 * n_leaves - 1
    leaf_row = np.full(n, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    return c0, c1, leaf_row


def test_cfg4_neighbourhood_equals_oracle(ctx, port):
    """BASELINE.json configs[3]: 1,000 cells x 5,000 bins x 1 allele, 1,996 candidates."""
    n_leaves = 1000
    prof = synth.make_profiles(n_leaves, 5000, 1, seed=0)
    bp = synth.bp_from_cn(prof["cn"], prof["seg"])
    c0, c1, leaf_row = tree_arrays(n_leaves, 1)
    n = 2 * n_leaves - 1
    leaf_bp = np.zeros((n, bp.shape[1]), np.int32)
    leaf_bp[n_leaves - 1:] = bp
    gp = Profiles(ctx, bp)
    gt = Tree(ctx, gp, c0, c1, leaf_row)
    pt = PortTree(port, c0, c1, leaf_bp)
    assert gt.small_rectilinear() == pt.small_rectilinear()
    eu, ev = pt.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    full = pt.greedy_nni(eu, ev, order)
    moves = gt.enumerate_moves(eu, ev, order)
    assert len(moves) == 2 * (n_leaves - 2)
    idx, score, allsc = gt.nni_best(moves, want_all=True)
    assert np.array_equal(allsc, full["scores"])
    assert idx == full["index"] and (idx < 0 or score == full["scores"][idx])
    if idx >= 0:
        assert gt.apply_nni(*moves[idx]) == pt.apply_nni(*full["move"]) == score
        S, E, sc = gt.intervals()
        assert np.array_equal(sc, pt.score) and np.array_equal(S, pt.start) and np.array_equal(E, pt.end)
    gt.free()
    gp.free()


def test_cfg5_shape_properties(ctx):
    """BASELINE.json configs[4]'s shape on one GPU: 10,000 cells x 10,000 bins x 2 alleles (19,999 nodes,
    19,996 candidates, 0.8 GB of intervals + 0.8 GB of UP intervals)."""
    import torch
    n_leaves = 10000
    bp = config_bp(n_leaves, 10000, 2)
    L = bp.shape[1]
    c0, c1, leaf_row = tree_arrays(n_leaves, 1)
    gp = Profiles(ctx, bp)
    gt = Tree(ctx, gp, c0, c1, leaf_row)
    root = gt.small_rectilinear()
    eu, ev = gt.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = gt.enumerate_moves(eu, ev, order)
    assert len(moves) == 2 * (n_leaves - 2)
    idx, score, allsc = gt.nni_best(moves, want_all=True)
    better = np.flatnonzero(allsc < root)
    if len(better):   # the reference's choice: strictly smallest, first in enumeration order
        assert idx == better[np.argmin(allsc[better])] and score == allsc[idx]
    else:
        assert idx == -1 and score == root
    # the winner and five arbitrary candidates, each applied for real, rescored from the leaves, and undone
    picks = [int(idx)] if idx >= 0 else []
    picks += [int(x) for x in np.random.default_rng(3).choice(len(moves), 5, replace=False)]
    for k in picks:
        u, w, v, z = (int(x) for x in moves[k])
        assert gt.apply_nni(u, w, v, z) == allsc[k]                  # path re-scoring (k_apply_move)
        t0, t1 = gt.topology()
        gt.set_topology(t0, t1)
        assert gt.small_rectilinear() == allsc[k]                    # bottom-up from the leaves (k_merge_level)
        assert gt.apply_nni(u, z, v, w) == root                      # the inverse move
    idx2, score2, allsc2 = gt.nni_best(moves, want_all=True)
    assert idx2 == idx and score2 == score and np.array_equal(allsc2, allsc)
    # position sharding: the {root, scores} vectors of two column halves add up to the whole
    cuts = [sharding.position_slice(L, r, 2) for r in range(2)]
    halves = [Profiles(ctx, np.ascontiguousarray(bp[:, lo:hi])) for lo, hi in cuts]
    trees = [Tree(ctx, p, c0, c1, leaf_row) for p in halves]
    assert sum(t.small_rectilinear() for t in trees) == root
    vecs = []
    for t in trees:
        address, n_vec = t.nni_scores_launch(moves)
        vecs.append(sharding.device_vector(address, n_vec, ctx.device))
    with torch.cuda.stream(torch.cuda.ExternalStream(ctx.stream(), device=ctx.device)):
        total = torch.stack(vecs).sum(0, dtype=torch.int32)
        for vec in vecs:
            vec.copy_(total)
    for t in trees:
        bi, bs, sc = t.nni_scores_finish(False, want_all=True)
        assert bi == idx and bs == score and np.array_equal(sc, allsc)
        t.free()
    for p in halves:
        p.free()
    gt.free()
    gp.free()


def test_cfg3_shape_distance_properties(ctx):
    """BASELINE.json configs[2]'s input on one GPU: 20,000 cells x 20,000 positions, 2.0e8 pairs."""
    n = 20000
    bp = config_bp(n, 10000, 2)
    gp = Profiles(ctx, bp)
    D = ctx.pinned((n, n), np.int32)
    gp.distance_matrix(D)
    gp.free()
    assert not D.diagonal().any()
    for r0 in range(0, n, 2000):                                      # symmetry, block by block
        assert np.array_equal(D[r0:r0 + 2000], D[:, r0:r0 + 2000].T)
    wide = bp.astype(np.int16)
    for i in (0, 1, 127, 128, 9999, 10000, 16383, 19871, 19999):       # rows against the definition
        row = np.concatenate([np.abs(wide[c0:c0 + 2000] - wide[i]).sum(1, dtype=np.int64) for c0 in range(0, n, 2000)])
        assert np.array_equal(D[i], row)
