"""GPU parity at the shapes the headline numbers are taken on, against the reference ITSELF.

Every comparison here drives oracle/_ref/libref_shim.so (the unmodified functions of
/root/reference/src/copy_number.cxx compiled in place, oracle/Makefile) where it exists, and the C
restatement (oracle/zcnt_oracle.c) beside it:
  * cfg2's own shape (1,000 x 10,000): the whole matrix against the port's pair loop, 64 random rows
    against the reference's operator- / breakpoint_magnitude (src/lazac.cxx:81-91);
  * cfg3 (20,000 x 20,000): 32 random rows plus the first and last row of 8 random tiles against
    the port, 4 of them against the reference;
  * cfg5 (10,000 cells x 20,000 positions) ON THE NJ SEED TREE bench.py measures: every candidate
    score of the neighbourhood, the winner and the greedy winner against the reference's candidate
    loop (src/copy_number.cxx:347-366), then the accepted move's scores and intervals.
LZG_PIN_CANDIDATES=k limits the cfg5 comparison to a k-candidate slice that contains the winner
and the deepest v (the full neighbourhood is about 40 s of the reference on one core and 3.2 GB).
"""
import os

import numpy as np
import pytest

from _inputs import config_bp, nj_seed_tree
from lazac_copy_number_b200 import nj
from lazac_copy_number_b200.api import Profiles, Tree

pytestmark = pytest.mark.gpu


def test_cfg2_shape_whole_matrix(ctx, port, ref):
    """BASELINE.json configs[1]: 1,000 cells x 5,000 bins x 2 alleles."""
    bp = config_bp(1000, 5000, 2)
    n = bp.shape[0]
    gp = Profiles(ctx, bp)
    D = gp.distance_matrix()
    gp.free()
    assert np.array_equal(D, port.distance_matrix(bp.astype(np.int32)))      # the whole matrix, pair by pair
    rows = np.sort(np.random.default_rng(2).choice(n, 64, replace=False)).astype(np.int32)
    assert np.array_equal(D[rows], ref.distance_rows_i8(bp, rows))           # the reference's own arithmetic


def test_cfg3_rows_against_reference(ctx, port, ref):
    """BASELINE.json configs[2]: 20,000 cells x 10,000 bins x 2 alleles on one GPU."""
    bp = config_bp(20000, 10000, 2)
    n = bp.shape[0]
    gp = Profiles(ctx, bp)
    D = ctx.pinned((n, n), np.int32)
    gp.distance_matrix(D)
    gp.free()
    rng = np.random.default_rng(3)
    rows = list(rng.choice(n, 32, replace=False))
    for t in rng.choice((n + 127) // 128, 8, replace=False):                 # tile edges: first and last row of 8 tiles
        rows += [int(t) * 128, min(n, (int(t) + 1) * 128) - 1]
    rows = np.array(sorted(set(int(r) for r in rows)), np.int32)
    want = port.distance_rows_i8(bp, rows)
    assert np.array_equal(D[rows], want)
    assert np.array_equal(D[:, rows].T, want)                                 # the mirrored halves
    few = rows[[0, len(rows) // 3, 2 * len(rows) // 3, len(rows) - 1]]
    assert np.array_equal(D[few], ref.distance_rows_i8(bp, few))


def test_cfg5_nj_seed_tree_neighbourhood_against_reference(ctx, ref):
    """BASELINE.json configs[4]'s shape on the tree bench.py's "nni_large" section uses."""
    n_leaves = 10000
    bp = config_bp(n_leaves, 10000, 2)
    gp = Profiles(ctx, bp)
    c0, c1, leaf_row, _ = nj_seed_tree(ctx, gp, n_leaves)
    gt = Tree(ctx, gp, c0, c1, leaf_row)
    root = gt.small_rectilinear()
    eu, ev = gt.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = gt.enumerate_moves(eu, ev, order)
    assert len(moves) == 2 * (n_leaves - 2)
    idx, score, allsc = gt.nni_best(moves, want_all=True)
    gidx, gscore, _ = gt.nni_best(moves, greedy=True)

    rt = ref.tree_i8(c0, c1, leaf_row, bp)
    assert rt.small_rectilinear() == root
    limit = int(os.environ.get("LZG_PIN_CANDIDATES", "0"))
    if limit and limit < len(moves):
        # a slice of the enumeration (whole edges, in enumeration order) that contains the winner and the deepest v
        dep = nj.depths(c0, c1)
        internal = [int(e) for e in order if c0[ev[e]] >= 0]                  # edge indices that yield candidates
        pos_of_edge = {e: k for k, e in enumerate(internal)}                   # candidates 2k, 2k + 1
        deepest = max(internal, key=lambda e: dep[ev[e]])
        keep = set(int(e) for e in np.random.default_rng(5).choice(internal, limit // 2, replace=False))
        keep.add(deepest)
        if idx >= 0:
            keep.add(internal[idx // 2])
        if gidx >= 0:
            keep.add(internal[gidx // 2])
        sub = np.array([e for e in internal if e in keep], np.int32)
        ref_moves, ref_scores = rt.candidate_scores(eu, ev, sub)
        pick = np.array([2 * pos_of_edge[int(e)] + z for e in sub for z in (0, 1)])
        assert np.array_equal(ref_moves, moves[pick])
        assert np.array_equal(ref_scores, allsc[pick])
    else:
        ref_moves, ref_scores = rt.candidate_scores(eu, ev, order)
        assert np.array_equal(ref_moves, moves)
        assert np.array_equal(ref_scores, allsc)
        better = np.flatnonzero(ref_scores < root)
        assert idx == (better[np.argmin(ref_scores[better])] if len(better) else -1)
        assert gidx == (better[0] if len(better) else -1)
    if idx >= 0:
        assert score == allsc[idx] and gscore == allsc[gidx]
        # accept the winner on both sides: every cumulative score, and the intervals along the re-scored path
        u, w, v, z = (int(x) for x in moves[idx])
        assert gt.apply_nni(u, w, v, z) == score
        rt.apply_nni(u, w, v, z)
        assert np.array_equal(gt.scores(), rt.scores())
        t0, t1 = gt.topology()
        r0, r1 = rt.topology()
        assert np.array_equal(t0, r0) and np.array_equal(t1, r1)
        parent = np.full(len(t0), -1)
        for a in range(len(t0)):
            if t0[a] >= 0:
                parent[t0[a]] = parent[t1[a]] = a
        path = []
        a = v
        while a >= 0:
            path.append(a)
            a = int(parent[a])
        S, E = gt.node_intervals(path)
        for k, a in enumerate(path):
            s, e, _, _ = rt.node(a)
            assert np.array_equal(S[k], s) and np.array_equal(E[k], e), a
    gt.free()
    gp.free()
