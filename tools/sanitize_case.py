#!/usr/bin/env python
"""Small invocations of every kernel family for compute-sanitizer (tools/gpu_sanitize.sh):
the smoke shape, a deep caterpillar tree (long heavy path + many light heads: the flag protocol of
k_up_pass, the last-CTA tail of k_apply_move), a distance matrix whose tiles are cut across CTAs
(atomic epilogue), the NJ seed tree, and a short device-side hill climb.  No torch import: the
sanitizer instruments every kernel of the process."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from lazac_copy_number_b200 import nj, synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles, Tree  # noqa: E402


def caterpillar(n_leaves):
    """root 0; internal nodes 0..n-2 in a chain, every one with a leaf hanging off; the last holds two leaves."""
    n = 2 * n_leaves - 1
    c0, c1 = np.full(n, -1, np.int32), np.full(n, -1, np.int32)
    for k in range(n_leaves - 1):
        leaf = n_leaves - 1 + k
        nxt = k + 1 if k + 1 < n_leaves - 1 else n - 1
        c0[k], c1[k] = min(nxt, leaf), max(nxt, leaf)
    leaf_row = np.full(n, -1, np.int32)
    leaf_row[n_leaves - 1:] = np.arange(n_leaves)
    return c0, c1, leaf_row


def neighbourhood(t, seed=0):
    eu, ev = t.edges()
    order = np.random.default_rng(seed).permutation(len(eu)).astype(np.int32)
    return t.enumerate_moves(eu, ev, order), order


def main():
    with Context(0) as ctx:
        # 1. smoke shape: distance (one 64 x 64 tile), full scoring, pass, move, labels
        bp = synth.random_bp(48, 800, seed=1, density=0.3)
        gp = Profiles(ctx, bp)
        D = gp.distance_matrix()
        assert np.array_equal(D[0], np.abs(bp - bp[0]).sum(1))
        c0, c1 = synth.random_tree(48, np.random.default_rng(2))
        leaf_row = np.full(95, -1, np.int32)
        leaf_row[47:] = np.arange(48)
        t = Tree(ctx, gp, c0, c1, leaf_row)
        root = t.small_rectilinear()
        moves, order = neighbourhood(t)
        idx, score, allsc = t.nni_best(moves, want_all=True)
        if idx >= 0:
            assert t.apply_nni(*moves[idx]) == score
            moves, order = neighbourhood(t)
            t.nni_best(moves)
        t.ancestral_labeling()
        if hasattr(t, "hill_climb"):
            t.hill_climb(order)
        t.free()
        # 2. deep caterpillar, perturbed so that light heads hang off the long path
        n_leaves = 300
        bp2 = synth.random_bp(n_leaves, 600, seed=3, density=0.2)
        gp2 = Profiles(ctx, bp2)
        c0, c1, leaf_row = caterpillar(n_leaves)
        c0, c1 = synth.perturb_tree(c0, c1, 200, np.random.default_rng(4))
        t2 = Tree(ctx, gp2, c0, c1, leaf_row)
        t2.small_rectilinear()
        for it in range(3):
            moves, order = neighbourhood(t2, it)
            idx, score, _ = t2.nni_best(moves)
            if idx < 0:
                break
            assert t2.apply_nni(*moves[idx]) == score
        if hasattr(t2, "hill_climb"):
            t2.hill_climb(order)
        t2.free()
        # 3. distance with tiles cut across CTAs (6 tiles over 148 persistent CTAs: atomic epilogues)
        bp3 = synth.random_bp(300, 5000, seed=5, density=0.2)
        gp3 = Profiles(ctx, bp3)
        D3 = gp3.distance_matrix()
        assert np.array_equal(D3[7], np.abs(bp3 - bp3[7]).sum(1)) and np.array_equal(D3, D3.T)
        # 4. NJ seed tree on it
        perm = np.random.default_rng(6).permutation(300).astype(np.int32)
        ctx.nj_seed_tree(D3, perm)
        gp3.free()
        gp2.free()
        gp.free()
    print("sanitize_case ok")


if __name__ == "__main__":
    main()
