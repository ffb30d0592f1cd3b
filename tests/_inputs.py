"""Seeded inputs of BASELINE.json's shapes, generated once per test session (the generator walks a
random tree in Python: tens of seconds at 20,000 cells)."""
import functools

import numpy as np

from lazac_copy_number_b200 import nj, search, synth


def bp_int8(cn, seg, rows=2000):
    out = np.empty(cn.shape, np.int8)
    for r0 in range(0, cn.shape[0], rows):
        out[r0:r0 + rows] = synth.bp_from_cn(cn[r0:r0 + rows], seg).astype(np.int8)
    return out


@functools.lru_cache(maxsize=None)
def config_bp(cells, bins, alleles, seed=0):
    """int8 breakpoint matrix [cells][bins * alleles] of the synthetic generator."""
    prof = synth.make_profiles(cells, bins, alleles, seed=seed)
    bp = bp_int8(prof["cn"], prof["seg"])
    bp.setflags(write=False)
    return bp


def nj_seed_tree(ctx, gp, n_leaves):
    """The seed tree bench.py's NNI sections start from: the GPU distance matrix of the profiles, Clearcut-exact NJ
    on the GPU (lzg_nj_seed_tree) with the taxon order of NJ_shuffle_distance_matrix for seed 0.
    Returns (child0, child1, leaf_row, D)."""
    D = gp.distance_matrix()
    perm = search.clearcut_shuffle(n_leaves, 0)
    left, right = ctx.nj_seed_tree(D, perm)
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
    return c0, c1, leaf_row, D
