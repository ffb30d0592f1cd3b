#!/usr/bin/env python
"""Device-side hill climb from the NJ seed tree of a synthetic input: per-iteration device times
(%globaltimer stamps of every pass), wall time of the whole call.
usage: python tools/prof_climb.py cells bins alleles [max_moves]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import nj, synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles, Tree  # noqa: E402

cells, bins, alleles = (int(x) for x in sys.argv[1:4])
max_moves = int(sys.argv[4]) if len(sys.argv) > 4 else 2 ** 31 - 1
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = np.empty(prof["cn"].shape, np.int8)
for r0 in range(0, cells, 2000):
    bp[r0:r0 + 2000] = synth.bp_from_cn(prof["cn"][r0:r0 + 2000], prof["seg"]).astype(np.int8)
with Context(0) as ctx:
    gp = Profiles(ctx, bp)
    D = gp.distance_matrix()
    perm = np.random.default_rng(0).permutation(cells).astype(np.int32)
    left, right = ctx.nj_seed_tree(D, perm)
    c0, c1, leaf_row = nj.tree_from_seed(left, right, perm)
    t = Tree(ctx, gp, c0, c1, leaf_row)
    root = t.small_rectilinear()
    order = np.random.default_rng(0).permutation(len(c0) - 1).astype(np.int32)
    base = t.clone()
    for rep in range(3):
        w = base.clone()
        ctx.sync()
        t0 = time.perf_counter()
        r = w.hill_climb(order, max_moves=max_moves, want_moves=False)
        wall = time.perf_counter() - t0
        st = w.climb_stamps().astype(np.int64)
        it = np.diff(st) / 1e3
        print("rep %d: root %d -> %d (returned %d), %d moves, %d passes, reason %d, wall %.3f ms = %.1f us per pass"
              % (rep, root, r["true_score"], r["returned_score"], r["n_moves"], r["n_passes"], r["reason"], 1e3 * wall, 1e6 * wall / max(1, r["n_passes"])))
        if len(it):
            print("   iteration (device stamps): median %.1f us, mean %.1f, min %.1f, p90 %.1f, max %.1f; first ten: %s"
                  % (np.median(it), it.mean(), it.min(), np.percentile(it, 90), it.max(), " ".join("%.0f" % x for x in it[:10])))
        w.free()
    base.free()
    t.free()
    gp.free()
