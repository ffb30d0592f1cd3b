#!/usr/bin/env python
"""Short driver for ncu captures of the NNI pass: random tree over synthetic leaves, a few
neighbourhood passes (UP intervals invalidated before each, as after an accepted move).
usage: python tools/prof_nni.py [cells bins alleles reps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import nj, synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles, Tree  # noqa: E402

cells, bins, alleles, reps = (int(x) for x in (sys.argv[1:5] + ["1000", "5000", "1", "3"][len(sys.argv) - 1:]))
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = synth.bp_from_cn(prof["cn"], prof["seg"]).astype(np.int8)
c0, c1 = synth.random_tree(cells, np.random.default_rng(1))
n = 2 * cells - 1
leaf_row = np.full(n, -1, np.int32)
leaf_row[cells - 1:] = np.arange(cells)
with Context(0) as ctx:
    gp = Profiles(ctx, bp)
    t = Tree(ctx, gp, c0, c1, leaf_row)
    root = t.small_rectilinear()
    eu, ev = t.edges()
    order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = t.enumerate_moves(eu, ev, order)
    dep = nj.depths(c0, c1)
    # T1 = T0 + one accepted move; the inverse of nni(u, w, v, z) is nni(u, z, v, w)
    m = tuple(int(x) for x in moves[len(moves) // 2])
    inv = (m[0], m[3], m[2], m[1])
    t.nni_best(moves)
    t.apply_nni(*m)
    eu1, ev1 = t.edges()
    moves1 = t.enumerate_moves(eu1, ev1, order)
    t.nni_best(moves1)
    t.apply_nni(*inv)
    t.nni_best(moves)
    times, applies = [], []
    for k in range(2 * reps):  # a hill-climb iteration: accept a move, score the new neighbourhood
        ctx.flush_l2()
        ctx.timer_start()
        t.apply_nni(*(m if k % 2 == 0 else inv))
        applies.append(ctx.timer_stop())
        ctx.timer_start()
        idx, score, _ = t.nni_best(moves1 if k % 2 == 0 else moves)
        times.append(ctx.timer_stop())
    cb, ub = t.last_nni_traffic()
    hot = []
    for _ in range(reps):  # UP intervals valid: scoring only
        ctx.flush_l2()
        ctx.timer_start()
        t.nni_best(moves)
        hot.append(ctx.timer_stop())
    print("nodes %d L %d candidates %d root %d best %d/%d mean depth %.1f" % (n, bp.shape[1], len(moves), root, idx, score, dep[moves[:, 2]].mean()))
    print("ms apply:", ["%.3f" % x for x in applies])
    print("ms pass after apply:", ["%.3f" % x for x in times], " ms scoring only:", ["%.3f" % x for x in hot])
    it = min(a + b for a, b in zip(applies, times))
    print("best iteration (apply + pass) %.3f ms; algorithmic bytes %d -> %.1f GB/s" % (it, cb + ub, (cb + ub) / (it * 1e-3) / 1e9))
    t.free()
    gp.free()
