import os, sys, time
sys.path.insert(0, "/root/repo")
import numpy as np
from lazac_copy_number_b200 import synth
from lazac_copy_number_b200.api import Context, Profiles, Tree
cells, bins, alleles = 10000, 10000, 2
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = synth.bp_from_cn(prof["cn"], prof["seg"]).astype(np.int8)
c0, c1 = synth.random_tree(cells, np.random.default_rng(1))
n = 2 * cells - 1
leaf_row = np.full(n, -1, np.int32); leaf_row[cells - 1:] = np.arange(cells)
with Context(0) as ctx:
    gp = Profiles(ctx, bp); t = Tree(ctx, gp, c0, c1, leaf_row); t.small_rectilinear()
    eu, ev = t.edges(); order = np.random.default_rng(0).permutation(len(eu)).astype(np.int32)
    moves = t.enumerate_moves(eu, ev, order)
    m = tuple(int(x) for x in moves[len(moves) // 2]); inv = (m[0], m[3], m[2], m[1])
    t.nni_best(moves); t.apply_nni(*m); e1 = t.edges(); moves1 = t.enumerate_moves(e1[0], e1[1], order); t.apply_nni(*inv)
    for k in range(8):
        ctx.flush_l2(); ctx.sync()
        w0 = time.perf_counter(); ctx.timer_start()
        t.apply_nni(*(m if k % 2 == 0 else inv)); w1 = time.perf_counter()
        t.nni_best(moves1 if k % 2 == 0 else moves)
        ms = ctx.timer_stop(); w2 = time.perf_counter()
        print("iteration %.3f ms device, wall apply %.0f us pass %.0f us" % (ms, 1e6*(w1-w0), 1e6*(w2-w1)), flush=True)
