#!/usr/bin/env python
"""Times lzg_nj_seed_tree (GPU, Clearcut-exact NJ) on L1 matrices of random profiles; for the first
size also the host restatement through `lazac nj`.  usage: python tools/time_nj.py [n ...]"""
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles  # noqa: E402

sizes = [int(x) for x in sys.argv[1:]] or [2000, 10000]
with Context(0) as ctx:
    for k, n in enumerate(sizes):
        if os.environ.get("NJ_RANDOM_PROFILES"):  # seconds instead of a minute of tree simulation in Python
            bp = (np.random.default_rng(0).integers(0, 40, (n, 4000)) == 0).astype(np.int8) * np.random.default_rng(1).integers(-2, 3, (n, 4000)).astype(np.int8)
        else:
            prof = synth.make_profiles(n, 2000, 2, seed=0)
            bp = synth.bp_from_cn(prof["cn"], prof["seg"]).astype(np.int8)
        gp = Profiles(ctx, bp)
        D = gp.distance_matrix()
        perm = np.random.default_rng(1).permutation(n).astype(np.int32)
        ctx.nj_seed_tree(D[:64, :64].copy(), np.arange(64, dtype=np.int32))  # warm-up
        t0 = time.perf_counter()
        left, right = ctx.nj_seed_tree(D, perm)
        dt = time.perf_counter() - t0
        print("n = %d: GPU NJ %.3f s (%.1f G matrix elements scanned/s incl. upload)" % (n, dt, n ** 3 / 6 / dt / 1e9), flush=True)
        t0 = time.perf_counter()
        l2, r2 = gp.nj_seed_tree(perm)
        dt = time.perf_counter() - t0
        print("   the matrix already on the device (lzg_nj_seed_tree_profiles): %.3f s, same tree: %s" % (dt, bool((l2 == left).all() and (r2 == right).all())), flush=True)
        gp.free()
        if k == 0 and not os.environ.get("NJ_SKIP_CLI"):
            lazac = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "host", "lazac")
            with tempfile.TemporaryDirectory() as d:
                src = os.path.join(d, "m.txt")
                with open(src, "w") as f:
                    f.write("%d\n" % n)
                    for i in range(n):
                        f.write("c%05d " % i + " ".join(map(str, D[i].tolist())) + "\n")
                for mode in ("host", "gpu"):
                    t0 = time.perf_counter()
                    subprocess.run([lazac, "nj", src, "-s", "1", "-o", os.path.join(d, mode)], check=True, env=dict(os.environ, LAZAC_NJ=mode))
                    print("  lazac nj (%s, incl. parsing the text matrix): %.2f s" % (mode, time.perf_counter() - t0), flush=True)
                same = open(os.path.join(d, "host"), "rb").read() == open(os.path.join(d, "gpu"), "rb").read()
                print("  identical Newick:", same)
