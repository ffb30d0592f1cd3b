#!/usr/bin/env python
"""Short driver for ncu captures: uploads a synthetic workload and runs the distance kernel a few
times.  usage: python tools/prof_distance.py [cells bins alleles reps]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles  # noqa: E402

cells, bins, alleles, reps = (int(x) for x in (sys.argv[1:5] + ["1000", "5000", "2", "5"][len(sys.argv) - 1:]))
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = synth.bp_from_cn(prof["cn"], prof["seg"]).astype(np.int8)
with Context(0) as ctx:
    gp = Profiles(ctx, bp)
    D = ctx.pinned((cells, cells), np.int32)
    times = []
    for _ in range(reps):
        ctx.flush_l2()
        ctx.timer_start()
        gp.distance_matrix(D)
        times.append(ctx.timer_stop())
    print("ms per call (kernel + D2H):", ["%.3f" % t for t in times], "checksum", int(D.sum()))
    gp.free()
