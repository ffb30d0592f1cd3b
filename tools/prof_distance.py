#!/usr/bin/env python
"""Short driver for ncu captures: uploads a synthetic workload and runs the distance kernel a few
times.  usage: python tools/prof_distance.py [cells bins alleles reps [device]]
With "device" the matrix stays in HBM (lzg_distance_device: ONE launch per matrix, what bench.py's
`value` times); without it lzg_distance_matrix runs (row groups + D2H for large matrices)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles  # noqa: E402

device = len(sys.argv) > 5 and sys.argv[5] == "device"
cells, bins, alleles, reps = (int(x) for x in (sys.argv[1:5] + ["1000", "5000", "2", "5"][len(sys.argv[1:5]):]))
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = np.empty(prof["cn"].shape, np.int8)
for r0 in range(0, cells, 2000):
    bp[r0:r0 + 2000] = synth.bp_from_cn(prof["cn"][r0:r0 + 2000], prof["seg"]).astype(np.int8)
del prof
with Context(0) as ctx:
    gp = Profiles(ctx, bp)
    if device:
        import torch
        d_D = torch.empty((cells, cells), dtype=torch.int32, device="cuda")
        times = []
        for _ in range(reps):
            ctx.flush_l2()
            ctx.timer_start()
            gp.distance_device_async(d_D.data_ptr(), 0, 1)
            times.append(ctx.timer_stop())
        ctx.sync()
        print("ms per call (device-resident):", ["%.3f" % t for t in times], "checksum", int(d_D.sum(dtype=torch.int64)), gp.distance_path())
        gp.free()
        sys.exit(0)
    D = ctx.pinned((cells, cells), np.int32)
    times = []
    for _ in range(reps):
        ctx.flush_l2()
        ctx.timer_start()
        gp.distance_matrix(D)
        times.append(ctx.timer_stop())
    print("ms per call (kernel + D2H):", ["%.3f" % t for t in times], "checksum", int(D.sum()))
    gp.free()
