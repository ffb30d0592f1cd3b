#!/usr/bin/env python
"""The super-block tile order of the distance schedule (LZG_DIST_SB, csrc/lzg_distance.cu) against
the row-major order: same matrix bit for bit (whole matrix in one launch, and the union of 8 shards),
and the time per launch of each.  usage: python tools/check_dist_order.py [cells bins alleles]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles  # noqa: E402

cells, bins, alleles = (int(x) for x in (sys.argv[1:4] + ["20000", "10000", "2"][len(sys.argv[1:4]):]))
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = np.empty(prof["cn"].shape, np.int8)
for r0 in range(0, cells, 2000):
    bp[r0:r0 + 2000] = synth.bp_from_cn(prof["cn"][r0:r0 + 2000], prof["seg"]).astype(np.int8)
del prof


def run(sb, n_shards, align=1):
    os.environ["LZG_DIST_SB"] = str(sb)
    os.environ["LZG_DIST_ALIGN"] = str(align)
    with Context(0) as ctx:   # a fresh context: schedules are cached per context
        gp = Profiles(ctx, bp)
        d = torch.zeros((cells, cells), dtype=torch.int32, device="cuda")
        ms = []
        for rep in range(3):
            tot = 0.0
            for sh in range(n_shards):
                ctx.flush_l2()
                ctx.timer_start()
                gp.distance_device_async(d.data_ptr(), sh, n_shards)
                tot += ctx.timer_stop()
            ms.append(tot)
        ctx.sync()
        gp.free()
    return d, min(ms)


ref, t0 = run(0, 1)
print("row-major      1 shard : %.3f ms" % t0, flush=True)
for sb, al in ((12, 0), (12, 1), (8, 1), (16, 1)):
    d, t = run(sb, 1, al)
    print("super-block %2d align %d 1 shard : %.3f ms  equal=%s" % (sb, al, t, bool(torch.equal(d, ref))), flush=True)
    del d
d, t = run(0, 8)
print("row-major      8 shards: %.3f ms (sum)  equal=%s" % (t, bool(torch.equal(d, ref))), flush=True)
del d
d, t = run(12, 8)
print("super-block 12 8 shards: %.3f ms (sum)  equal=%s" % (t, bool(torch.equal(d, ref))), flush=True)
