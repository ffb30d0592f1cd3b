#!/usr/bin/env python
"""lzg_distance_matrix_multi end to end (host int8 profiles -> whole host matrix) on 1 .. G GPUs of the box.
usage: python tools/prof_multi.py cells bins alleles"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402  (device count only)

from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200.api import Context, MultiContext, Profiles  # noqa: E402

cells, bins, alleles = (int(x) for x in sys.argv[1:4])
prof = synth.make_profiles(cells, bins, alleles, seed=0)
bp = np.empty(prof["cn"].shape, np.int8)
for r0 in range(0, cells, 2000):
    bp[r0:r0 + 2000] = synth.bp_from_cn(prof["cn"][r0:r0 + 2000], prof["seg"]).astype(np.int8)
n, L = bp.shape
G = torch.cuda.device_count()
ref = None
with Context(0) as ctx:
    h_bp = ctx.pinned(bp.shape, np.int8)
    h_bp[...] = bp
    h_D = ctx.pinned((n, n), np.int32)
    for rep in range(3):
        t0 = time.perf_counter()
        p = Profiles(ctx, h_bp)
        p.distance_matrix(h_D)
        p.free()
        dt = time.perf_counter() - t0
    print("1 GPU, lzg_profiles_upload_i8 + lzg_distance_matrix: %.2f ms" % (1e3 * dt))
    ref = h_D.copy()
    for g in sorted({1, 2, 4, 8, G} & set(range(1, G + 1))):
        with MultiContext(list(range(g))) as m:
            ts = []
            for rep in range(5):
                h_D[...] = -1 if rep == 0 else h_D
                t0 = time.perf_counter()
                m.distance_matrix(h_bp, out=h_D)
                ts.append(time.perf_counter() - t0)
            ok = np.array_equal(h_D, ref)
            print("%d GPU(s), lzg_distance_matrix_multi_i8: %s ms, best %.2f ms, equal to the 1-GPU matrix: %s" % (g, " ".join("%.2f" % (1e3 * x) for x in ts), 1e3 * min(ts), ok))
