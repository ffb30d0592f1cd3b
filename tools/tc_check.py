#!/usr/bin/env python
"""Tensor-core distance (lzg_distance_tc.cu) against k_zcnt_distance (LZG_TC=0) on the same profiles, then timing.
usage: python tools/tc_check.py [n L ...]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lazac_copy_number_b200 import synth  # noqa: E402
from lazac_copy_number_b200.api import Context, Profiles  # noqa: E402

shapes = [int(x) for x in sys.argv[1:]] or [256, 300, 700, 1000, 1500, 3000]
shapes = list(zip(shapes[0::2], shapes[1::2]))
with Context(0) as ctx:
    for n, L in shapes:
        bp = synth.random_bp(n, L, seed=n + L, lo=-3, hi=3, density=0.05).astype(np.int8)
        os.environ["LZG_TC"] = "0"
        p = Profiles(ctx, bp)
        want = p.distance_matrix()
        p.free()
        os.environ["LZG_TC"] = "1"
        p = Profiles(ctx, bp)
        got = p.distance_matrix()
        same = bool(np.array_equal(got, want))
        bad = int((got != want).sum())
        print("n %d L %d: equal %s (%d of %d differ) %s" % (n, L, same, bad, n * n, p.distance_path()), flush=True)
        if not same:
            ii, jj = np.nonzero(got != want)
            print("   first:", [(int(a), int(b), int(got[a, b]), int(want[a, b])) for a, b in list(zip(ii, jj))[:8]])
            print("   rows differing:", np.unique(ii)[:20], "cols:", np.unique(jj)[:20])
        ts = []
        for _ in range(3):
            ctx.sync()
            t0 = time.perf_counter()
            p.distance_matrix()
            ts.append(time.perf_counter() - t0)
        p.free()
        import torch
        d = torch.empty((n, n), dtype=torch.int32, device="cuda:0")
        for tag, env in (("tensor core", "1"), ("VABSDIFF4", "0")):
            os.environ["LZG_TC"] = env
            q = Profiles(ctx, bp)
            q.distance_device_async(d.data_ptr())
            ctx.sync()
            ks = []
            for _ in range(5):
                ctx.timer_start()
                q.distance_device_async(d.data_ptr())
                ks.append(ctx.timer_stop())
            q.free()
            print("   %s: device-resident %.3f ms (best of 5)" % (tag, min(ks)), flush=True)
        print("   host matrix best %.2f ms" % (1e3 * min(ts)), flush=True)
