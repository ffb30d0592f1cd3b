#!/usr/bin/env python
"""Times the distance kernel (device-resident, CUDA events) for LZG_SEG_COST values (fixed cost of a stream-K segment in the schedule).
usage: python tools/sweep_distance.py 0 1 2 12 4      (experiments; not part of the product)"""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

CHILD = r'''
import sys, json, numpy as np, torch
sys.path.insert(0, %r)
from lazac_copy_number_b200 import synth
from lazac_copy_number_b200.api import Context, Profiles
out = {}
with Context(0) as ctx:
    for name, (n, L) in {"cfg2": (1000, 10000), "mid": (6000, 4096), "n3000": (3000, 10000), "odd": (777, 1111)}.items():
        bp = synth.random_bp(n, L, seed=1, lo=-2, hi=2, density=0.01).astype(np.int8)
        gp = Profiles(ctx, bp)
        d = torch.empty((n, n), dtype=torch.int32, device="cuda")
        ts = []
        for k in range(8):
            ctx.flush_l2()
            ctx.timer_start()
            gp.distance_device_async(d.data_ptr())
            ts.append(ctx.timer_stop())
        ms = float(np.median(ts[3:]))
        P = n * (n + 1) // 2
        out[name] = {"ms": round(ms, 4), "frac_sad_peak": round(P * L / 4 / (ms * 1e-3) / 1.8304e13, 4)}
        chk = int(d[n // 2].sum().item()) == int(np.abs(bp.astype(np.int32) - bp[n // 2].astype(np.int32)).sum())
        out[name]["row_ok"] = chk
        gp.free()
print(json.dumps(out))
''' % ROOT

for v in sys.argv[1:] or ["0"]:
    env = dict(os.environ, LZG_SEG_COST=v)
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    print("seg_cost", v, r.stdout.strip().splitlines()[-1] if r.returncode == 0 and r.stdout.strip() else "FAILED " + r.stderr[-400:], flush=True)
