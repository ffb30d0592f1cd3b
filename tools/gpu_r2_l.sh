#!/bin/bash
# round 2, call L: NJ timing only (fused against two-barrier form, with the per-CTA phase timers)
mkdir -p gpurun_out
NJ_SKIP_CLI=1 NJ_RANDOM_PROFILES=1 LZG_NJ_DEBUG=1 timeout 600 python tools/time_nj.py 10000 > gpurun_out/r02l_nj.log 2>&1; grep -v "n 64\|: 0\.[0-9] ms\|n 64:" gpurun_out/r02l_nj.log
