#!/bin/bash
# round 2, call K: NJ kernels -- parity, then timing of the forms (LZG_NJ_PERSISTENT = 3 default, 2 fused, 1 persistent, 0 per join)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_nj.py -x -q ) > gpurun_out/r02k_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02k_pytest.log
NJ_SKIP_CLI=1 NJ_RANDOM_PROFILES=1 LZG_NJ_PERSISTENT=1 LZG_NJ_DEBUG=1 timeout 600 python tools/time_nj.py 10000 > gpurun_out/r02k_nj.log 2>&1; grep -v "lazac nj\|identical\|n 64\|: 0\.[0-9] ms" gpurun_out/r02k_nj.log
for mode in 1 2; do echo "LZG_NJ_PERSISTENT=$mode"; NJ_SKIP_CLI=1 NJ_RANDOM_PROFILES=1 LZG_NJ_PERSISTENT=$mode timeout 600 python tools/time_nj.py 10000 10000 7000 5000 2000 1000 2>&1 | grep "^n ="; done | tee gpurun_out/r02k_nj_modes.log
