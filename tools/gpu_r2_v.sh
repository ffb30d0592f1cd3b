#!/bin/bash
# round 2, call V: both tensor-core kernels after a change -- parity (pairs, then single CTAs), device-resident cfg3 timing
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_distance_tc.py tests/test_gpu_multi.py -x -q ) > gpurun_out/r02v_pytest.log 2>&1; echo "pairs rc=$?"; tail -1 gpurun_out/r02v_pytest.log
( LZG_TC_2CTA=0 timeout 600 python -m pytest tests/test_gpu_distance_tc.py tests/test_gpu_multi.py -x -q ) > gpurun_out/r02v_pytest1.log 2>&1; echo "single rc=$?"; tail -1 gpurun_out/r02v_pytest1.log
echo pairs; LZG_TC_TIMING=1 timeout 300 python tools/prof_distance.py 20000 10000 2 3 device 2>&1 | tail -1
echo single; LZG_TC_2CTA=0 LZG_TC_TIMING=1 timeout 300 python tools/prof_distance.py 20000 10000 2 3 device 2>&1 | tail -1
