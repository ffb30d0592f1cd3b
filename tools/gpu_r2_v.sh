#!/bin/bash
# round 2, call V: k_tc_distance experiments -- parity, then device-resident cfg3 timing per variant
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_distance_tc.py tests/test_gpu_multi.py -x -q ) > gpurun_out/r02v_pytest.log 2>&1; echo "tc rc=$?"; tail -2 gpurun_out/r02v_pytest.log
echo default; LZG_TC_TIMING=1 timeout 600 python tools/prof_distance.py 20000 10000 2 4 device 2>&1 | tail -1
echo nostore; LZG_TC_NOSTORE=1 LZG_TC_TIMING=1 timeout 600 python tools/prof_distance.py 20000 10000 2 4 device 2>&1 | tail -1
