#!/bin/bash
# round 2, call P: tensor-core distance -- its parity tests, the distance / multi / fullsize / pins suites on top of it, cfg3 timing
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_distance_tc.py -x -q ) > gpurun_out/r02p_pytest_tc.log 2>&1; echo "tc rc=$?"; tail -5 gpurun_out/r02p_pytest_tc.log
( timeout 1500 python -m pytest tests/test_gpu_distance.py tests/test_gpu_int16.py tests/test_gpu_multi.py tests/test_gpu_fullsize.py tests/test_gpu_reference_pins.py tests/test_zz_no_alloc.py -x -q -k "not cfg5" ) > gpurun_out/r02p_pytest_rest.log 2>&1; echo "rest rc=$?"; tail -5 gpurun_out/r02p_pytest_rest.log
timeout 600 python tools/prof_distance.py 20000 10000 2 5 device > gpurun_out/r02p_prof_cfg3.log 2>&1; tail -8 gpurun_out/r02p_prof_cfg3.log
LZG_TC=0 timeout 600 python tools/prof_distance.py 20000 10000 2 3 device > gpurun_out/r02p_prof_cfg3_alu.log 2>&1; tail -2 gpurun_out/r02p_prof_cfg3_alu.log
timeout 600 python tools/prof_distance.py 20000 10000 2 3 > gpurun_out/r02p_prof_cfg3_host.log 2>&1; tail -2 gpurun_out/r02p_prof_cfg3_host.log
