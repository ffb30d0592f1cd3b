#!/bin/bash
# round 2, call J: persistent NJ kernel -- parity, timing at n = 10,000 against the per-join launches
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_nj.py tests/test_gpu_host_cli.py -x -q --durations=5 ) > gpurun_out/r02j_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02j_pytest.log
LZG_NJ_DEBUG=1 timeout 600 python tools/time_nj.py > gpurun_out/r02j_nj_persistent.log 2>&1; cat gpurun_out/r02j_nj_persistent.log
LZG_NJ_PERSISTENT=0 timeout 600 python tools/time_nj.py > gpurun_out/r02j_nj_perjoin.log 2>&1; cat gpurun_out/r02j_nj_perjoin.log
