#!/bin/bash
# round 2, call O: NJ entry points (host / device / profiles) and the CLI that uses them; timing through each
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_nj.py tests/test_gpu_host_cli.py tests/test_gpu_distance.py tests/test_zz_no_alloc.py -x -q ) > gpurun_out/r02o_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02o_pytest.log
NJ_SKIP_CLI=1 NJ_RANDOM_PROFILES=1 timeout 600 python tools/time_nj.py 10000 10000 5000 2000 > gpurun_out/r02o_nj.log 2>&1; grep "^n =\|device" gpurun_out/r02o_nj.log
