#!/bin/bash
# round 2, call B: the device-side hill climb -- parity tests (graph and queued batches), CLI goldens, timing at cfg4 / cfg5
mkdir -p gpurun_out
export LZG_CLIMB_DEBUG=1
( time timeout 900 python -m pytest tests/test_gpu_climb.py -x -q --durations=5 ) > gpurun_out/r02b_pytest_climb.log 2>&1; echo "climb rc=$?" | tee -a gpurun_out/r02b_pytest_climb.log
tail -30 gpurun_out/r02b_pytest_climb.log
unset LZG_CLIMB_DEBUG
( time timeout 900 python -m pytest tests/test_gpu_host_cli.py tests/test_gpu_tree.py tests/test_gpu_golden.py -x -q --durations=5 ) > gpurun_out/r02b_pytest_cli.log 2>&1; echo "cli rc=$?" | tee -a gpurun_out/r02b_pytest_cli.log
tail -15 gpurun_out/r02b_pytest_cli.log
timeout 600 python tools/prof_climb.py 1000 5000 1 > gpurun_out/r02b_climb_cfg4.log 2>&1; echo "cfg4 rc=$?"; cat gpurun_out/r02b_climb_cfg4.log
LZG_CLIMB_GRAPH=0 timeout 600 python tools/prof_climb.py 1000 5000 1 > gpurun_out/r02b_climb_cfg4_batches.log 2>&1; echo "cfg4 batches rc=$?"; cat gpurun_out/r02b_climb_cfg4_batches.log
timeout 900 python tools/prof_climb.py 10000 10000 2 > gpurun_out/r02b_climb_cfg5.log 2>&1; echo "cfg5 rc=$?"; cat gpurun_out/r02b_climb_cfg5.log
