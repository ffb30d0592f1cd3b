#!/bin/bash
# round 2, call G: 16-bit device format + restructured k_apply2 -- whole GPU suite, climb timing
mkdir -p gpurun_out
export LZG_PIN_CANDIDATES=${LZG_PIN_CANDIDATES:-2000}
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=6 ) > gpurun_out/r02g_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/r02g_pytest_gpu.log
timeout 600 python tools/prof_climb.py 1000 5000 1 > gpurun_out/r02g_climb_cfg4.log 2>&1; cat gpurun_out/r02g_climb_cfg4.log
timeout 900 python tools/prof_climb.py 10000 10000 2 > gpurun_out/r02g_climb_cfg5.log 2>&1; cat gpurun_out/r02g_climb_cfg5.log
