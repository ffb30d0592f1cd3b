#!/bin/bash
# round 2, call D (2 GPUs): multi-GPU distance from one process -- parity with one GPU, CLI --devices, timing at cfg3
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r02d_gpus.txt
( timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host_cli.py tests/test_gpu_distance.py -x -q --durations=5 ) > gpurun_out/r02d_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02d_pytest.log
timeout 600 python tools/prof_multi.py 20000 10000 2 > gpurun_out/r02d_multi_cfg3.log 2>&1; cat gpurun_out/r02d_multi_cfg3.log
