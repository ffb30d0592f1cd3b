#!/bin/bash
# round 2, call A: the reference pins + whole GPU suite, integer-pipe microbenchmark, sanitizers, ncu of the NNI kernels at cfg4
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.csv
free -g | head -2 >> gpurun_out/gpu.csv; nproc >> gpurun_out/gpu.csv
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=12 ) > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02a_pytest_gpu.log
bash tools/gpu_microbench.sh
sed -i 's/timeout 900/timeout 420/' tools/gpu_sanitize.sh; bash tools/gpu_sanitize.sh
timeout 200 python tools/prof_nni.py 1000 5000 1 3 > gpurun_out/r02a_prof_nni_cfg4_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_nni_items|k_up_pass|k_apply_move|k_nni_argmin' -s 8 -c 4 -f \
    -o gpurun_out/r02a_prof_nni_cfg4 python tools/prof_nni.py 1000 5000 1 3 > gpurun_out/r02a_prof_nni_cfg4_ncu.log 2>&1
echo "ncu nni cfg4 rc=$?"
cat gpurun_out/r02a_prof_nni_cfg4_plain.log
tail -25 gpurun_out/r02a_pytest_gpu.log
