#!/bin/bash
# ncu --set full captures of the two dominant kernels (one gpurun call; plain run first each time).
# Usage (on the GPU box, from the repo root): bash tools/gpu_prof.sh
mkdir -p gpurun_out
timeout 300 python tools/prof_distance.py 1000 5000 2 3 > gpurun_out/prof_dist_cfg2_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_zcnt_distance -s 1 -c 2 -f \
    -o gpurun_out/prof_dist_cfg2 python tools/prof_distance.py 1000 5000 2 3 > gpurun_out/prof_dist_cfg2_ncu.log 2>&1
echo "ncu dist cfg2 rc=$?"
timeout 300 python tools/prof_distance.py 8000 10000 2 2 > gpurun_out/prof_dist_8k_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_zcnt_distance -s 1 -c 1 -f \
    -o gpurun_out/prof_dist_8k python tools/prof_distance.py 8000 10000 2 2 > gpurun_out/prof_dist_8k_ncu.log 2>&1
echo "ncu dist 8k rc=$?"
timeout 300 python tools/prof_nni.py 1000 5000 1 3 > gpurun_out/prof_nni_cfg4_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_nni_pass -s 1 -c 2 -f \
    -o gpurun_out/prof_nni_cfg4 python tools/prof_nni.py 1000 5000 1 3 > gpurun_out/prof_nni_cfg4_ncu.log 2>&1
echo "ncu nni cfg4 rc=$?"
timeout 300 python tools/prof_nni.py 4000 10000 2 2 > gpurun_out/prof_nni_4k_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_nni_pass -s 1 -c 1 -f \
    -o gpurun_out/prof_nni_4k python tools/prof_nni.py 4000 10000 2 2 > gpurun_out/prof_nni_4k_ncu.log 2>&1
echo "ncu nni 4k rc=$?"
cat gpurun_out/prof_*_plain.log
