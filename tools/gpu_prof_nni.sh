#!/bin/bash
# ncu --set full of k_nni_pass at cfg5's shape: launches = [full pass, UP pass + 1 item, scoring only]
# usage: bash tools/gpu_prof_nni.sh [tag]   (LZG_PASS_DEBUG is passed through)
tag=${1:-cfg5}
mkdir -p gpurun_out
timeout 300 python tools/prof_nni.py 10000 10000 2 1 > gpurun_out/prof_nni_${tag}_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_nni_pass -c 3 -f \
    -o gpurun_out/prof_nni_${tag} python tools/prof_nni.py 10000 10000 2 1 > gpurun_out/prof_nni_${tag}_ncu.log 2>&1
echo "ncu nni ${tag} rc=$?"
cat gpurun_out/prof_nni_${tag}_plain.log
