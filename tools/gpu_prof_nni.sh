#!/bin/bash
# ncu --set full of the NNI pass kernels at cfg5's shape (tools/prof_nni.py 10000 10000 2 1)
mkdir -p gpurun_out
timeout 300 python tools/prof_nni.py 10000 10000 2 1 > gpurun_out/prof_nni_cfg5_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_nni_items|k_up_pass' -s 4 -c 2 -f \
    -o gpurun_out/prof_nni_cfg5 python tools/prof_nni.py 10000 10000 2 1 > gpurun_out/prof_nni_cfg5_ncu.log 2>&1
echo "ncu nni cfg5 rc=$?"
cat gpurun_out/prof_nni_cfg5_plain.log
