#!/bin/bash
# round 2, call N: the bench as the driver runs it at N=1 (reference arm first), then NJ timing of the three forms
mkdir -p gpurun_out
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r02n_bench_reference.json 2> gpurun_out/r02n_bench_reference.err ); echo "ref rc=$?"
( time timeout 900 python bench.py > gpurun_out/r02n_bench_n1.json 2> gpurun_out/r02n_bench_n1.err ); echo "bench rc=$?"; tail -4 gpurun_out/r02n_bench_n1.err
cut -c1-600 gpurun_out/r02n_bench_reference.json; cut -c1-3000 gpurun_out/r02n_bench_n1.json
for mode in 3 2 1 0; do echo "LZG_NJ_PERSISTENT=$mode"; NJ_SKIP_CLI=1 NJ_RANDOM_PROFILES=1 LZG_NJ_PERSISTENT=$mode timeout 600 python tools/time_nj.py 10000 5000 2000 1000 2>&1 | grep "^n ="; done | tee gpurun_out/r02n_nj_modes.log
