#!/bin/bash
# One gpurun call for the cfg3-headline bench: both arms, the launch list of the same command,
# and an ncu --set full capture of k_zcnt_distance at cfg3 (device-resident, one launch per matrix).
# Usage (on the GPU box, from the repo root): bash tools/gpu_final.sh
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 300 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/prof_dist_cfg3_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_zcnt_distance -s 1 -c 1 -f \
    -o gpurun_out/prof_dist_cfg3 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/prof_dist_cfg3_ncu.log 2>&1
echo "ncu dist cfg3 rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
    -k 'regex:^k_(zcnt|zero|fill|bias|nni|up_|apply|merge|level|gather|label|unpack)' \
    python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
cat gpurun_out/prof_dist_cfg3_plain.log; cat gpurun_out/bench_reference.json; cut -c1-2500 gpurun_out/bench_n1.json
