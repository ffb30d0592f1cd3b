#!/bin/bash
# round 2, call U: evidence for the final kernels -- ncu --set full of k_tc_distance at cfg3, launch list of the bench command
mkdir -p gpurun_out
timeout 300 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02u_prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_tc_distance -s 1 -c 1 -f \
    -o gpurun_out/r02u_ncu_tc_cfg3 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02u_ncu_tc.log 2>&1
echo "ncu tc rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02u_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r02u_launches.csv \
    -k 'regex:^k_' python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02u_ncu_launches.log 2>&1
echo "ncu launches rc=$?"; wc -l gpurun_out/r02u_launches.csv
