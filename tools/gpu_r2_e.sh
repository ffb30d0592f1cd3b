#!/bin/bash
# round 2, call E: incremental climb -- parity tests, CLI goldens, timing at cfg4 / cfg5
mkdir -p gpurun_out
export LZG_CLIMB_DEBUG=1
( timeout 600 python -m pytest tests/test_gpu_climb.py -x -q ) > gpurun_out/r02e_pytest_climb.log 2>&1; echo "climb rc=$?"; tail -25 gpurun_out/r02e_pytest_climb.log
unset LZG_CLIMB_DEBUG
( timeout 900 python -m pytest tests/test_gpu_host_cli.py tests/test_gpu_tree.py -x -q ) > gpurun_out/r02e_pytest_cli.log 2>&1; echo "cli rc=$?"; tail -5 gpurun_out/r02e_pytest_cli.log
timeout 600 python tools/prof_climb.py 1000 5000 1 > gpurun_out/r02e_climb_cfg4.log 2>&1; cat gpurun_out/r02e_climb_cfg4.log
timeout 900 python tools/prof_climb.py 10000 10000 2 > gpurun_out/r02e_climb_cfg5.log 2>&1; cat gpurun_out/r02e_climb_cfg5.log
for cfg in "1000 5000 1" "10000 10000 2"; do
  tag=$(echo $cfg | tr ' ' '_')
  LZG_CLIMB_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02e_launches_$tag.csv \
    -k 'regex:^k_(nni|up_|apply|dirty|items|cand)' python tools/prof_climb.py $cfg > gpurun_out/r02e_ncu_$tag.log 2>&1
  python tools/summarize_launches.py gpurun_out/r02e_launches_$tag.csv | tee gpurun_out/r02e_launches_$tag.txt
done
