#!/bin/bash
# round 2, call C: climb with the fused arg-min + shared-memory cost reductions; per-kernel times of one iteration (ncu launch list, queued-batch mode)
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_climb.py tests/test_gpu_host_cli.py -x -q ) > gpurun_out/r02c_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02c_pytest.log
timeout 600 python tools/prof_climb.py 1000 5000 1 > gpurun_out/r02c_climb_cfg4.log 2>&1; cat gpurun_out/r02c_climb_cfg4.log
LZG_CLIMB_FUSE=0 timeout 600 python tools/prof_climb.py 1000 5000 1 > gpurun_out/r02c_climb_cfg4_nofuse.log 2>&1; cat gpurun_out/r02c_climb_cfg4_nofuse.log
timeout 900 python tools/prof_climb.py 10000 10000 2 > gpurun_out/r02c_climb_cfg5.log 2>&1; cat gpurun_out/r02c_climb_cfg5.log
for cfg in "1000 5000 1" "10000 10000 2"; do
  tag=$(echo $cfg | tr ' ' '_')
  LZG_CLIMB_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02c_launches_$tag.csv \
    -k 'regex:^k_(nni|up_|apply)' python tools/prof_climb.py $cfg > gpurun_out/r02c_ncu_$tag.log 2>&1
  python tools/summarize_launches.py gpurun_out/r02c_launches_$tag.csv | tee gpurun_out/r02c_launches_$tag.txt
done
