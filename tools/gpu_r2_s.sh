#!/bin/bash
# round 2, call S: tensor-core kernel after a change -- parity tests, per-kernel durations at cfg3, device-resident timing
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_distance_tc.py tests/test_gpu_multi.py -x -q ) > gpurun_out/r02s_pytest.log 2>&1; echo "tc rc=$?"; tail -3 gpurun_out/r02s_pytest.log
timeout 600 python tools/prof_distance.py 20000 10000 2 5 device > gpurun_out/r02s_prof_cfg3.log 2>&1; tail -2 gpurun_out/r02s_prof_cfg3.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 12 --csv --log-file gpurun_out/r02s_launches.csv python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02s.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r02s_launches.csv")) if len(r)>10]
h=rows[0]
for r in rows[1:]:
    print(r[h.index("Kernel Name")][:40], r[h.index("Metric Value")])
PY
