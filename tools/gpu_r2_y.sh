#!/bin/bash
# round 2, call Y: range kernel rewrite -- parity tests, smoke, per-kernel durations at cfg3
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_distance_tc.py tests/test_gpu_multi.py -x -q ) > gpurun_out/r02y_pytest.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/r02y_pytest.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02y_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02y_smoke.log
LZG_TC_TIMING=1 timeout 300 python tools/prof_distance.py 20000 10000 2 4 device 2>&1 | tail -1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 8 --csv --log-file gpurun_out/r02y_launches.csv python tools/prof_distance.py 20000 10000 2 1 device > gpurun_out/r02y.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r02y_launches.csv")) if len(r)>10]
h=rows[0]
for r in rows[1:]:
    print(r[h.index("Kernel Name")][:40], r[h.index("Metric Value")])
PY
