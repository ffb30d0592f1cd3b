#!/bin/bash
# round 2, call Q: bench (N=1) on the tensor-core path, launch list, ncu --set full of k_tc_distance at cfg3
mkdir -p gpurun_out
( time timeout 900 python bench.py > gpurun_out/r02q_bench_n1.json 2> gpurun_out/r02q_bench_n1.err ); echo "bench rc=$?"; tail -4 gpurun_out/r02q_bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02q_bench_n1.json").read().splitlines()[-1])
print("value %.4g ms %.3f e2e %.3f ms launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"]))
print(json.dumps(d["roofline"])[:1800])
print(json.dumps(d.get("cfg2",{}).get("roofline",{}))[:300])
print(json.dumps(d.get("nni",{}).get("tail",{}))[:900])
PY
timeout 300 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02q_prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_tc_distance -s 1 -c 1 -f \
    -o gpurun_out/r02q_ncu_tc_cfg3 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02q_ncu_tc.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/*.ncu-rep
