#!/bin/bash
# round 2, call X: after the CTA-pair kernel -- whole GPU suite, smoke, bench N=1, ncu --set full of k_tc_distance2 at cfg3
mkdir -p gpurun_out
( timeout 2400 python -m pytest tests -x -q -m gpu ) > gpurun_out/r02x_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02x_pytest_gpu.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02x_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02x_smoke.log
( time timeout 900 python bench.py > gpurun_out/r02x_bench_n1.json 2> gpurun_out/r02x_bench_n1.err ); echo "bench rc=$?"; tail -3 gpurun_out/r02x_bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02x_bench_n1.json").read().splitlines()[-1])
print("value %.4g ms %.3f e2e %.3f ms launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"]))
r=d["roofline"]; print({k:r.get(k) for k in ("kernel","achieved","peak","frac","frac_of_2x_measured_bf16","kernel_ms","prepare_ms","executed_frac","tiles")})
PY
timeout 300 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02x_prof_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_tc_distance2 -s 1 -c 1 -f \
    -o gpurun_out/r02x_ncu_tc2_cfg3 python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/r02x_ncu_tc2.log 2>&1
echo "ncu rc=$?"
