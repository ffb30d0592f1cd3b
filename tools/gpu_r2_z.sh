#!/bin/bash
# round 2, call Z: last single-GPU check of the round -- whole GPU suite, smoke, bench N=1
mkdir -p gpurun_out
( timeout 2400 python -m pytest tests -x -q -m gpu ) > gpurun_out/r02z_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02z_pytest_gpu.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02z_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02z_smoke.log
( time timeout 900 python bench.py > gpurun_out/r02z_bench_n1.json 2> gpurun_out/r02z_bench_n1.err ); echo "bench rc=$?"; tail -3 gpurun_out/r02z_bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02z_bench_n1.json").read().splitlines()[-1])
print("value %.4g ms %.3f e2e %.3f ms launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"]))
r=d["roofline"]; print({k:r.get(k) for k in ("kernel","achieved","peak","frac","kernel_ms","prepare_ms","executed_frac")})
PY
