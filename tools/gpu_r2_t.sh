#!/bin/bash
# round 2, call T: whole GPU suite, smoke, bench (both arms) at N=1 after the tensor-core path
mkdir -p gpurun_out
( timeout 2400 python -m pytest tests -x -q -m gpu --durations=6 ) > gpurun_out/r02t_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -12 gpurun_out/r02t_pytest_gpu.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02t_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02t_smoke.log
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r02t_bench_reference.json 2> gpurun_out/r02t_bench_reference.err ); echo "ref rc=$?"
( time timeout 900 python bench.py > gpurun_out/r02t_bench_n1.json 2> gpurun_out/r02t_bench_n1.err ); echo "bench rc=$?"; tail -3 gpurun_out/r02t_bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02t_bench_n1.json").read().splitlines()[-1])
print("value %.4g ms %.3f e2e %.3f ms launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["gpu_launches"]))
r=d["roofline"]; print({k:r[k] for k in ("achieved","peak","frac","kernel_ms","prepare_ms","executed_frac","tiles")})
print(json.dumps(d.get("nni",{}).get("tail",{}))[:700])
print("nni", d["nni"]["ms_per_step"], "cfg4", d.get("nni_cfg4",{}).get("ms_per_step"))
PY
