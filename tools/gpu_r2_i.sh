#!/bin/bash
# round 2, call I (N GPUs): phased multi-GPU distance -- parity, CLI byte tests (new writers), timing per phase count, bench under torchrun
N=${1:-2}
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_host_cli.py tests/test_gpu_distance.py -x -q ) > gpurun_out/r02i_pytest_n$N.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02i_pytest_n$N.log
for ph in 1 2 4 8; do echo "phases $ph"; LZG_MULTI_PHASES=$ph timeout 600 python tools/prof_multi.py 20000 10000 2 2>&1 | grep "GPU(s)"; done | tee gpurun_out/r02i_multi_phases_n$N.log
timeout 600 python tools/prof_multi.py 20000 10000 2 > gpurun_out/r02i_multi_cfg3_n$N.log 2>&1; cat gpurun_out/r02i_multi_cfg3_n$N.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 \
    > gpurun_out/r02i_bench_n$N.json 2> gpurun_out/r02i_bench_n$N.err ); echo "bench rc=$?"; tail -3 gpurun_out/r02i_bench_n$N.err; wc -l gpurun_out/r02i_bench_n$N.json
python - <<PY
import json
d=json.loads(open("gpurun_out/r02i_bench_n$N.json").read().splitlines()[-1])
print("value", d["value"], "ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["e2e"].get("matrix_equal_to_1gpu"))
print("nni", d["nni"]["ms_per_step"], d["nni"]["climb"])
PY
