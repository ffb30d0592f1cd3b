#!/bin/bash
# round 2, call R (N GPUs): multi-GPU distance with the tensor-core path -- parity, timing per GPU count, bench under torchrun
N=${1:-2}
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_multi.py -x -q ) > gpurun_out/r02r_pytest_n$N.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02r_pytest_n$N.log
timeout 600 python tools/prof_multi.py 20000 10000 2 > gpurun_out/r02r_multi_cfg3_n$N.log 2>&1; cat gpurun_out/r02r_multi_cfg3_n$N.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 \
    > gpurun_out/r02r_bench_n$N.json 2> gpurun_out/r02r_bench_n$N.err ); echo "bench rc=$?"; tail -3 gpurun_out/r02r_bench_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02r_bench_n$N.json").read().splitlines()[-1])
print("value %.4g ms %.3f e2e %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"]), d["e2e"].get("matrix_equal_to_1gpu"), d.get("n1_same_workload"))
print(json.dumps(d["roofline"])[:700])
print("nni", d["nni"]["ms_per_step"], d["nni"]["climb"])
PY
