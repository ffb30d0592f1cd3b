#!/bin/bash
# round 2: bench.py under torchrun at N GPUs only (no tests), as the driver launches it
N=${1:-8}
mkdir -p gpurun_out
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 \
    > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err ); echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r02_bench_n$N.json").read().splitlines()[-1])
print("value %.4g ms %.3f e2e %.3f" % (d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"]), d["e2e"].get("matrix_equal_to_1gpu"), d.get("n1_same_workload"))
print("nni", d["nni"]["ms_per_step"], d["nni"]["climb"]["equal_to_1_rank_climb"])
PY
