#!/bin/bash
# The driver's scaling run: bench.py at N = 1, 2, 4, 8 back to back on one 8-GPU box.
# usage (gpurun --gpus 8): bash tools/gpu_scale.sh [steps]
steps=${1:-10}
mkdir -p gpurun_out
timeout 600 python bench.py --gpus 1 --steps $steps --warmup 3 --no-nni-large > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo "N=1 rc=$?"
for N in 2 4 8; do
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + N)) \
        bench.py --gpus $N --steps $steps --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err; echo "N=$N rc=$?"
done
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open("gpurun_out/scale_n%d.json" % n).read().strip().splitlines()[-1])
        x = d.get("nni", {})
        print("N=%d %s: %.3e pairs/s (%.2f ms, frac %.2f); nni %s: %.3e cand/s (pass %.3f ms, iteration %.3f ms)" % (
            n, d["config"]["workload"], d["value"], d["ms_per_step"], d["roofline"]["frac"], x.get("config", {}).get("workload"),
            x.get("value", 0), x.get("ms_per_step", 0), x.get("iteration", {}).get("ms_per_step", 0)))
    except Exception as e:
        print("N=%d: %r" % (n, e))
PY
