#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench (cfg3 headline + cfg2 + NNI), cfg2 alone, ncu launch list
# (every kernel but the NJ seed tree's ~22000 k_nj_* launches, which would fill the window;
# profiles/r01_launches_bench_cfg2_v4_with_nj.txt holds a capture that includes them).
# Usage (from the repo root, on the GPU box): bash tools/gpu_check.sh
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm,clocks.max.mem,memory.total --format=csv > gpurun_out/gpu.csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --workload cfg2 --steps 20 --warmup 3 --no-nni --no-cpu > gpurun_out/bench_cfg2_alone.json 2> gpurun_out/bench_cfg2.err; echo "bench cfg2 rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
    -k 'regex:^k_(zcnt|zero|fill|bias|nni|up_|apply|merge|level|gather|label|unpack)' \
    python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
echo "ncu rc=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log | tail -2; cat gpurun_out/bench_n1.json; cat gpurun_out/bench_cfg2_alone.json; cat gpurun_out/bench_reference.json
