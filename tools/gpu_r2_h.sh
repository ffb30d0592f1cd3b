#!/bin/bash
# round 2, call H (N GPUs): bench.py under torchrun as the driver launches it (distance strong-scaled + multi-GPU e2e, sharded climb), multi tests
N=${1:-2}
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_climb.py -x -q ) > gpurun_out/r02h_pytest_n$N.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02h_pytest_n$N.log
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 \
    > gpurun_out/r02h_bench_n$N.json 2> gpurun_out/r02h_bench_n$N.err ); echo "bench rc=$?"; tail -15 gpurun_out/r02h_bench_n$N.err; cut -c1-1500 gpurun_out/r02h_bench_n$N.json
timeout 600 python tools/prof_multi.py 20000 10000 2 > gpurun_out/r02h_multi_cfg3_n$N.log 2>&1; cat gpurun_out/r02h_multi_cfg3_n$N.log
