#!/bin/bash
# NNI iteration loop on the GPU box: tree parity tests, then timings of the pass at two sizes.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_tree.py -m gpu -x -q > gpurun_out/pytest_tree.log 2>&1; rc=$?; echo "pytest tree rc=$rc"
tail -15 gpurun_out/pytest_tree.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 600 python -m pytest tests/test_gpu_golden.py tests/test_gpu_host_cli.py -m gpu -x -q 2>&1 | tail -5
for e in 0 1; do
echo "LZG_EAGER_UP=$e"
LZG_EAGER_UP=$e timeout 300 python tools/prof_nni.py 1000 5000 1 3 2>&1 | tail -4
LZG_EAGER_UP=$e timeout 300 python tools/prof_nni.py 10000 10000 2 2 2>&1 | tail -4
done
