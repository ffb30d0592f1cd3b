#!/bin/bash
# round 2, call M: the no-alloc entry point alone in a fresh process, then the whole GPU suite (it is collected last), smoke
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q ) > gpurun_out/r02m_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/r02m_noalloc_alone.log
( timeout 2400 python -m pytest tests -x -q -m gpu --durations=8 ) > gpurun_out/r02m_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -14 gpurun_out/r02m_pytest_gpu.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02m_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r02m_smoke.log
