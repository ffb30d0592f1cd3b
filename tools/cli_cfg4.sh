#!/bin/bash
# End to end: `lazac nni` (C++ host over the C ABI) on BASELINE.json configs[3]'s input -- synthetic
# 1,000 cells x 5,000 bins (the CSV carries two alleles; -a is the aggression, the reference CLI
# reads cn_a only), NJ seed tree, whole search.  usage: bash tools/cli_cfg4.sh [cells bins iterations]
cells=${1:-1000}; bins=${2:-5000}; iters=${3:-3}
mkdir -p gpurun_out
d=$(mktemp -d)
python - <<PY
import sys, time
sys.path.insert(0, ".")
from lazac_copy_number_b200 import synth
t = time.time()
prof = synth.make_profiles($cells, $bins, 2, seed=0)
synth.write_cn_csv("$d/in.csv", prof)
print("csv written in %.1f s" % (time.time() - t))
PY
ls -la $d/in.csv
t0=$(date +%s%N)
./host/lazac nni $d/in.csv -o $d/out -a 1.0 -i $iters -s 7 > gpurun_out/cli_cfg4.log 2> gpurun_out/cli_cfg4.err
rc=$?
t1=$(date +%s%N)
echo "lazac nni rc=$rc wall $(( (t1 - t0) / 1000000 )) ms"
grep -m1 "No seed tree" gpurun_out/cli_cfg4.log | cut -c1-26
grep -m1 "iteration 0:" gpurun_out/cli_cfg4.log | cut -c1-26
grep "Candidate tree scores" gpurun_out/cli_cfg4.log | tail -1 | cut -c1-120
tail -1 gpurun_out/cli_cfg4.log
ls -la $d | awk '{print $5, $9}'
rm -rf $d
