#!/bin/bash
# round 2, call F: whole GPU suite, smoke, both bench arms, ncu --set full captures (distance cfg3 / cfg2, climb kernels cfg4 / cfg5),
# launch list of the bench command.  Summaries are written straight into profiles-ready text under gpurun_out/.
mkdir -p gpurun_out
export LZG_PIN_CANDIDATES=${LZG_PIN_CANDIDATES:-3000}
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 ) > gpurun_out/r02f_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/r02f_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02f_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02f_smoke.log
( time timeout 900 python bench.py > gpurun_out/r02f_bench_n1.json 2> gpurun_out/r02f_bench_n1.err ); echo "bench rc=$?"; tail -5 gpurun_out/r02f_bench_n1.err
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02f_bench_reference.json 2> gpurun_out/r02f_bench_reference.err ); echo "ref rc=$?"
# ncu --set full: the distance kernel at cfg3 and cfg2 (device-resident, one launch per matrix)
for cfg in "20000 10000 2 2:cfg3" "1000 5000 2 3:cfg2"; do
  a=${cfg%%:*}; tag=${cfg##*:}
  timeout 300 python tools/prof_distance.py $a device > gpurun_out/r02f_prof_dist_${tag}_plain.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_zcnt_distance -s 1 -c 1 -f \
      -o gpurun_out/r02f_dist_$tag python tools/prof_distance.py $a device > gpurun_out/r02f_prof_dist_${tag}_ncu.log 2>&1
  echo "ncu dist $tag rc=$?"
  python tools/ncu_summary.py gpurun_out/r02f_dist_$tag.ncu-rep > gpurun_out/r02_ncu_dist_$tag.txt
done
# ncu --set full: the climb's kernels (queued mode: the graph's kernels one by one): full UP pass, all items, then the first iterations
for cfg in "10000 10000 2:cfg5" "1000 5000 1:cfg4"; do
  a=${cfg%%:*}; tag=${cfg##*:}
  LZG_CLIMB_GRAPH=0 timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^k_(up_df|items_wl|apply2|up_delta|dirty_scan|cand_assemble|nni_argmin)' -c 13 -f \
      -o gpurun_out/r02f_climb_$tag python tools/prof_climb.py $a 3 > gpurun_out/r02f_prof_climb_${tag}_ncu.log 2>&1
  echo "ncu climb $tag rc=$?"
  python tools/ncu_summary.py gpurun_out/r02f_climb_$tag.ncu-rep > gpurun_out/r02_ncu_climb_$tag.txt
done
# launch list of the bench command (every kernel but the NJ seed tree's)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r02f_launches.csv \
    -k 'regex:^k_(zcnt|zero|fill|bias|nni|up_|apply|merge|level|gather|label|unpack|dirty|items|cand)' \
    python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r02f_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python tools/summarize_launches.py gpurun_out/r02f_launches.csv --by-name > gpurun_out/r02_launches_bench.txt; cat gpurun_out/r02_launches_bench.txt
cut -c1-3000 gpurun_out/r02f_bench_n1.json; echo; cat gpurun_out/r02f_bench_reference.json
