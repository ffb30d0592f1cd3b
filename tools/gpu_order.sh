#!/bin/bash
# A/B of the distance tile order on the GPU box: equality + times, then DRAM bytes per launch under ncu.
mkdir -p gpurun_out
timeout 600 python tools/check_dist_order.py > gpurun_out/dist_order.log 2>&1; echo "order rc=$?"
for al in 0 1; do
LZG_DIST_ALIGN=$al timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none \
    -k regex:k_zcnt_distance -s 1 -c 1 --csv --log-file gpurun_out/dist_order_ncu_al$al.csv \
    python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/dist_order_ncu_al$al.log 2>&1; echo "ncu align=$al rc=$?"
done
cat gpurun_out/dist_order.log; tail -4 gpurun_out/dist_order_ncu_al0.csv | cut -d, -f13-; tail -4 gpurun_out/dist_order_ncu_al1.csv | cut -d, -f13-
