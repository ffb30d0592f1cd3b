#!/bin/bash
# DRAM bytes per cfg3 matrix for several super-block edges of the distance plan (LZG_DIST_SB), under ncu.
mkdir -p gpurun_out
for sb in 6 8 16 24; do
LZG_DIST_SB=$sb timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none \
    -k regex:k_zcnt_distance -s 1 -c 1 --csv --log-file gpurun_out/dist_sb$sb.csv \
    python tools/prof_distance.py 20000 10000 2 2 device > gpurun_out/dist_sb$sb.log 2>&1
echo "SB=$sb rc=$? $(grep -E 'dram__bytes_read|time_duration|hit_rate' gpurun_out/dist_sb$sb.csv | awk -F'","' '{printf "%s=%s ", $(NF-2), $NF}')"
done
