#!/bin/bash
# one full ncu capture of each binning kernel of an iteration (k_preprocess, k_row_table, k_group_count, k_scan_*, k_place, k_rank_groups)
OUT=gpurun_out; TAG=${1:-ncubin}; mkdir -p $OUT
export GS2D_STATS=0
timeout 300 python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_quick.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_preprocess|k_group_count|k_scan|k_place|k_rank' -s 18 -c 6 -f \
    -o $OUT/${TAG}_bin python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; cat $OUT/${TAG}_quick.log
