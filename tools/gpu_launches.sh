#!/bin/bash
# ncu launch list (device time of every launch) of a short run
OUT=gpurun_out; TAG=${1:-ll}; mkdir -p $OUT
export GS2D_STATS=0
timeout 300 python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_quick.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu1.log 2>&1
echo "launch list exit $?"; cat $OUT/${TAG}_quick.log
