#!/bin/bash
# one full ncu capture of each blend kernel (after the same command exited 0 without ncu)
OUT=gpurun_out; TAG=${1:-ncu}; mkdir -p $OUT
export GS2D_STATS=0
timeout 300 python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_quick.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_forward|k_backward' -s 8 -c 2 -f \
    -o $OUT/${TAG}_blend python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; cat $OUT/${TAG}_quick.log
