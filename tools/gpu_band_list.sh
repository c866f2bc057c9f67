#!/bin/bash
# launch list (ncu, serialised) of one rank's band of an 8-rank run, emulated on one GPU (GS2D_BAND=r/n)
OUT=gpurun_out; mkdir -p $OUT; B=${1:-3/8}
GS2D_STATS=0 GS2D_BAND=$B timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/band_launches.csv python tools/quick_time.py 4096 1000000 3 > $OUT/band_ncu.log 2>&1; echo "ncu exit $?"
python profiles/launch_list.py $OUT/band_launches.csv | tail -16
GS2D_STATS=0 GS2D_BAND=$B timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-260
