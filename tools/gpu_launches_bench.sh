#!/bin/bash
# ncu launch list of the bench command itself (short): per-launch device time of every kernel bench.py launches
OUT=gpurun_out; TAG=${1:-llb}; mkdir -p $OUT
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_bench.log 2>&1
echo "launch list exit $?"; wc -l $OUT/${TAG}_launches_bench.csv
