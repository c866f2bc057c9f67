#!/bin/bash
# One GPU-box call: parity tests, bench line, ncu launch list and one full ncu capture of each blend kernel.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh <tag>'
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.csv 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cat $OUT/${TAG}_bench.json | cut -c1-600
if [ "$2" != "noncu" ]; then
export GS2D_STATS=0   # profile the variant the timed run uses (work counters off)
timeout 300 python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_quick.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu1.log 2>&1
echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_forward|k_backward' -s 8 -c 2 -f \
    -o $OUT/${TAG}_blend python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
fi
cat $OUT/${TAG}_quick.log
