#!/bin/bash
# round 2, call A: parity tests (both arithmetic modes), variant A/B, bench line, full ncu capture of the blend kernels
TAG=${1:-r2a}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.csv 2>&1
timeout 1200 python -m pytest tests -m gpu -q -s -p no:cacheprovider > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -15 $OUT/${TAG}_pytest.log
AB_ITERS=20 timeout 600 python tools/variant_ab.py $OUT/${TAG}_ab.json "GS2D_EXACT=1,GS2D_ROWTAB=0" "GS2D_EXACT=1" default "GS2D_ROWTAB=0" "GS2D_SIDE_STREAM=0" > $OUT/${TAG}_ab.log 2>&1
echo "ab exit $?"; cut -c1-400 $OUT/${TAG}_ab.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cut -c1-300 $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
export GS2D_STATS=0
timeout 300 python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_quick.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_forward|k_backward|k_row_table|k_preprocess' -s 12 -c 4 -f \
    -o $OUT/${TAG}_blend python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
cat $OUT/${TAG}_quick.log
