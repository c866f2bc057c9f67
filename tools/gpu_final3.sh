#!/bin/bash
# round 2, final profiles: GPU suite, bench line, full ncu capture of one launch of each hot kernel, launch list of the bench command
TAG=${1:-r2f}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench_n1.json 2> $OUT/${TAG}_bench_n1.err; echo "bench exit $?"; cut -c1-400 $OUT/${TAG}_bench_n1.json
export GS2D_STATS=0
timeout 300 python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_quick.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_forward|k_backward|k_row_table|k_preprocess|k_rank_groups|k_group_scan|k_adam' -s 21 -c 7 -f \
    -o $OUT/${TAG}_kernels python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"; cat $OUT/${TAG}_quick.log
unset GS2D_STATS
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $OUT/${TAG}_launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_bench.log 2>&1
echo "launch list exit $?"; wc -l $OUT/${TAG}_launches_bench.csv
