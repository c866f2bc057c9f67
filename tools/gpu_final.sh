#!/bin/bash
# Round-end style check on one box: GPU parity tests, smoke, the bench line, the ncu launch list of the bench command.
TAG=${1:-final}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests -m gpu -x -q --durations=8 > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -14 $OUT/${TAG}_pytest.log
timeout 300 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cut -c1-1500 $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
if [ "$2" != "noncu" ]; then
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_bench.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu1.log 2>&1; echo "launch list exit $?"
fi
