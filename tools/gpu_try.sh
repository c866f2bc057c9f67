#!/bin/bash
# quick GPU check of a kernel change: parity tests, then timing (optionally a sweep of the blend register caps)
OUT=gpurun_out; TAG=${1:-try}; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee $OUT/${TAG}_pytest.log
if [ "$2" == "sweep" ]; then
for f in 32 24 20; do for b in 20 16; do
GS2D_FWD_MINB=$f GS2D_BWD_MINB=$b timeout 120 python tools/quick_time.py 4096 1000000 20 2>&1 | tail -1 | tee -a $OUT/${TAG}_sweep.log
done; done
fi
GS2D_STATS=0 timeout 120 python tools/quick_time.py 4096 1000000 20 2>&1 | tail -1 | tee -a $OUT/${TAG}_sweep.log
