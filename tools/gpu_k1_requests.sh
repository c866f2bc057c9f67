#!/bin/bash
# L2 requests / atomics of k_preprocess per binning mode (before/after of "one store per arrival")
OUT=gpurun_out; mkdir -p $OUT
for sl in 1 2; do
  GS2D_SLOTS=$sl GS2D_STATS=0 timeout 600 ncu --set full --clock-control none -k regex:'k_preprocess' -s 3 -c 1 -f -o $OUT/k1_slots$sl python tools/quick_time.py 4096 1000000 3 > $OUT/k1_slots$sl.log 2>&1; echo "ncu slots=$sl exit $?"
done
