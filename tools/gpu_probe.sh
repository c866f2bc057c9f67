#!/bin/bash
# K1 knock-out timings (GS2D_K1_PROBE: results wrong, timing only)
for pr in 0 1 2 3; do
  echo -n "K1_PROBE=$pr: "
  GS2D_K1_PROBE=$pr GS2D_STATS=0 timeout 200 python tools/quick_time.py 4096 1000000 10 2>&1 | tail -1 | cut -c40-200
done | tee gpurun_out/probe_k1.log
