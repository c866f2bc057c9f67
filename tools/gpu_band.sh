#!/bin/bash
for b in "" "3/8"; do
  echo -n "band=$b: "
  GS2D_STATS=0 GS2D_BAND=$b timeout 300 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
done
