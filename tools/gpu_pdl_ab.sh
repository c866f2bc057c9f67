#!/bin/bash
# One short GPU-box call: smoke() of the default path, then the A/B of the forward -> backward overlap (GS2D_PDL).
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader > $OUT/pdl_smi.csv 2>&1
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/pdl_smoke.log 2>&1; echo "smoke exit $?" | tee -a $OUT/pdl_smoke.log
tail -2 $OUT/pdl_smoke.log
timeout 200 python tools/pdl_ab.py 4096 1000000 30 $OUT/pdl_ab.json > $OUT/pdl_ab.log 2>&1; echo "ab exit $?" | tee -a $OUT/pdl_ab.log
cat $OUT/pdl_ab.log | cut -c1-260
