#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
AB_ITERS=30 timeout 300 python tools/variant_ab.py $OUT/r2r_ab.json default GS2D_SCAN_FUSED=0 default GS2D_SCAN_FUSED=0 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_equal'), r.get('big_image_equal'), r.get('mse_max_rel'), r.get('error'))"
for v in 1 0; do for b in 3/8 1/2; do
  echo -n "band $b FUSED=$v: "
  GS2D_SCAN_FUSED=$v GS2D_STATS=0 GS2D_BAND=$b timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-160
done; done
