#!/bin/bash
# round 2, group-row binning: GPU suite, same-box A/B of GS2D_SLOTS=2 (default) vs 1, whole image and one rank's band of 8
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/r2b_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/r2b_pytest.log
AB_ITERS=30 timeout 300 python tools/variant_ab.py $OUT/r2b_ab.json default GS2D_SLOTS=1 default GS2D_SLOTS=1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_equal'), r.get('big_image_equal'), r.get('mse_max_rel'), r.get('error'))"
for sl in 2 1 2 1; do
  echo -n "band 3/8 slots=$sl: "
  GS2D_SLOTS=$sl GS2D_STATS=0 GS2D_BAND=3/8 timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-260
done | tee $OUT/r2b_band.log
