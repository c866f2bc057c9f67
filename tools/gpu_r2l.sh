#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r2l}
timeout 900 python -m pytest tests -m gpu -x -q -k "band or slot_binning or switches or crop or edge or row_table or huge or capacity" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
AB_ITERS=30 timeout 300 python tools/variant_ab.py $OUT/${TAG}_ab.json default GS2D_SLOTS=1 default 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_equal'), r.get('big_image_equal'), r.get('mse_max_rel'), r.get('error'))"
for v in 1 0; do
  echo -n "band 3/8 BAND_LIST=$v: "
  GS2D_BAND_LIST=$v GS2D_STATS=0 GS2D_BAND=3/8 timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
done | tee $OUT/${TAG}_band.log
for v in 1 0; do
  echo -n "band 1/2 BAND_LIST=$v: "
  GS2D_BAND_LIST=$v GS2D_STATS=0 GS2D_BAND=1/2 timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
done | tee -a $OUT/${TAG}_band.log
bash tools/gpu_band_list.sh 3/8 | head -10
