#!/bin/bash
# usage: gpurun -- 'bash tools/gpu_r2d.sh TAG [pytest-args...]': binning tests, same-box A/B vs GS2D_SLOTS=1, launch list
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r2d}; shift
timeout 900 python -m pytest tests -m gpu -x -q "$@" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/${TAG}_pytest.log
AB_ITERS=30 timeout 300 python tools/variant_ab.py $OUT/${TAG}_ab.json default GS2D_SLOTS=1 default 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_equal'), r.get('big_image_equal'), r.get('mse_max_rel'), r.get('error'))"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches.csv python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu.log 2>&1; echo "ncu exit $?"
python profiles/launch_list.py $OUT/${TAG}_launches.csv | tail -14
