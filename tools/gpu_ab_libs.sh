#!/bin/bash
# compile-time A/B on one box: the same variant list against libgs2d_base.so (previous commit) and libgs2d.so
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ablib}; shift
for rep in 1 2; do
GS2D_LIB=$PWD/2dgaussiansplatting_b200/libgs2d_base.so timeout 100 python tools/variant_ab.py $OUT/${TAG}_base$rep.json "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print('base', r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_sha'), r.get('big_image_sha'), r.get('error'))"
timeout 100 python tools/variant_ab.py $OUT/${TAG}_new$rep.json "$@" 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print('new ', r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_sha'), r.get('big_image_sha'), r.get('error'))"
done
