#!/bin/bash
# compile-time A/B on one box: the same variant list against libgs2d_base.so (previous commit) and libgs2d.so,
# then (optionally) the GPU test suite against the new library. Boxes differ by +-1.5 %, so only numbers of ONE call
# compare. Build the base first (it travels with the snapshot, *.so is git-ignored):
#   rm -rf /tmp/base && mkdir /tmp/base && git archive HEAD | tar -x -C /tmp/base && \
#   make -C /tmp/base/2dgaussiansplatting_b200/csrc -j8 && \
#   cp /tmp/base/2dgaussiansplatting_b200/libgs2d.so 2dgaussiansplatting_b200/libgs2d_base.so
# usage: gpurun -- 'bash tools/gpu_ab_libs.sh TAG REPS pytest|nopytest VARIANT...'
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ablib}; REPS=${2:-1}; PYT=${3:-nopytest}; shift 3
show() { python -c "
import sys, json
for l in sys.stdin:
    try: r = json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print('$1', r.get('variant'), r.get('graph_ms_per_it'), r.get('groups_ms'), r.get('init_image_sha'), r.get('big_image_sha'), r.get('error'))"; }
for rep in $(seq 1 $REPS); do
GS2D_LIB=$PWD/2dgaussiansplatting_b200/libgs2d_base.so timeout 100 python tools/variant_ab.py $OUT/${TAG}_base$rep.json "$@" 2>&1 | show base
timeout 100 python tools/variant_ab.py $OUT/${TAG}_new$rep.json "$@" 2>&1 | show "new "
done
if [ "$PYT" == "pytest" ]; then
timeout 300 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/${TAG}_pytest.log
fi
