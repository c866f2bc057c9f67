#!/bin/bash
# short GPU call: selected tests + variant A/B (+ optional ncu of the named kernels)
# usage: gpu_quick.sh TAG "pytest -k expression" "variant ..." [ncu kernel regex]
TAG=${1:-q}; KEXPR=${2:-}; shift 2
OUT=gpurun_out; mkdir -p $OUT
if [ -n "$KEXPR" ]; then
timeout 900 python -m pytest tests -m gpu -q -s -p no:cacheprovider -k "$KEXPR" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -4 $OUT/${TAG}_pytest.log
fi
NCU=${NCU_KERNELS:-}
AB_ITERS=${AB_ITERS:-20} timeout 900 python tools/variant_ab.py $OUT/${TAG}_ab.json "$@" > $OUT/${TAG}_ab.log 2>&1
echo "ab exit $?"; python - <<P
import json
for r in json.load(open("$OUT/${TAG}_ab.json")):
    print(r["variant"], "| graph %.4f plain %.4f |" % (r.get("graph_ms_per_it", 0), r.get("plain_ms_per_it", 0)), r.get("groups_ms"), "| img_eq", r.get("init_image_equal"), r.get("big_image_equal"), "mse_rel %.2e" % r.get("mse_max_rel", -1), r.get("error", ""))
P
if [ -n "$NCU" ]; then
export GS2D_STATS=0
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$NCU" -s ${NCU_SKIP:-12} -c ${NCU_COUNT:-4} -f \
    -o $OUT/${TAG}_ncu python tools/quick_time.py 4096 1000000 3 > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"
fi
