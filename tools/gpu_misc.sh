#!/bin/bash
for v in "" "GS2D_SIDE_STREAM=0"; do
  echo -n "[$v] "; env $v GS2D_STATS=0 timeout 300 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-190
done
AB_ITERS=30 timeout 600 python tools/variant_ab.py gpurun_out/prio_ab.json default "GS2D_SIDE_STREAM=0" > gpurun_out/prio_ab.log 2>&1; python - <<P
import json
for r in json.load(open("gpurun_out/prio_ab.json")):
    print(r["variant"], "| graph %.4f plain %.4f |" % (r.get("graph_ms_per_it", 0), r.get("plain_ms_per_it", 0)), r.get("groups_ms"))
P
