#!/bin/bash
# BASELINE configs[3] and [4] on one GPU (with their CPU baselines)
OUT=gpurun_out; mkdir -p $OUT
for c in c5 c4; do
  timeout 1500 python bench.py --config $c --steps 20 --warmup 5 > $OUT/r2_bench_${c}_n1.json 2> $OUT/r2_bench_${c}_n1.err; echo "bench $c exit $?"
  python - <<P
import json
try:
    d = json.load(open("$OUT/r2_bench_${c}_n1.json"))
    print("$c value %.2f e2e %.2f exact %.2f | ms %.3f | frac %.3f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"], d["roofline"]["frac"]), d.get("image_iters_per_sec"), d["cpu_baseline"])
    print({n: round(v["ms"], 4) for n, v in d["roofline"]["kernels"].items()})
except Exception as e:
    print("no line", e)
P
  tail -2 $OUT/r2_bench_${c}_n1.err
done
