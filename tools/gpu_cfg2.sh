#!/bin/bash
# C4 and C5 on one GPU with the final kernels (GPU arm only: the CPU baselines of these configs are in r2_bench_c{4,5}_n1.json)
OUT=gpurun_out; mkdir -p $OUT
for cfg in c4 c5; do
  timeout 600 python bench.py --config $cfg --steps 20 --warmup 5 --no-cpu-baseline > $OUT/r2h_bench_${cfg}_n1.json 2> $OUT/r2h_bench_${cfg}_n1.err; echo "bench $cfg exit $?"
  python - <<P
import json
try:
    d = json.load(open("$OUT/r2h_bench_${cfg}_n1.json"))
    print("value %.2f e2e %.2f exact %.2f | ms %.4f | frac %.3f" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"], d["roofline"]["frac"]), d.get("image_iters_per_sec"))
except Exception as e:
    print("no line:", e)
P
done
