#!/bin/bash
# driver-style bench line at N = 4 (C3, 20 steps)
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 4 --steps 20 --warmup 5 > $OUT/r2g_bench_n4.json 2> $OUT/r2g_bench_n4.err
echo "bench N=4 exit $?"
python - <<P
import json
try:
    d = json.load(open("$OUT/r2g_bench_n4.json"))
    k = d["roofline"]["kernels"]
    print("value %.1f e2e %.1f exact %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"]), {n: round(v["ms"], 4) for n, v in k.items() if "ms" in v})
except Exception as e:
    print("no line:", e)
P
