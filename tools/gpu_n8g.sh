#!/bin/bash
# one short 8-GPU call: correctness at N = 8 (8-rank fit == single-GPU fit, every rank's copy exact for its band) and the
# driver-style bench line (C3, 20 steps)
TAG=${1:-r2g}; OUT=gpurun_out; mkdir -p $OUT
MGPU_ITERS=30 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 tools/mgpu_check.py > $OUT/${TAG}_check.log 2>&1; echo "check exit $?"
grep -E "world=|comm mode|identical|ok|OK" $OUT/${TAG}_check.log | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 8 --steps 20 --warmup 5 > $OUT/${TAG}_bench_n8.json 2> $OUT/${TAG}_bench_n8.err
echo "bench N=8 exit $?"
python - <<P
import json
try:
    d = json.load(open("$OUT/${TAG}_bench_n8.json"))
    k = d["roofline"]["kernels"]
    ex = d.get("exchange") or {}
    print("value %.1f e2e %.1f exact %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"]), {n: round(v["ms"], 4) for n, v in k.items() if "ms" in v}, ex.get("sparse"), ex.get("nvlink_mb_per_step_per_rank"))
except Exception as e:
    print("no line:", e)
P
