#!/bin/bash
# final 8-GPU call: C3 at N = 8 and 4 (driver-style: 20 steps), C4 and C5 at N = 8
TAG=${1:-r2}; OUT=gpurun_out; mkdir -p $OUT
run() { n=$1; cfg=$2; steps=$3; name=${TAG}_bench_${cfg}_n${n}
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus $n --config $cfg --steps $steps --warmup 5 > $OUT/$name.json 2> $OUT/$name.err
  echo "bench $cfg N=$n exit $?"
  python - <<P
import json
try:
    d = json.load(open("$OUT/$name.json"))
    k = d["roofline"]["kernels"]
    ex = d.get("exchange") or {}
    print("value %.1f e2e %.1f exact %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"]), {n: round(v["ms"], 4) for n, v in k.items() if "ms" in v}, ex.get("sparse"), ex.get("nvlink_mb_per_step_per_rank"), d.get("image_iters_per_sec"))
except Exception as e:
    print("no line:", e)
P
}
run 8 c3 20
run 4 c3 20
run 8 c4 20
run 8 c5 20
