#!/bin/bash
# 8-GPU call: correctness at N=8 + bench at N=8 and N=4 (run with gpurun --gpus 8)
TAG=${1:-n8}; OUT=gpurun_out; mkdir -p $OUT
MGPU_ITERS=30 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 tools/mgpu_check.py > $OUT/${TAG}_check.log 2>&1; echo "check exit $?"
grep -E "rank|world=|comm mode|identical|ok" $OUT/${TAG}_check.log | tail -14
run() { n=$1; shift; v="$*"; name=${TAG}_bench_n${n}_$(echo ${v:-default} | tr ' =' '__')
  env $v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus $n --steps ${STEPS:-100} --warmup 5 > $OUT/$name.json 2> $OUT/$name.err
  echo "bench N=$n [$v] exit $?"
  python - <<P
import json
try:
    d = json.load(open("$OUT/$name.json"))
    k = d["roofline"]["kernels"]
    ex = d.get("exchange") or {}
    print("value %.1f e2e %.1f exact %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"]), {n: round(v["ms"], 4) for n, v in k.items()}, ex.get("sparse"), ex.get("nvlink_mb_per_step_per_rank"))
except Exception as e:
    print("no line:", e)
P
}
run 8
run 8 GS2D_PEER_SPARSE=0
run 8 GS2D_COMM_GRAPH=0
run 4
