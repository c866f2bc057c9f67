#!/bin/bash
# 8-GPU call: breakdown of the peer update (GS2D_PEER_DEBUG=1 prints per-rank averages at destroy)
TAG=${1:-n8b}; OUT=gpurun_out; mkdir -p $OUT
run() { n=$1; shift; v="$*"; name=${TAG}_n${n}_$(echo ${v:-default} | tr ' =' '__')
  env $v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus $n --steps ${STEPS:-60} --warmup 5 > $OUT/$name.json 2> $OUT/$name.err
  echo "bench N=$n [$v] exit $?"
  grep "peer update over" $OUT/$name.err | sort | head -8
  python - <<P
import json
try:
    d = json.load(open("$OUT/$name.json"))
    k = d["roofline"]["kernels"]
    print("value %.1f e2e %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["ms_per_step"]), {n: round(v["ms"], 4) for n, v in k.items() if n in ("blend_overlapped", "place_rank", "k_adam", "preprocess_emit", "exchange")})
except Exception as e:
    print("no line:", e)
P
}
run 8 GS2D_PEER_DEBUG=1
run 8 GS2D_PEER_DEBUG=1 GS2D_PEER_SPARSE=0
run 8 GS2D_PEER_DEBUG=1 GS2D_PEER_TRAFFIC=0
