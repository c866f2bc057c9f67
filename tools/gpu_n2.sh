#!/bin/bash
# 2-GPU call: multi-rank tests + bench at N=2 (run with gpurun --gpus 2)
TAG=${1:-n2}; OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -s -p no:cacheprovider > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
grep -E "rank|passed|failed|world=|comm mode" $OUT/${TAG}_pytest.log | tail -30
for v in "" "GS2D_PEER_SPARSE=0" "GS2D_COMM_GRAPH=0"; do
  env $v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps ${STEPS:-50} --warmup 5 > $OUT/${TAG}_bench_${v:-default}.json 2> $OUT/${TAG}_bench_${v:-default}.err
  echo "bench [$v] exit $?"
  python - <<P
import json
try:
    d = json.load(open("$OUT/${TAG}_bench_${v:-default}.json"))
    k = d["roofline"]["kernels"]
    print("value %.1f e2e %.1f exact %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"]), {n: round(v["ms"], 4) for n, v in k.items()}, d.get("exchange"))
except Exception as e:
    print("no line:", e)
P
done
tail -5 $OUT/${TAG}_bench_default.err
