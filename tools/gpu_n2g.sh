#!/bin/bash
# short 2-GPU call: default exchange path at two ranks (fit == single-GPU fit) + driver-style bench line
OUT=gpurun_out; mkdir -p $OUT
MGPU_ITERS=20 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 tools/mgpu_check.py > $OUT/r2g_check_n2.log 2>&1; echo "check exit $?"
grep -E "world=|comm mode|identical|all ranks" $OUT/r2g_check_n2.log | tail -5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 2 --steps 20 --warmup 5 > $OUT/r2g_bench_n2.json 2> $OUT/r2g_bench_n2.err
echo "bench N=2 exit $?"
python - <<P
import json
d = json.load(open("$OUT/r2g_bench_n2.json"))
print("value %.1f e2e %.1f exact %.1f | ms %.4f |" % (d["value"], d["e2e"]["value"], d["exact_mode"]["value"], d["ms_per_step"]), (d.get("exchange") or {}).get("sparse"))
P
