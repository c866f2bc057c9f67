#!/bin/bash
# One multi-GPU bench line, launched the way the driver launches it.
# usage: gpurun --gpus N --timeout 600 -- 'bash tools/gpu_bench_n.sh N <tag>'
N=${1:-2}
TAG=${2:-n$N}
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench exit $?"
cut -c1-400 gpurun_out/${TAG}_bench.json
tail -3 gpurun_out/${TAG}_bench.err
