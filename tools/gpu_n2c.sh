#!/bin/bash
TAG=${1:-n2c}; OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2
for v in "GS2D_PEER_PROBE=0" "GS2D_PEER_PROBE=0 GS2D_PEER_SPARSE=0"; do
  name=$(echo $v | tr ' =' '__')
  env GS2D_PEER_DEBUG=1 $v QM_ITERS=30 QM_TAG="$v" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29553 tools/quick_multi.py > $OUT/${TAG}_$name.log 2>&1
  echo "[$v] exit $?: $(grep 'ms/it' $OUT/${TAG}_$name.log)"; grep "peer update over" $OUT/${TAG}_$name.log | sort | sed -n '1p'
done
