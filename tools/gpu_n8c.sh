#!/bin/bash
TAG=${1:-n8c}; OUT=gpurun_out; mkdir -p $OUT
for probe in 0 1 2 3 4 8 11; do
  GS2D_PEER_DEBUG=1 GS2D_PEER_PROBE=$probe QM_TAG="probe=$probe" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node ${NPROC:-8} --master-addr 127.0.0.1 --master-port 29553 tools/quick_multi.py > $OUT/${TAG}_probe$probe.log 2>&1
  echo "probe $probe exit $?: $(grep 'ms/it' $OUT/${TAG}_probe$probe.log)"; grep "peer update over" $OUT/${TAG}_probe$probe.log | sort | sed -n '1p;8p'
done
