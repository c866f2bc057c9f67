#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $OUT/qi_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/qi_pytest.log
GS2D_STATS=0 timeout 300 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
for k in 1 8; do GS2D_PEER_SELF=$k GS2D_PEER_DEBUG=1 python tools/peer_self.py 2>&1 | tail -1; done
