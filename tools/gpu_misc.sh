#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
for k in 1 2 8; do GS2D_PEER_SELF=$k GS2D_PEER_DEBUG=1 python tools/peer_self.py 2>&1 | tail -2; done
GS2D_PEER_SELF=8 QM_ITERS=3 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_adam_peer' -s 3 -c 1 -f -o $OUT/peer_self python tools/peer_self.py > $OUT/peer_self_ncu.log 2>&1; echo "ncu exit $?"
