#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q -k "band or slot_binning or crop or edge or row_table" > $OUT/r2k_pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/r2k_pytest.log
for v in 1 0 1 0; do
  echo -n "band 3/8 BAND_LIST=$v: "
  GS2D_BAND_LIST=$v GS2D_STATS=0 GS2D_BAND=3/8 timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
done | tee $OUT/r2k_band.log
for v in 1 0; do
  echo -n "band 1/2 BAND_LIST=$v: "
  GS2D_BAND_LIST=$v GS2D_STATS=0 GS2D_BAND=1/2 timeout 200 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
done | tee -a $OUT/r2k_band.log
bash tools/gpu_band_list.sh 3/8 | head -10
