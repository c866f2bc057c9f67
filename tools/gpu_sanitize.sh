#!/bin/bash
# compute-sanitizer over every kernel path: memcheck (out-of-bounds / misaligned), racecheck (the
# warp-private shared-memory reuse of the blend kernels: reduce_group, the staged records and coverage words),
# initcheck (reads of device memory nobody wrote), synccheck. Logs -> gpurun_out/<tag>_sanitize_<tool>.log; the
# summaries are committed under profiles/.
TAG=${1:-san}; OUT=gpurun_out; mkdir -p $OUT
python tools/sanitize_case.py > $OUT/${TAG}_sanitize_plain.log 2>&1 || { echo "case fails without the sanitizer"; tail -5 $OUT/${TAG}_sanitize_plain.log; exit 1; }
for tool in memcheck racecheck initcheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 7 python tools/sanitize_case.py > $OUT/${TAG}_sanitize_$tool.log 2>&1
  echo "$tool exit $?: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $OUT/${TAG}_sanitize_$tool.log | tail -1)"
done
