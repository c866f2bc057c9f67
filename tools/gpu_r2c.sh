#!/bin/bash
# round 2: GPU suite + ncu launch list of a short bench (per-kernel times of the group-row binning)
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/r2c_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/r2c_pytest.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/r2c_launches.csv python tools/quick_time.py 4096 1000000 3 > $OUT/r2c_ncu.log 2>&1; echo "ncu exit $?"
python profiles/launch_list.py $OUT/r2c_launches.csv | tail -20
