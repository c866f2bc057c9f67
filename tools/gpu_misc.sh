#!/bin/bash
GS2D_STATS=0 timeout 300 python tools/quick_time.py 4096 1000000 30 2>&1 | tail -1 | cut -c40-200
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider -k "slot_binning or culling or capacity or binning_switches" 2>&1 | tail -2
