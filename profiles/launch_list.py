#!/usr/bin/env python
"""Print one iteration of an ncu launch list (gpu__time_duration.sum per launch). usage: launch_list.py X.csv [iteration]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
it = int(sys.argv[2]) if len(sys.argv) > 2 else 3
for i, r in enumerate(rows):
    if r and r[0] == 'ID':
        hdr = r; start = i + 1; break
ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
seq = [(r[ki].split('(')[0], float(r[vi].replace(',', '')) * (1e-3 if r[ui] in ('ns', 'nsecond') else 1.0)) for r in rows[start:] if len(r) > vi]
idx = [i for i, (k, v) in enumerate(seq) if 'k_preprocess' in k]
a, b = idx[it], idx[it + 1]
tot = sum(v for k, v in seq[a:b])
for k, v in seq[a:b]:
    print(f"{v:9.1f} us {v/tot*100:5.1f}%  {k}")
print(f"{tot:9.1f} us total (serialised, cold-cache: compare shares)")
