#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` output: instructions executed per CUDA source line.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > cs.csv ; python profiles/ncu_lines.py cs.csv [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
fn = None
hdr = None
per = {}
for r in rows:
    if not r:
        continue
    if r[0] == "Function Name":
        fn = r[1].split("(")[0]
        per.setdefault(fn, [])
        continue
    if r[0] == "Line No":
        hdr = r
        ie, te, sm = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
        pe = hdr.index("Predicated-On Thread Instructions Executed")
        continue
    if fn and hdr and r[0] not in ("", "File Path") and r[2] == "-":
        try:
            per[fn].append((int(r[ie]), int(r[te]), int(r[pe]), int(r[sm]), r[0], r[1].strip()[:110]))
        except ValueError:
            pass
for fn, lines in per.items():
    tot = sum(l[0] for l in lines)
    tots = sum(l[3] for l in lines)
    print(f"=== {fn}: {tot/1e6:.1f} M warp-instructions, {tots} samples")
    for ie_, te_, pe_, s_, ln, src in sorted(lines, reverse=True)[:top]:
        print(f"{ie_/tot*100:5.1f}% inst  {s_/max(tots,1)*100:5.1f}% smp  act {te_/max(ie_,1):5.1f} pred {pe_/max(ie_,1):5.1f}  L{ln:>4} {src}")
