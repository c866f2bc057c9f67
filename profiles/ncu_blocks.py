#!/usr/bin/env python
"""Summarise an ncu report of the blend kernels: headline metrics and SASS grouped into runs of equal execution count
(= basic blocks of the hot loops), per tile.
usage: python profiles/ncu_blocks.py X.ncu-rep [n_tiles] [min_total]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
ntiles = float(sys.argv[2]) if len(sys.argv) > 2 else 65536.0
min_total = float(sys.argv[3]) if len(sys.argv) > 3 else 150.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'smsp__warps_eligible.avg.per_cycle_active', 'smsp__warps_active.avg.per_cycle_active']
keys += [h for h in hdr if 'smsp__average_warps_issue_stalled' in h and h.endswith('_per_issue_active.ratio')]
names = [r[hdr.index('Kernel Name')].split('(')[0] for r in rows[2:]]
print("kernels:", names)
for k in keys:
    if k in hdr:
        i = hdr.index(k)
        vals = [r[i] for r in rows[2:]]
        if any(float(v.replace(',', '')) > 0.02 for v in vals):
            print(f"{k} [{units[i]}]: " + "  ".join(vals))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kname, hdr2, per = None, None, {}
for r in csv.reader(io.StringIO(src)):
    if not r:
        continue
    if r[0] == "Kernel Name":
        kname = r[1].split('(')[0]; per[kname] = []; continue
    if r[0] == "Address":
        hdr2 = r; ie = r.index("Instructions Executed"); te = r.index("Thread Instructions Executed"); sm = r.index("# Samples"); continue
    if kname and r[0].startswith("0x"):
        per[kname].append((r[1].strip(), int(r[ie]), int(r[te]), int(r[sm])))
for kname, ins in per.items():
    T = sum(i[1] for i in ins) / ntiles; S = sum(i[3] for i in ins)
    print(f"=== {kname}: {T:.0f} warp-instructions per tile, {len(ins)} SASS lines")
    runs = []
    for n, (txt, e, t, s) in enumerate(ins):
        c = e / ntiles
        if runs and abs(runs[-1]['c'] - c) <= 0.02 * max(c, 1):
            u = runs[-1]; u['n'] += 1; u['end'] = n; u['tot'] += c; u['smp'] += s; u['thr'] += t
        else:
            runs.append(dict(start=n, end=n, c=c, n=1, tot=c, smp=s, thr=t, first=txt))
    for u in runs:
        if u['tot'] >= min_total:
            act = u['thr'] / max(u['tot'] * ntiles, 1)
            print(f"  {u['start']:4d}-{u['end']:4d} n={u['n']:3d} x {u['c']:7.1f}/tile = {u['tot']:8.1f} ({u['tot']/T*100:4.1f}%) lanes={act:4.1f} samples={u['smp']/max(S,1)*100:4.1f}%  {u['first'][:48]}")
