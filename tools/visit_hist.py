#!/usr/bin/env python
"""Developer statistic: how full are the 8x4 block visits of the forward blend? (GS2D_VISIT_HIST=1, counting variant)
usage: GS2D_VISIT_HIST=1 python tools/visit_hist.py [size] [splats] [iters]"""
import ctypes as C, importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("GS2D_VISIT_HIST", "1")
gs2d = importlib.import_module("2dgaussiansplatting_b200")
S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 3
tgt = np.rint(gs2d.synth_target(S, S, seed=1) * 255).astype(np.uint8)
with gs2d.Fitter(S, S, N) as fit:
    fit.set_target_u8(tgt)
    fit.restart()
    fit.set_stats(True)
    fit.step(K)
    out = (C.c_ulonglong * 66)()
    fit.lib.gs2d_debug_visit_hist.argtypes = [C.POINTER(C.c_ulonglong * 66), C.c_int]
    fit.lib.gs2d_debug_visit_hist(C.byref(out), 1)
h = np.array(out[:33], np.float64); t = np.array(out[33:], np.float64)
tot = h.sum()
print(f"{tot/K:.3e} block visits per iteration; mean covered-and-live lanes per visit {np.dot(h, np.arange(33))/tot:.2f}")
c = np.cumsum(h) / tot
for k in (1, 2, 4, 8, 12, 16, 20, 24, 28, 31, 32):
    print(f"  visits with <= {k:2d} active lanes: {c[k]*100:5.1f} %   (lane-slots in them: {np.dot(h[:k+1], np.arange(k+1))/np.dot(h, np.arange(33))*100:5.1f} % of the live pairs)")
ct = np.cumsum(t) / t.sum()
print("visits by live pixels left in the tile (of 256):")
for k in (1, 2, 4, 8, 16, 24, 32):
    print(f"  <= {8*k:3d} live pixels: {ct[k]*100:5.1f} % of the visits")
