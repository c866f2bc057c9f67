#!/usr/bin/env python
"""Check of the adaptive binning switch (gs2d_ctx::slot_binning): a fit whose Gaussians overflow every group / slot row must
switch to the triple path at its first synchronisation point and give the same loss curve as GS2D_SLOTS=0; a normal
fit must stay on the slot path. usage: GS2D_DEBUG_BINNING=1 python tools/adapt_check.py"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
rng = np.random.default_rng(0)
big = np.zeros(2000, gs2d.SPLAT_DTYPE)
big["pos"] = rng.uniform(0, 511, (2000, 2))
big["sx"], big["sy"] = rng.uniform(200, 1024, 2000), rng.uniform(200, 1024, 2000)
big["rot"], big["color"], big["opacity"] = rng.uniform(0, 3.14, 2000), rng.uniform(0, 1, (2000, 3)), rng.uniform(0.1, 0.3, 2000)
tb, tn = gs2d.synth_target(512, 512, seed=4), gs2d.synth_target(1024, 768, seed=5)
out = {}
for slots in (2, 1, 0):
    gs2d.set_tuning("GS2D_SLOTS", slots)
    with gs2d.Fitter(512, 512, 2000) as f:
        f.set_target(tb)
        f.upload(big)
        out["big", slots] = np.concatenate([f.step(1)[:, 0] for _ in range(4)] + [f.step(3)[:, 0]])
    with gs2d.Fitter(1024, 768, 60000) as f:
        f.set_target(tn)
        f.restart()
        out["normal", slots] = np.concatenate([f.step(1)[:, 0] for _ in range(3)] + [f.step(4)[:, 0]])
gs2d.set_tuning("GS2D_SLOTS", 2)
for k in ("big", "normal"):
    for slots in (2, 1):
        a, b = out[k, slots], out[k, 0]
        rel = float(np.max(np.abs(a - b) / b))
        print(k, f"GS2D_SLOTS={slots}: mse", a[0], "...", a[-1], "max rel diff adaptive vs GS2D_SLOTS=0:", rel, flush=True)
        assert a[0] == b[0] and rel < 1e-6, (a, b)
print("adapt_check OK")
