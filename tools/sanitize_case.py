#!/usr/bin/env python
"""Small fit that walks every kernel path, for compute-sanitizer (tools/gpu_sanitize.sh): Gaussians of all sizes (some
taller than the row table: in-kernel footprint solve; some huge: slot overflow -> triples -> k_place), opacity
optimised, two tile-row bands one after the other, stand-alone forward / backward, graph replay and plain launches,
both arithmetic modes, a batch of two images."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
W, H, N = 200, 150, 400
rng = np.random.default_rng(1)
s = np.zeros(N, gs2d.SPLAT_DTYPE)
s["pos"] = rng.uniform(-5, [W + 5, H + 5], (N, 2))
s["sx"], s["sy"] = rng.uniform(1, 14, N), rng.uniform(1, 14, N)
s["sx"][::9] = rng.uniform(40, 300, len(s["sx"][::9]))
s["rot"], s["color"], s["opacity"] = rng.uniform(-3, 3, N), rng.uniform(0, 1, (N, 3)), rng.uniform(0.1, 1, N)
target = gs2d.synth_target(W, H, seed=2)
for exact in (0, 1):
    gs2d.set_tuning("GS2D_EXACT", exact)
    with gs2d.Fitter(W, H, N, n_images=2) as fit:
        for i in range(2):
            fit.set_target(target, image=i)
            fit.upload(s, image=i)
        fit.optimize_opacity = True
        fit.forward(); fit.backward(); fit.adam()
        for r in range(2):
            fit.set_band(r, 2)
            fit.forward(); fit.backward()
        fit.set_band(0, 1)
        mse = fit.step(3)
        gs2d.set_tuning("GS2D_GRAPH", 0)
        fit.set_profiling(True)
        mse2 = fit.step(2)
        fit.set_profiling(False)
        gs2d.set_tuning("GS2D_GRAPH", 1)
        gs2d.set_tuning("GS2D_PDL", 0)
        mse3 = fit.step(1)
        gs2d.set_tuning("GS2D_PDL", 1)
        img = fit.image(1)
        print("exact" if exact else "relaxed", "mse", mse[:, 0], mse2[:, 0], mse3[:, 0], "image finite", bool(np.isfinite(img).all()))
print("sanitize_case OK")
