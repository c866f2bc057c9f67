#!/usr/bin/env python
"""Device-timed fit iterations/s of the BASELINE.json configs that are not the bench.py headline (C1, C2, C4, C5) on
ONE GPU, with the CPU oracle timed beside C1/C2 (1 thread). Developer tool: writes JSON lines to stdout.
usage: python tools/configs_bench.py [c1 c2 c4 c5 c5share]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
gs2d = importlib.import_module("2dgaussiansplatting_b200")

def golden_target(cfg):
    rgb = np.load(os.path.join(ROOT, "tests", "golden", f"{cfg}_target_rgb8.npz"))["rgb"]
    h, w, _ = rgb.shape
    return w, h, np.concatenate([rgb, np.full((h, w, 1), 255, np.uint8)], axis=2)

def gpu_rate(W, H, N, B, targets_u8, iters, warm=10):
    with gs2d.Fitter(W, H, N, n_images=B) as fit:
        for i in range(B):
            fit.set_target_u8(targets_u8[i % len(targets_u8)], image=i)
        fit.restart()
        fit.step(warm)
        fit.set_stats(False)
        fit.step(iters, want_mse=False)   # profiling off: the iteration is replayed as a CUDA graph
        fit.sync()
        ms = fit.last_step_ms() / iters
        fit.set_profiling(True)           # per-group times: plain launches bracketed by events
        fit.step(max(iters // 4, 1), want_mse=False)
        fit.sync()
        prof, _ = fit.get_profile()
        mse = fit.last_mse()
    return ms, prof, float(mse[0])

def cpu_rate(W, H, N, target_u8, iters):
    import oracle_binding as ob
    st = ob.State(W, H, N)
    tgt = (target_u8.astype(np.float32) / np.float32(255.0))
    st.step(tgt)
    t0 = time.perf_counter()
    for _ in range(iters):
        st.step(tgt)
    return (time.perf_counter() - t0) / iters

def synth_u8(S, seed):
    return np.rint(gs2d.synth_target(S, S, seed=seed) * 255.0).astype(np.uint8)

which = sys.argv[1:] or ["c1", "c2", "c4", "c5", "c5share"]
for c in which:
    if c in ("c1", "c2"):
        W, H, t = golden_target(c)
        ms, prof, mse = gpu_rate(W, H, 1024, 1, [t], 2000, warm=50)
        cpu_s = cpu_rate(W, H, 1024, t, 100)
        out = {"config": c, "image": [W, H], "n_gaussians": 1024, "gpu_ms_per_it": ms, "gpu_it_s": 1e3 / ms,
               "cpu_it_s_1thread": 1.0 / cpu_s, "speedup": cpu_s * 1e3 / ms, "groups_ms": prof}
    elif c == "c4":
        t = synth_u8(8192, 1)
        ms, prof, mse = gpu_rate(8192, 8192, 4_000_000, 1, [t], 20, warm=4)
        out = {"config": "c4 on ONE gpu", "image": [8192, 8192], "n_gaussians": 4_000_000, "gpu_ms_per_it": ms,
               "gpu_it_s": 1e3 / ms, "groups_ms": prof}
    else:
        B = 256 if c == "c5" else 32
        ts = [synth_u8(512, 1 + i) for i in range(8)]
        ms, prof, mse = gpu_rate(512, 512, 16384, B, ts, 30, warm=5)
        out = {"config": f"c5: {B} independent 512x512 images x 16384 Gaussians on ONE gpu" +
                         (" (= one rank's share of the 8-GPU replicas run)" if B == 32 else ""),
               "gpu_ms_per_it": ms, "batch_it_s": 1e3 / ms, "image_it_s": B * 1e3 / ms, "groups_ms": prof}
    print(json.dumps(out), flush=True)
