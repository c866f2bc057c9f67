#!/usr/bin/env python
"""A/B of the forward -> backward overlap (GS2D_PDL, gs2d_blend.cuh) on one GPU: same fit, same iterations, with
and without the overlap, replayed as a CUDA graph and as plain launches, on the whole image and on one rank's band of
an 8-rank run (GS2D_BAND-style: gs2d_set_band without a communicator). Prints one JSON line per variant: ms per
iteration (CUDA events around the K iterations), the mse of the first and last iteration, and whether the mse
sequence equals the default variant's (iteration 0 bit for bit, the rest to 1e-6: gradient atomics are unordered).
usage: python tools/pdl_ab.py [size] [splats] [iters] [out.json]"""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 30
OUT = sys.argv[4] if len(sys.argv) > 4 else None
rng = np.random.default_rng(0)
tgt = rng.integers(0, 256, (S, S, 4), dtype=np.uint8)
results = []
t00 = time.time()
with gs2d.Fitter(S, S, N) as fit:
    fit.set_target_u8(tgt)
    fit.set_stats(False)
    for band in (None, (3, 8)):
        if band:
            fit.set_band(*band)
        ref = None
        for graph in (1, 0):
            for pdl in (0, 1, 0, 1):
                rec = {"band": band, "graph": graph, "pdl": pdl}
                try:
                    gs2d.set_tuning("GS2D_GRAPH", graph)
                    gs2d.set_tuning("GS2D_PDL", pdl)
                    fit.restart()
                    fit.step(3)
                    mse = fit.step(K)[:, 0]
                    rec["ms_per_it"] = fit.last_step_ms() / K
                    rec["mse_first"], rec["mse_last"] = float(mse[0]), float(mse[-1])
                    if ref is None:
                        ref = mse
                    rec["mse0_equal"] = bool(mse[0] == ref[0])
                    rec["mse_max_rel"] = float(np.max(np.abs(mse - ref) / ref))
                except Exception as e:  # keep going: the other variants still tell something
                    rec["error"] = repr(e)
                rec["t"] = round(time.time() - t00, 1)
                print(json.dumps(rec), flush=True)
                results.append(rec)
if OUT:
    with open(OUT, "w") as f:
        json.dump(results, f, indent=1)
