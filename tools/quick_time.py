#!/usr/bin/env python
"""Quick per-kernel-group timing of the fit iteration (developer tool; the judged numbers come from bench.py).
usage: [GS2D_FWD_MINB=..] [GS2D_BWD_MINB=..] [GS2D_STATS=0] python tools/quick_time.py [size] [splats] [iters] [n_images]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 20
B = int(sys.argv[4]) if len(sys.argv) > 4 else 1
rng = np.random.default_rng(0)
tgt = rng.integers(0, 256, (S, S, 4), dtype=np.uint8)
with gs2d.Fitter(S, S, N, n_images=B) as fit:
    for i in range(B):
        fit.set_target_u8(tgt, image=i)
    fit.restart()
    if os.environ.get('GS2D_OPT_OPACITY') == '1':  # the "Optimize opacity" checkbox (main.cpp:825): k_backward<.., OPACITY = true>
        fit.optimize_opacity = True
    if os.environ.get('GS2D_BAND'):  # "r/n": time one rank's band of an n-rank run on a single GPU (no all-reduce)
        r, n = map(int, os.environ['GS2D_BAND'].split('/'))
        fit.set_band(r, n)
    fit.step(3)
    fit.set_profiling(True)
    fit.set_stats(os.environ.get('GS2D_STATS', '1') != '0')
    fit.step(K, want_mse=False)
    fit.sync()
    ms = fit.last_step_ms() / K
    prof, n = fit.get_profile()
    st = fit.stats()
    print(f"MINB fwd={os.environ.get('GS2D_FWD_MINB','-')} bwd={os.environ.get('GS2D_BWD_MINB','-')} stats={os.environ.get('GS2D_STATS','1')} {S}x{S} N={N} B={B}: "
          f"{ms:.3f} ms/it ({1e3/ms:.1f} it/s) | " + " ".join(f"{k}={v:.3f}" for k, v in prof.items()) +
          f" | keys/it={st['keys_sum']/max(st['n_forwards'],1):.3e} live/it={st['live_pairs_sum']/max(st['n_forwards'],1):.3e}"
          f" staged/it={st['staged_sum']/max(st['n_forwards'],1):.3e}")
