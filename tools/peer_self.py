#!/usr/bin/env python
"""The peer-memory update path on ONE GPU (GS2D_PEER_SELF=k: this rank is its own only peer and owns 1/k of the quads),
for timing / ncu of k_adam_peer and the barrier kernels without a second process. Developer tool."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
S, N, K = 4096, 1_000_000, int(os.environ.get("QM_ITERS", 10))
tgt = np.random.default_rng(0).integers(0, 256, (S, S, 4), dtype=np.uint8)
with gs2d.Fitter(S, S, N) as fit:
    fit.comm_init(0, 1, bytes(128))
    fit.set_target_u8(tgt)
    fit.restart()
    fit.step(3)
    fit.set_profiling(True)
    fit.step(K, want_mse=False)
    fit.sync()
    prof, n = fit.get_profile()
    print({k: round(v, 4) for k, v in prof.items()}, fit.last_step_ms() / K)
