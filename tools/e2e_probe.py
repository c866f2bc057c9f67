"""Where does the end-to-end loop of bench.py lose time against the device-resident run? Same iterations of the C3 fit
(restart + W warm-up), K steps, wall clock per step for the loop with parts of it left out.
usage: python tools/e2e_probe.py [S N K W]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
from gs2d_loader import gs2d  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 20
W = int(sys.argv[4]) if len(sys.argv) > 4 else 5

pin = torch.empty((S, S, 4), dtype=torch.uint8, pin_memory=True)
pin.numpy()[...] = np.rint(gs2d.synth_target(S, S, seed=1) * 255.0).astype(np.uint8)
fit = gs2d.Fitter(S, S, N)
fit.set_target_u8_ptr(pin.data_ptr())
fit.restart()
fit.step(3)


def run(name, ticket, upload, adopt=1, block=False):
    gs2d.set_tuning("GS2D_ADOPT_TARGET", adopt)
    fit.set_target_u8_ptr(pin.data_ptr())
    fit.restart()
    fit.step(W)
    if upload:
        fit.set_target_u8_ptr_async(pin.data_ptr())
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    prev = None
    marks = []
    for _ in range(K):
        fit.step(1, want_mse=False)
        tk = fit.mse_ticket() if ticket else None
        if upload:
            fit.set_target_u8_ptr_async(pin.data_ptr())
        if prev is not None:
            fit.mse_wait(prev)
            marks.append(time.perf_counter())
        prev = tk
    if prev is not None:
        fit.mse_wait(prev)
        marks.append(time.perf_counter())
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    d = np.diff(np.asarray(marks)) * 1e3 if len(marks) > 2 else np.zeros(1)
    print("%-34s %.4f ms/step to last mse, %.4f incl. final sync | between mse arrivals: median %.4f min %.4f max %.4f"
          % (name, (t1 - t0) * 1e3 / K, (t2 - t0) * 1e3 / K, np.median(d), d.min(), d.max()), flush=True)


# device-resident reference: K iterations in one call, CUDA events of the library
fit.restart()
fit.step(W - 1)
torch.cuda.synchronize()
fit.step(1, want_mse=False)
fit.step(K, want_mse=False)
fit.sync()
print("device-resident (events)           %.4f ms/step" % (fit.last_step_ms() / K), flush=True)
fit.restart()
fit.step(W)
torch.cuda.synchronize()
t0 = time.perf_counter()
fit.step(K, want_mse=False)
fit.sync()
print("device-resident (wall, K in a call) %.4f ms/step" % ((time.perf_counter() - t0) * 1e3 / K), flush=True)

run("steps only", False, False)
run("steps + tickets", True, False)
run("steps + uploads", False, True)
run("steps + tickets + uploads (e2e)", True, True)
run("the same, staging copied (old)", True, True, adopt=0)
run("steps + tickets + uploads (e2e)", True, True)
