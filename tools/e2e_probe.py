#!/usr/bin/env python
"""Where the end-to-end step loses time against the device-resident one (developer tool): wall-clock per step of
loops that add the upload, the conversion and the loss read-back one at a time.
usage: python tools/e2e_probe.py [size] [splats] [steps]"""
import importlib, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 30
rng = np.random.default_rng(0)
tgt = rng.integers(0, 256, (S, S, 4), dtype=np.uint8)
pin = torch.from_numpy(tgt.copy()).pin_memory()


def timed(name, body, fit):
    fit.restart()  # every loop times the SAME iterations of the fit (the work per iteration drifts as it converges)
    fit.step(5)
    fit.sync()
    t0 = time.perf_counter()
    body()
    fit.sync()
    print(f"{name:58s} {(time.perf_counter() - t0) * 1e3 / K:.3f} ms/step", flush=True)


with gs2d.Fitter(S, S, N) as fit:
    fit.set_target_u8(tgt)
    fit.restart()
    fit.set_stats(False)
    fit.step(5)

    def a():
        fit.step(K, want_mse=False)

    def b():
        for _ in range(K):
            fit.step(1, want_mse=False)

    def c():
        prev = None
        for _ in range(K):
            fit.step(1, want_mse=False)
            t = fit.mse_ticket()
            if prev is not None:
                fit.mse_wait(prev)
            prev = t
        fit.mse_wait(prev)

    def d():
        for _ in range(K):
            fit.step(1, want_mse=False)
            fit.set_target_u8_ptr_async(pin.data_ptr())

    def e():
        prev = None
        fit.set_target_u8_ptr_async(pin.data_ptr())
        for _ in range(K):
            fit.step(1, want_mse=False)
            t = fit.mse_ticket()
            fit.set_target_u8_ptr_async(pin.data_ptr())
            if prev is not None:
                fit.mse_wait(prev)
            prev = t
        fit.mse_wait(prev)

    def f():
        for _ in range(K):
            fit.step(1, want_mse=False)
            fit.last_mse()

    timed("one gs2d_step(K) call", a, fit)
    timed("K gs2d_step(1) calls", b, fit)
    acc = {}

    def call(name, fn, *args):
        t0 = time.perf_counter()
        r = fn(*args)
        acc[name] = acc.get(name, 0.0) + time.perf_counter() - t0
        return r

    def e_host():
        prev = None
        fit.set_target_u8_ptr_async(pin.data_ptr())
        for _ in range(K):
            call("step", fit.step, 1, False)
            t = call("ticket", fit.mse_ticket)
            call("upload", fit.set_target_u8_ptr_async, pin.data_ptr())
            if prev is not None:
                call("wait", fit.mse_wait, prev)
            prev = t
        fit.mse_wait(prev)

    timed("e2e loop with host times per call", e_host, fit)
    print("   host ms per step: " + ", ".join(f"{k} {v * 1e3 / K:.3f}" for k, v in acc.items()))
    timed("K x (step, ticket, wait previous)", c, fit)
    timed("K x (step, async upload)", d, fit)
    timed("K x (step, ticket, async upload, wait previous) = e2e", e, fit)
    timed("K x (step, last_mse)", f, fit)
    # the H2D copy alone
    dst = torch.empty(S * S * 4, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        dst.copy_(pin.view(-1), non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    print(f"H2D of {S * S * 4 / 1e6:.0f} MB from pinned memory alone: {dt * 1e3:.3f} ms ({S * S * 4 / dt / 1e9:.1f} GB/s)")
