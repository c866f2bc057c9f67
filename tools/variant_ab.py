#!/usr/bin/env python
"""A/B of tuning-knob variants of the native library on one GPU (developer tool; judged numbers come from bench.py).
usage: python tools/variant_ab.py OUT.json VARIANT [VARIANT ...]     VARIANT = "NAME=V,NAME=V" or "default"
For every variant, in ONE process (gs2d_set_tuning): the 4096^2 / 1 M fit restarted, 3 warm-up + K iterations
  * replayed as a CUDA graph (profiling off): ms per iteration = what bench.py's `value` measures;
  * as plain launches with the per-group CUDA events on: the kernel-group breakdown;
and two correctness probes against the first variant: sha256 of the image of a forward pass from the reference's
init() (the lists are ordered, so it must be identical bit for bit) and of a 512^2 fit with 2000 huge Gaussians
(200..1024 px: hundreds of entries per bucket, capacity growth), plus the mse sequence (1e-6: unordered atomics)."""
import hashlib, importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
OUT = sys.argv[1]
VARIANTS = sys.argv[2:] or ["default"]
S, N, K = int(os.environ.get("AB_SIZE", 4096)), int(os.environ.get("AB_SPLATS", 1_000_000)), int(os.environ.get("AB_ITERS", 30))
KNOBS = sorted({kv.split("=")[0] for v in VARIANTS if v != "default" for kv in v.split(",")})
DEFAULTS = {"GS2D_PDL": 1, "GS2D_SLOTS": 2, "GS2D_GRAPH": 1, "GS2D_CULL": 1, "GS2D_EXACT": 0, "GS2D_ROWTAB": 1,
            "GS2D_SIDE_STREAM": 1}


def apply(variant):
    for k in KNOBS:
        gs2d.set_tuning(k, DEFAULTS.get(k, 0))
    if variant != "default":
        for kv in variant.split(","):
            k, v = kv.split("=")
            gs2d.set_tuning(k, int(v))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


rng = np.random.default_rng(0)
tgt = rng.integers(0, 256, (S, S, 4), dtype=np.uint8)
big = np.zeros(2000, gs2d.SPLAT_DTYPE)
big["pos"] = rng.uniform(0, 511, (2000, 2))
big["sx"], big["sy"] = rng.uniform(200, 1024, 2000), rng.uniform(200, 1024, 2000)
big["rot"], big["color"], big["opacity"] = rng.uniform(0, 3.14, 2000), rng.uniform(0, 1, (2000, 3)), rng.uniform(0.1, 0.3, 2000)
tgt_big = gs2d.synth_target(512, 512, seed=4)
results, ref = [], None
t00 = time.time()
with gs2d.Fitter(S, S, N) as fit:
    fit.set_target_u8(tgt)
    for variant in VARIANTS:
        rec = {"variant": variant}
        try:
            apply(variant)
            fit.set_stats(True)
            fit.set_profiling(False)
            fit.restart()
            fit.forward()
            rec["init_image_sha"] = sha(fit.image())
            st = fit.stats()
            rec["keys"], rec["live"] = int(st["keys"]), int(st["live_pairs"])
            fit.set_stats(False)
            fit.restart()
            fit.step(3)
            mse = fit.step(K)[:, 0]
            rec["graph_ms_per_it"] = fit.last_step_ms() / K
            fit.set_profiling(True)
            fit.restart()
            fit.step(3)
            fit.set_profiling(True)
            fit.step(K)
            prof, n = fit.get_profile()
            rec["plain_ms_per_it"] = fit.last_step_ms() / K
            rec["groups_ms"] = {k: round(v, 4) for k, v in prof.items()}
            fit.set_profiling(False)
            with gs2d.Fitter(512, 512, 2000) as fb:
                fb.set_target(tgt_big)
                fb.upload(big)
                fb.forward()
                rec["big_image_sha"] = sha(fb.image())
                rec["big_mse"] = [float(x) for x in fb.step(2)[:, 0]]
            rec["mse_first"], rec["mse_last"] = float(mse[0]), float(mse[-1])
            if ref is None:
                ref = (rec["init_image_sha"], rec["big_image_sha"], mse, rec["big_mse"])
            rec["init_image_equal"] = rec["init_image_sha"] == ref[0]
            rec["big_image_equal"] = rec["big_image_sha"] == ref[1]
            rec["mse_max_rel"] = float(np.max(np.abs(mse - ref[2]) / ref[2]))
            rec["big_mse_max_rel"] = float(np.max(np.abs(np.array(rec["big_mse"]) - np.array(ref[3])) / np.array(ref[3])))
        except Exception as e:  # keep going: the other variants still tell something
            rec["error"] = repr(e)
        rec["t"] = round(time.time() - t00, 1)
        print(json.dumps(rec), flush=True)
        results.append(rec)
with open(OUT, "w") as f:
    json.dump(results, f, indent=1)
