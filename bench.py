#!/usr/bin/env python
"""bench.py — fit iterations/s (forward + backward + Adam) of the B200-native 2D Gaussian-splatting engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--size 4096] [--splats 1000000]

Workload (BASELINE.json metric / configs[2]): synthetic 4096x4096 RGB target, 1 M Gaussians initialised with the
reference's own init() formula (main.cpp:280-305), lr 0.05, opacity not optimised. One "step" = one whole loop body of
the reference (main.cpp:414-809): bin + sort, forward blend, loss, backward, [gradient all-reduce], Adam + clamps.
N > 1 (launched by torchrun, one process per GPU): the image is split into tile-row bands, one per rank, and the
per-Gaussian gradients are all-reduced over NCCL before Adam -> the total work is fixed ("strong" scaling).

Prints ONE JSON line (rank 0). `value` = device-timed iterations/s with all inputs resident in HBM; `e2e` = the same
metric through the C ABI with HOST buffers: every step uploads the RGBA8 target from pinned host memory
(gs2d_set_target_rgba8, the reference's own ingest format, prp.cpp:179-193 + :796-807) and reads the mse back.

--impl reference: the reference's own CPU implementation of the path, single thread (the reference is
single-threaded), on the box's host cores. oracle/_ref/ref_main (the unmodified main.cpp) is hard-wired to one JPEG
and 1024 splats, so at this size the function-ised restatement oracle/gs2d_oracle.c is timed (bit-identical to
main.cpp, tests/test_oracle.py); each step is a bounded sample (1/16 of the workload at equal density).
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_FWD, FLOP_BWD = 27.0, 86.0  # per live (pixel, splat) pair, hoisted counts of SURVEY.md §8(d)
ADAM_BYTES = 288.0               # per Gaussian: 8 planes x (p,g,m,v read + p,m,v,g write) + opacity plane p,g rw
BIN_BYTES_PER_KEY = 13.0         # scan + rank passes per (tile, Gaussian) list entry, see run_ours
SAMPLE_DIV = 16                  # CPU sample = 1/16 of the workload (same Gaussian density)


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index, period=0.01):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml

            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # pragma: no cover
            self.err = repr(e)

    NAMES = {
        0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
        0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown",
        0x100: "display_clock_setting",
    }

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.NAMES.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def cpu_sample(size, splats, iters):
    """Time the CPU oracle (the reference's algorithm, 1 thread) on 1/SAMPLE_DIV of the workload at equal density.
    Returns (equivalent full-workload it/s, description, seconds per sample iteration)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    gs2d = importlib.import_module("2dgaussiansplatting_b200")
    side = int(round(size / SAMPLE_DIV ** 0.5))
    n = max(1, splats // SAMPLE_DIV)
    target = gs2d.synth_target(side, side, seed=1)
    st = ob.State(side, side, n)
    st.step(target)  # page in
    t0 = time.perf_counter()
    for _ in range(iters):
        st.step(target)
    dt = (time.perf_counter() - t0) / iters
    desc = (f"{iters} iterations of a {side}x{side} window with {n} Gaussians (1/{SAMPLE_DIV} of the workload at equal "
            f"density, {dt:.2f} s each); value = sample it/s / {SAMPLE_DIV}")
    return 1.0 / (dt * SAMPLE_DIV), desc, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    t_first = time.perf_counter()
    # warm-up steps and timed steps are both sample iterations
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    gs2d = importlib.import_module("2dgaussiansplatting_b200")
    side = int(round(args.size / SAMPLE_DIV ** 0.5))
    n = max(1, args.splats // SAMPLE_DIV)
    target = gs2d.synth_target(side, side, seed=1)
    st = ob.State(side, side, n)
    for _ in range(args.warmup):
        st.step(target)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st.step(target)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = 1.0 / (dt * SAMPLE_DIV)
    sample = (f"each step = one iteration of a {side}x{side} window with {n} Gaussians (1/{SAMPLE_DIV} of the "
              f"workload at equal density, {dt:.2f} s); value = sample it/s / {SAMPLE_DIV}")
    line = {
        "impl": "reference", "metric": "fit_iters_per_sec", "value": value, "unit": "it/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * SAMPLE_DIV * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": value, "unit": "it/s", "cores": 1, "kind": "port", "sample": sample,
                         "host_cores_available": os.cpu_count()},
        "e2e": {"value": value, "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t_first,
    }
    _emit(line)
    return 0


def workload_config(args, n_gpus):
    return {
        "workload": f"synthetic {args.size}x{args.size} RGB target, {args.splats} Gaussians (BASELINE configs[2])",
        "image": [args.size, args.size], "n_gaussians": args.splats, "init": "reference init() main.cpp:280-305",
        "optimize_opacity": False, "lr": 0.05,
        "parallelism": "single GPU" if n_gpus == 1 else f"{n_gpus} tile-row bands + NCCL all-reduce of 9 gradient planes",
        "l2_policy": "no flush: per-iteration inputs (target 268 MB, image 268 MB, key lists >250 MB) exceed the 126 MB L2",
    }


def run_ours(args):
    import torch

    gs2d = importlib.import_module("2dgaussiansplatting_b200")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one process per GPU)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the product has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    S, N, K, W = args.size, args.splats, args.steps, args.warmup
    target = gs2d.synth_target(S, S, seed=1)
    pinned_u8 = torch.empty((S, S, 4), dtype=torch.uint8, pin_memory=True)
    pinned_u8.numpy()[...] = np.rint(target * 255.0).astype(np.uint8)  # 8-bit target, as a decoded JPEG would be

    fit = gs2d.Fitter(S, S, N, device=local_rank)
    if world > 1:
        box = [gs2d.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        fit.comm_init(rank, world, box[0])
    fit.set_target_u8_ptr(pinned_u8.data_ptr())
    fit.restart()

    peaks = measured_peaks()
    fp32_peak = hbm_peak_local = None
    if rank == 0:
        fp32_peak, hbm_peak_local = gs2d.microbench(local_rank)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- counting pass (untimed): live (pixel, Gaussian) pairs and list entries of the SAME K iterations ---------
    # The counting variant of the forward blend costs ~10 % (profiles/r1_ncu_blend_v4.txt), so the work counters the
    # roofline needs are taken on an identical, untimed run of the same iterations and the timed run has them off.
    fit.step(max(W, 3))  # warm-up (>= 3), also grows the key buffers to their steady size
    fit.set_stats(True)
    fit.set_profiling(True)  # resets the running sums
    fit.step(K, want_mse=False)
    fit.sync()
    stats = fit.stats()
    fit.set_profiling(False)
    fit.set_stats(False)
    fit.restart()

    # ---- device-resident run: `value` ---------------------------------------------------------------------------
    fit.step(max(W, 3))
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    fit.set_profiling(True)
    launches0 = fit.launch_count()
    t0 = time.perf_counter()
    fit.step(K, want_mse=False)
    fit.sync()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = fit.last_step_ms()  # CUDA events on the library's stream around the K iterations
    prof, n_prof = fit.get_profile()
    fit.set_profiling(False)
    clocks = sampler.stop()
    launches = fit.launch_count() - launches0
    barrier()
    # ---- the same iterations once more with the forward -> backward overlap off (untimed for `value`) -----------
    # In the timed run the backward grid starts inside the forward grid's tail (programmatic dependent launch,
    # gs2d_blend.cuh), so CUDA events can only bracket the PAIR. The per-kernel durations the roofline explains
    # itself with come from this pass: same fit, same iterations, the two grids serialised.
    gs2d.set_tuning("GS2D_PDL", 0)
    fit.restart()
    fit.step(max(W, 3))
    fit.set_profiling(True)
    fit.step(K, want_mse=False)
    fit.sync()
    sep_ms = fit.last_step_ms() / K
    prof_sep, _ = fit.get_profile()
    fit.set_profiling(False)
    gs2d.set_tuning("GS2D_PDL", 1)
    barrier()
    t = torch.tensor([dev_ms, wall_ms], dtype=torch.float64, device="cuda")
    live = torch.tensor([float(stats["live_pairs_sum"]), float(stats["keys_sum"]), float(stats["n_forwards"])],
                        dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(live, op=dist.ReduceOp.SUM)
        live[2] /= world
    dev_ms, wall_ms = float(t[0]), float(t[1])
    ms_per_step = dev_ms / K
    value = 1e3 / ms_per_step
    n_fw = max(float(live[2]), 1.0)
    live_per_it, keys_per_it = float(live[0]) / n_fw, float(live[1]) / n_fw

    # ---- end to end through the C ABI with host buffers: `e2e` --------------------------------------------------
    # Every step copies its input (the RGBA8 target; this rank's band of it when sharded) from pinned host memory and
    # reads its result (the mse) back. Software pipeline of depth one, as a caller streaming targets would write it:
    # the upload of step i+1 is issued while step i runs (gs2d_set_target_rgba8_async: own copy stream, two staging
    # buffers), and the mse of step i is waited for (gs2d_mse_wait on its ticket) after step i+1 has been enqueued,
    # so the device does not idle across the host round trip. Every step's upload and read-back are inside the
    # timed region; the last step's read-back closes it.
    # Both loops run the SAME iterations of the fit as the device-resident pass (restart + the same warm-up): the
    # work per iteration drifts as the fit converges, so a pass over later iterations would not be comparable.
    Ke = max(3, min(K, args.e2e_steps)) if args.e2e_steps > 0 else K
    h2d_bytes = fit.target_bytes_per_upload()
    fit.restart()
    fit.step(max(W, 3))
    fit.set_target_u8_ptr_async(pinned_u8.data_ptr())  # pipeline fill: the target of the first timed step
    barrier()
    t0 = time.perf_counter()
    mse_last, prev = None, None
    for _ in range(Ke):
        fit.step(1, want_mse=False)                         # enqueue the step (its target was uploaded before)
        ticket = fit.mse_ticket()                           # D2H of this step's result, enqueued behind it
        fit.set_target_u8_ptr_async(pinned_u8.data_ptr())   # H2D of the next step's input, overlapping this step
        if prev is not None:
            mse_last = fit.mse_wait(prev)                   # blocks on the PREVIOUS step's result only
        prev = ticket
    mse_last = fit.mse_wait(prev)
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / Ke
    # the same without the overlap (blocking upload, then the step), for the record
    fit.restart()
    fit.step(max(W, 3))
    barrier()
    t0 = time.perf_counter()
    for _ in range(Ke):
        fit.set_target_u8_ptr(pinned_u8.data_ptr())
        fit.step(1)
    torch.cuda.synchronize()
    e2e_serial_ms = (time.perf_counter() - t0) * 1e3 / Ke
    te = torch.tensor([e2e_ms, e2e_serial_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms, e2e_serial_ms = float(te[0]), float(te[1])

    if rank != 0:
        fit.close()
        if dist is not None:
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel ------------------------------------------------------------------------
    hbm_peak = peaks["hbm_gbs"] if peaks else 6650.0
    hbm_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"
    kern = {}
    blend_ms = prof["forward"] + prof["backward"]
    if blend_ms > 0:
        # timed region: k_forward + k_backward overlapped (+ k_finalize_loss and two small memsets, ~15 us)
        a = (FLOP_FWD + FLOP_BWD) * live_per_it / world / (blend_ms * 1e-3) / 1e12
        kern["blend_overlapped"] = {"ms": blend_ms, "bound": "fp32_issue", "achieved": a, "peak": fp32_peak,
                                    "unit": "TFLOP/s", "frac": a / fp32_peak,
                                    "what": "k_forward + k_backward as launched in the timed region: the backward grid "
                                            "fills the forward grid's tail (programmatic dependent launch), CUDA "
                                            "events bracket the pair; 27 + 86 flop per live pair"}
    if prof_sep["forward"] > 0:
        a = FLOP_FWD * live_per_it / world / (prof_sep["forward"] * 1e-3) / 1e12
        kern["k_forward"] = {"ms": prof_sep["forward"], "bound": "fp32_issue", "achieved": a, "peak": fp32_peak,
                             "unit": "TFLOP/s", "frac": a / fp32_peak, "pass": "serialised (GS2D_PDL=0), untimed"}
    if prof_sep["backward"] > 0:
        a = FLOP_BWD * live_per_it / world / (prof_sep["backward"] * 1e-3) / 1e12
        kern["k_backward"] = {"ms": prof_sep["backward"], "bound": "fp32_issue", "achieved": a, "peak": fp32_peak,
                              "unit": "TFLOP/s", "frac": a / fp32_peak, "pass": "serialised (GS2D_PDL=0), untimed"}
    kern["serialised_ms_per_step"] = {"ms": sep_ms}
    if prof["sort"] > 0:
        # k_scan_* + k_place + k_rank with slot binning (gs2d_binning.cu): per bucket (3.4 entries) the 16-byte slot row,
        # the 2-byte counter twice (scan, k_rank), the 4-byte offset written and read = 30 B, i.e. 9 B per list entry,
        # + 4 B per entry written in index order
        b = keys_per_it / world * BIN_BYTES_PER_KEY
        a = b / (prof["sort"] * 1e-3) / 1e9
        kern["place_rank"] = {"ms": prof["sort"], "bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s",
                              "frac": a / hbm_peak, "bytes_per_list_entry": BIN_BYTES_PER_KEY}
    if prof["adam"] > 0:
        a = ADAM_BYTES * N / (prof["adam"] * 1e-3) / 1e9
        kern["k_adam"] = {"ms": prof["adam"], "bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s",
                          "frac": a / hbm_peak}
    kern["preprocess_emit"] = {"ms": prof["preprocess_emit"]}
    kern["allreduce"] = {"ms": prof["allreduce"]}
    dom = max((k for k in kern if "frac" in kern[k] and "pass" not in kern[k]), key=lambda k: kern[k]["ms"])
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        traffic = tj["k_forward"] + tj["k_backward"] if dom == "blend_overlapped" else tj.get(dom)
    except Exception:
        pass
    roofline = {"kernel": dom, "bound": kern[dom]["bound"], "achieved": kern[dom]["achieved"], "peak": kern[dom]["peak"],
                "unit": kern[dom]["unit"], "frac": kern[dom]["frac"], "traffic": traffic,
                "peak_source": ("gs2d_microbench_fp32 on this device (FP32 FMA issue rate; MEASURED_PEAKS.json has no "
                                "FP32 figure)") if kern[dom]["bound"] == "fp32_issue" else hbm_src,
                "algorithmic_work": "27 (fwd) / 86 (bwd) flop per live (pixel,splat) pair x live pairs counted on the device; "
                                    "traffic = dram bytes of both blend kernels per iteration (ncu, serialised capture)",
                "live_pairs_per_iteration": live_per_it, "keys_per_iteration": keys_per_it,
                "hbm_copy_gbs_this_device": hbm_peak_local, "kernels": kern}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, desc, _ = cpu_sample(S, N, args.cpu_iters)
        cpu = {"value": v, "unit": "it/s", "cores": 1, "kind": "port", "sample": desc,
               "host_cores_available": os.cpu_count()}

    line = {
        "metric": "fit_iters_per_sec", "value": value, "unit": "it/s", "n_gpus": world, "steps": K, "warmup": max(W, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
        "clocks": clocks, "wall_ms_per_step": wall_ms / K,
        "e2e": {"value": 1e3 / e2e_ms, "unit": "it/s", "h2d_bytes_per_step": h2d_bytes * world,
                "d2h_bytes_per_step": (8 + 40) * world, "ms_per_step": e2e_ms, "steps": Ke,
                "serial_ms_per_step": e2e_serial_ms,
                "what": "per step and rank: gs2d_step(1) enqueued, gs2d_set_target_rgba8_async of the next step's "
                        "target from pinned host memory (this rank's band), gs2d_mse_ticket for this step (D2H of the "
                        "loss and the 40-byte device state) and gs2d_mse_wait (blocking) for the previous one; "
                        "serial_ms_per_step = blocking gs2d_set_target_rgba8 then gs2d_step(1, &mse)"},
        "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
        "last_mse": float(mse_last[0]) if mse_last is not None else None,
    }
    _emit(line)
    fit.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def _emit(line):
    """Print the ONE JSON line on the real stdout (fd 1 is pointed at stderr while the run is in progress so that
    banners printed by native libraries, e.g. "NCCL version ...", cannot get in front of it)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=4096)
    ap.add_argument("--splats", type=int, default=1_000_000)
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end loops (0 = --steps)")
    ap.add_argument("--cpu-iters", type=int, default=10)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 40:
            args.steps = 40  # bounded: ~1 s per sample step on one core
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
