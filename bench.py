#!/usr/bin/env python
"""bench.py — fit iterations/s (forward + backward + Adam) of the B200-native 2D Gaussian-splatting engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c3|c4|c5]

Workload (default c3 = BASELINE.json metric / configs[2]): synthetic 4096x4096 RGB target, 1 M Gaussians initialised
with the reference's own init() formula (main.cpp:280-305), lr 0.05, opacity not optimised. One "step" = one whole
loop body of the reference (main.cpp:414-809): binning, forward blend, loss, backward, [gradient exchange], Adam +
clamps. N > 1 (launched by torchrun, one process per GPU): the image is split into tile-row bands, one per rank, and
the per-Gaussian gradients are exchanged before Adam -> the total work is fixed ("strong" scaling).
--config c4: 8192^2 / 4 M Gaussians (configs[3]); --config c5: 256 independent 512^2 images with 16 K Gaussians each
(configs[4]; N > 1 = replicas only, 256 / N images per rank, no collective).

Prints ONE JSON line (rank 0). `value` = device-timed iterations/s with all inputs resident in HBM, profiling OFF (the
iteration runs as the library runs it for a user: one CUDA graph per iteration on one GPU); the per-kernel-group times
the roofline explains itself with come from separate, untimed passes over the same iterations. `e2e` = the same metric
through the C ABI with HOST buffers: every step uploads the RGBA8 target from pinned host memory
(gs2d_set_target_rgba8_async, the reference's own ingest format, prp.cpp:179-193 + :796-807) and reads the mse back.

--impl reference: the reference's own CPU implementation of the path, single thread (the reference is
single-threaded: main.cpp never calls prlib's ParallelFor / ThreadPool), on the box's host cores, on the SAME config:
every warm-up and timed step is one real iteration of the full problem (4096^2 / 1 M: ~12 s each).
oracle/_ref/ref_main (the unmodified main.cpp) is hard-wired to one JPEG and 1024 splats, so the function-ised
restatement oracle/gs2d_oracle.c is timed (bit-identical to main.cpp, tests/test_oracle.py).
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

CONFIGS = {
    # name: (image side, Gaussians per image, images, BASELINE.json configs[] index)
    "c3": (4096, 1_000_000, 1, 2),
    "c4": (8192, 4_000_000, 1, 3),
    "c5": (512, 16_384, 256, 4),
}


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


def oracle_state(args):
    """The CPU oracle's loop-carried state and target for this config (one image: the reference fits one image)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_binding as ob

    gs2d = importlib.import_module("2dgaussiansplatting_b200")
    target = np.rint(gs2d.synth_target(args.size, args.size, seed=1) * 255.0).astype(np.float32) / np.float32(255.0)
    return ob.State(args.size, args.size, args.splats), np.ascontiguousarray(target.astype(np.float32))


def cpu_sample(args, iters):
    """Time the CPU oracle (the reference's algorithm, 1 thread) on REAL iterations of this config (for c5: of one of
    the 256 images). Returns (equivalent whole-workload it/s, description)."""
    st, target = oracle_state(args)
    t0 = time.perf_counter()
    for _ in range(iters):
        st.step(target)
    dt = (time.perf_counter() - t0) / iters
    if args.images > 1:
        desc = (f"{iters} iterations of ONE of the {args.images} images ({args.size}x{args.size}, {args.splats} Gaussians, "
                f"{dt:.3f} s each); value = 1 / ({args.images} x that): the images are independent")
        return 1.0 / (dt * args.images), desc
    desc = (f"{iters} real iterations of the full workload ({args.size}x{args.size}, {args.splats} Gaussians, "
            f"{dt:.2f} s each, from init(); no extrapolation)")
    return 1.0 / dt, desc


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    t_first = time.perf_counter()
    st, target = oracle_state(args)
    # every warm-up and every timed step is one real iteration of the full problem (c5: of all its images in turn
    # would take 256 x as long - there the step is one image and the value is divided, stated in `sample`)
    for _ in range(args.warmup):
        st.step(target)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st.step(target)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    per_step = dt * args.images
    value = 1.0 / per_step
    if args.images > 1:
        sample = (f"each step = one iteration of ONE of the {args.images} independent images ({dt:.3f} s); "
                  f"ms_per_step and value are scaled by {args.images}")
    else:
        sample = (f"each step = one real iteration of the full {args.size}x{args.size} / {args.splats}-Gaussian problem "
                  f"({dt:.2f} s), continuing the fit from init(); nothing extrapolated")
    line = {
        "impl": "reference", "metric": "fit_iters_per_sec", "value": value, "unit": "it/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "strong" if args.images == 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1, None),
        "cpu_baseline": {"value": value, "unit": "it/s", "cores": 1, "kind": "port", "sample": sample,
                         "host_cores_available": os.cpu_count()},
        "e2e": {"value": value, "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t_first,
    }
    _emit(line)
    return 0


def workload_config(args, n_gpus, comm_mode):
    """comm_mode: (n_ranks, peer_memory) as gs2d_comm_mode reports it for the context that ran, or None."""
    if n_gpus == 1:
        par = "single GPU"
    elif args.images > 1:
        par = f"{n_gpus} replicas x {args.images // n_gpus} independent images, no collective"
    elif comm_mode is not None and comm_mode[1]:
        par = (f"{n_gpus} tile-row bands; gradient exchange = ONE fused kernel per rank over NVLink peer memory (CUDA "
               f"IPC): reduce-scatter of the 9 gradient planes + Adam on the owned slice + all-gather of the "
               f"parameters; flag barriers over peer mailboxes; NCCL only for the rendezvous (gs2d_comm_mode: peer_memory=1)")
    else:
        par = (f"{n_gpus} tile-row bands; gradient exchange = ncclAllReduce of the 9 gradient planes + the loss, Adam "
               f"replicated (gs2d_comm_mode: peer_memory=0)")
    img = f"{args.images} x " if args.images > 1 else ""
    return {
        "workload": f"synthetic {img}{args.size}x{args.size} RGB target, {args.splats} Gaussians"
                    f"{' each' if args.images > 1 else ''} (BASELINE configs[{CONFIGS[args.config][3]}])",
        "image": [args.size, args.size], "n_gaussians": args.splats, "n_images": args.images,
        "init": "reference init() main.cpp:280-305", "optimize_opacity": False, "lr": 0.05, "parallelism": par,
        "arithmetic": "relaxed (GS2D_EXACT=0, the library default: FMA-contracted blend chain, results within FP32 "
                      "round-off of the reference; `exact_mode` in this line = the bit-exact kernels)",
        "l2_policy": "no flush: per-iteration inputs (target + image 2 x 268 MB, row table 512 MB, key lists >250 MB at "
                     "c3) exceed the 126 MB L2",
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

    S, N, K, W = args.size, args.splats, args.steps, max(args.warmup, 3)
    replicas = args.images > 1          # c5: independent images, replicas only
    n_img = args.images // world if replicas else 1
    if replicas and args.images % world:
        raise SystemExit(f"--config c5: {args.images} images do not divide over {world} ranks")
    img0 = rank * n_img if replicas else 0
    # 8-bit targets, as a decoded JPEG would be; one pinned buffer per image of this rank
    pinned = []
    for i in range(n_img):
        t = torch.empty((S, S, 4), dtype=torch.uint8, pin_memory=True)
        t.numpy()[...] = np.rint(gs2d.synth_target(S, S, seed=1 + img0 + i) * 255.0).astype(np.uint8)
        pinned.append(t)

    fit = gs2d.Fitter(S, S, N, n_images=n_img, device=local_rank)
    comm_mode = None
    if world > 1 and not replicas:
        box = [gs2d.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        fit.comm_init(rank, world, box[0])
        comm_mode = fit.comm_mode()
        if os.environ.get("GS2D_PEER_ADAM", "1") != "0" and not comm_mode[1]:
            print(f"[bench rank {rank}] WARNING: the peer-memory exchange was requested but not agreed (CUDA IPC / P2P "
                  f"unavailable?): running the NCCL all-reduce path", file=sys.stderr)

    def upload_all():
        for i, t in enumerate(pinned):
            fit.set_target_u8_ptr(t.data_ptr(), image=i)

    upload_all()
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

    def timed_pass(profile=False, stats=False):
        """restart + W warm-up iterations, then K iterations enqueued back to back behind ONE more untimed iteration
        that is already running when they are enqueued: the CUDA events around the K iterations (on the library's
        stream) then start at an iteration boundary with the device busy - and, with N ranks, behind that iteration's
        device-side barrier, i.e. aligned across the ranks. Returns device ms for the K iterations."""
        fit.restart()
        fit.step(W - 1)
        barrier()
        if stats:
            fit.set_stats(True)
        if profile:
            fit.set_profiling(True)  # resets the running sums; also keeps the iteration out of the CUDA graph
        fit.step(1, want_mse=False)
        fit.step(K, want_mse=False)
        fit.sync()
        ms = fit.last_step_ms()
        out = {"ms": ms}
        if profile:
            prof, n_prof = fit.get_profile()
            # K + 1 iterations were profiled (the aligning one included): means per iteration
            out["prof"] = prof
            fit.set_profiling(False)
        if stats:
            out["stats"] = fit.stats()
            fit.set_stats(False)
        barrier()
        return out

    # ---- counting pass (untimed): live (pixel, Gaussian) pairs and list entries of the SAME iterations ------------
    # The counting variant of the forward blend is slower, so the work counters the roofline needs are taken on an
    # identical, untimed run of the same iterations and the timed run has them off.
    fit.step(3)  # grows the key buffers to their steady size, builds the graph
    cnt = timed_pass(profile=True, stats=True)
    stats = cnt["stats"]

    # ---- device-resident run: `value` (profiling off: the library's own launch path, graph replay on one GPU) -----
    fit.restart()
    fit.step(W - 1)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    if comm_mode is not None:
        fit.comm_traffic(reset=True)
    fit.step(1, want_mse=False)
    launches0 = fit.launch_count()
    t0 = time.perf_counter()
    fit.step(K, want_mse=False)
    fit.sync()
    wall_ms = (time.perf_counter() - t0) * 1e3
    dev_ms = fit.last_step_ms()  # CUDA events on the library's stream around exactly the K iterations
    clocks = sampler.stop()
    launches = fit.launch_count() - launches0
    nvlink = None
    if comm_mode is not None:
        rd, wr, sparse = fit.comm_traffic()
        nvlink = [rd / (K + 1), wr / (K + 1), 1.0 if sparse else 0.0]  # the aligning iteration is counted too
    barrier()

    # ---- per-kernel-group times (untimed for `value`): the same iterations with CUDA events between the groups ----
    prof = timed_pass(profile=True)["prof"]
    # ... and once more with the forward -> backward overlap off: in the runs above the backward grid starts inside
    # the forward grid's tail (programmatic dependent launch, gs2d_blend.cuh), so events can only bracket the PAIR
    gs2d.set_tuning("GS2D_PDL", 0)
    sep = timed_pass(profile=True)
    prof_sep, sep_ms = sep["prof"], sep["ms"] / K
    gs2d.set_tuning("GS2D_PDL", 1)
    # ---- the bit-exact kernels (GS2D_EXACT=1), same iterations, for the record ------------------------------------
    gs2d.set_tuning("GS2D_EXACT", 1)
    exact_ms = timed_pass()["ms"]
    exact_prof = timed_pass(profile=True)["prof"]
    gs2d.set_tuning("GS2D_EXACT", 0)

    t = torch.tensor([dev_ms, wall_ms, exact_ms], dtype=torch.float64, device="cuda")
    live = torch.tensor([float(stats["live_pairs_sum"]), float(stats["keys_sum"]), float(stats["n_forwards"])],
                        dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(live, op=dist.ReduceOp.SUM)
        live[2] /= world
    dev_ms, wall_ms, exact_ms = float(t[0]), float(t[1]), float(t[2])
    exchange = None
    if nvlink is not None:
        tv = torch.tensor(nvlink, dtype=torch.float64, device="cuda")
        dist.all_reduce(tv, op=dist.ReduceOp.SUM)
        plane_mb = 9 * 4 * (N + 7) / 1e6
        exchange = {
            "path": "peer_memory" if comm_mode[1] else "nccl_allreduce", "sparse": bool(tv[2] > 0),
            "nvlink_mb_per_step_per_rank": {"gradients_read": float(tv[0]) / world / 1e6,
                                            "parameters_written": float(tv[1]) / world / 1e6},
            "dense_equivalent_mb_per_step_per_rank": {"gradients_read": plane_mb * (world - 1) / world,
                                                      "parameters_written": plane_mb * (world - 1) / world},
            "what": "payload counted by the exchange kernel itself (16-byte quads), mean over the timed iterations; "
                    "dense = every rank reads / writes its 1/N slice of all 9 planes from / to every peer",
        }
    ms_per_step = dev_ms / K
    value = 1e3 / ms_per_step
    n_fw = max(float(live[2]), 1.0)
    live_per_it, keys_per_it = float(live[0]) / n_fw, float(live[1]) / n_fw  # whole job (all ranks, all images)

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
    h2d_bytes = fit.target_bytes_per_upload() * n_img
    e2e_ms = e2e_serial_ms = None
    mse_last = None
    if n_img == 1:
        def pipelined(n_steps):
            """n_steps of the end-to-end loop; the first step's target has been uploaded (asynchronously) before."""
            prev = None
            for _ in range(n_steps):
                fit.step(1, want_mse=False)                        # enqueue the step (its target was uploaded before)
                ticket = fit.mse_ticket()                          # D2H of this step's result, enqueued behind it
                fit.set_target_u8_ptr_async(pinned[0].data_ptr())  # H2D of the next step's input, overlapping this step
                if prev is not None:
                    fit.mse_wait(prev)                             # blocks on the PREVIOUS step's result only
                prev = ticket
            return fit.mse_wait(prev)

        # the W warm-up steps go through the same loop: what a streaming caller pays once (the pinned result ring, the
        # staging buffers, one captured graph per staging buffer) is set up there, as the kernels' warm-up is
        fit.restart()
        fit.set_target_u8_ptr_async(pinned[0].data_ptr())
        pipelined(W)  # leaves the target of the first timed step uploaded: the pipeline is full
        barrier()
        t0 = time.perf_counter()
        mse_last = pipelined(Ke)
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / Ke
    # the same without the overlap (blocking upload of every image, then the step with its loss read back); for the
    # image batch (c5) this IS the end-to-end number: the asynchronous upload holds one image at a time
    fit.restart()
    fit.step(W)
    barrier()
    t0 = time.perf_counter()
    for _ in range(Ke):
        upload_all()
        mse_last = fit.step(1)[0]
    torch.cuda.synchronize()
    e2e_serial_ms = (time.perf_counter() - t0) * 1e3 / Ke
    if e2e_ms is None:
        e2e_ms = e2e_serial_ms
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
    per_rank_live = live_per_it / world

    def blend_entry(ms, flop, **kw):
        a = flop * per_rank_live / (ms * 1e-3) / 1e12
        return dict({"ms": ms, "bound": "fp32_issue", "achieved": a, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": a / fp32_peak}, **kw)

    blend_ms = prof["forward"] + prof["backward"]
    if blend_ms > 0:
        kern["blend_overlapped"] = blend_entry(
            blend_ms, FLOP_FWD + FLOP_BWD,
            what="k_forward + k_backward as the library launches them: the backward grid fills the forward grid's "
                 "tail (programmatic dependent launch), CUDA events bracket the pair; 27 + 86 flop per live pair. "
                 "Measured in an untimed pass over the same iterations (events between the kernel groups keep the "
                 "iteration out of the CUDA graph, so `value` is timed without them)")
    if prof_sep["forward"] > 0:
        kern["k_forward"] = blend_entry(prof_sep["forward"], FLOP_FWD, **{"pass": "serialised (GS2D_PDL=0), untimed"})
    if prof_sep["backward"] > 0:
        kern["k_backward"] = blend_entry(prof_sep["backward"], FLOP_BWD, **{"pass": "serialised (GS2D_PDL=0), untimed"})
    eb = exact_prof["forward"] + exact_prof["backward"]
    if eb > 0:
        kern["blend_overlapped_exact_mode"] = blend_entry(eb, FLOP_FWD + FLOP_BWD, **{"pass": "GS2D_EXACT=1, untimed"})
    kern["serialised_ms_per_step"] = {"ms": sep_ms}
    if prof["sort"] > 0:
        # k_group_scan + k_place + k_rank_groups with group rows (gs2d_binning.cu): per (group of 4 tiles, index slice)
        # pair (13 list entries) the 64-byte row and the 4-byte counter read twice (scan, rank) and four 4-byte offsets
        # written and read = ~9 B per list entry, + 4 B per entry written in index order. k_row_table runs beside these
        # kernels on a second stream.
        b = keys_per_it / world * BIN_BYTES_PER_KEY
        a = b / (prof["sort"] * 1e-3) / 1e9
        kern["place_rank"] = {"ms": prof["sort"], "bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s",
                              "frac": a / hbm_peak, "bytes_per_list_entry": BIN_BYTES_PER_KEY}
    if prof["adam"] > 0:
        a = ADAM_BYTES * N * n_img / (prof["adam"] * 1e-3) / 1e9
        kern["k_adam"] = {"ms": prof["adam"], "bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s",
                          "frac": a / hbm_peak}
    kern["preprocess_emit"] = {"ms": prof["preprocess_emit"]}
    kern["exchange"] = {"ms": prof["allreduce"],
                        "what": "barrier 1 + loss all-reduce (peer path) / gradient all-reduce (NCCL path)"}
    dom = max((k for k in kern if "frac" in kern[k] and "pass" not in kern[k]), key=lambda k: kern[k]["ms"])
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if args.config == "c3" and world == 1:
            traffic = tj["k_forward"] + tj["k_backward"] if dom == "blend_overlapped" else tj.get(dom)
            traffic_src = tj.get("source")
    except Exception:
        pass
    roofline = {"kernel": dom, "bound": kern[dom]["bound"], "achieved": kern[dom]["achieved"], "peak": kern[dom]["peak"],
                "unit": kern[dom]["unit"], "frac": kern[dom]["frac"], "traffic": traffic,
                "traffic_source": traffic_src,
                "peak_source": ("gs2d_microbench_fp32 on this device (FP32 FMA issue rate; MEASURED_PEAKS.json has no "
                                "FP32 figure)") if kern[dom]["bound"] == "fp32_issue" else hbm_src,
                "algorithmic_work": "27 (fwd) / 86 (bwd) flop per live (pixel,splat) pair x live pairs counted on the device; "
                                    "traffic = dram bytes of both blend kernels per iteration from a committed ncu capture "
                                    "(see traffic_source), not measured in this run",
                "live_pairs_per_iteration": live_per_it, "keys_per_iteration": keys_per_it,
                "hbm_copy_gbs_this_device": hbm_peak_local, "kernels": kern}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        v, desc = cpu_sample(args, args.cpu_iters)
        cpu = {"value": v, "unit": "it/s", "cores": 1, "kind": "port", "sample": desc,
               "host_cores_available": os.cpu_count()}

    line = {
        "metric": "fit_iters_per_sec", "value": value, "unit": "it/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if replicas else "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, world, comm_mode),
        "clocks": clocks, "wall_ms_per_step": wall_ms / K,
        "e2e": {"value": 1e3 / e2e_ms, "unit": "it/s", "h2d_bytes_per_step": h2d_bytes * world,
                "d2h_bytes_per_step": (8 * n_img + 48) * world, "ms_per_step": e2e_ms, "steps": Ke,
                "serial_ms_per_step": e2e_serial_ms,
                "what": "per step and rank: gs2d_step(1) enqueued, gs2d_set_target_rgba8_async of the next step's "
                        "target from pinned host memory (this rank's band), gs2d_mse_ticket for this step (D2H of the "
                        "loss and the 48-byte device state) and gs2d_mse_wait (blocking) for the previous one; the W "
                        "warm-up steps run through the same loop (untimed); wall clock over the K steps; "
                        "serial_ms_per_step = blocking gs2d_set_target_rgba8 then gs2d_step(1, &mse)"},
        "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "exchange": exchange,
        "exact_mode": {"value": 1e3 * K / exact_ms, "unit": "it/s", "ms_per_step": exact_ms / K,
                       "what": "GS2D_EXACT=1: blend kernels in the reference's operand order without FMA contraction "
                               "(image bit-identical to the CPU reference); same iterations, device-timed like `value`"},
        "last_mse": float(mse_last[0]) if mse_last is not None else None,
    }
    if replicas:
        line["image_iters_per_sec"] = value * args.images
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c3", choices=sorted(CONFIGS), help="BASELINE.json workload (default: the metric's)")
    ap.add_argument("--size", type=int, default=0, help="override the image side (development)")
    ap.add_argument("--splats", type=int, default=0, help="override the Gaussians per image (development)")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end loops (0 = --steps)")
    ap.add_argument("--cpu-iters", type=int, default=0, help="iterations of the cpu_baseline sample (0 = per config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    side, n, images, _ = CONFIGS[args.config]
    args.size = args.size or side
    args.splats = args.splats or n
    args.images = images
    if args.cpu_iters <= 0:
        # ~10-30 s of CPU work: c3 ~12 s per iteration, c4 ~50 s, c5 ~0.2 s per image-iteration
        args.cpu_iters = {"c3": 2, "c4": 1, "c5": 20}[args.config]
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
