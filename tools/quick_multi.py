#!/usr/bin/env python
"""Multi-rank timing probe (developer tool; launch with torch.distributed.run): C3 fit, N ranks, K iterations; prints the
device time per iteration (max over ranks). With GS2D_PEER_DEBUG=1 the library prints the breakdown of the peer update."""
import importlib, os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
gs2d = importlib.import_module("2dgaussiansplatting_b200")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
S, N, K = int(os.environ.get("QM_SIZE", 4096)), int(os.environ.get("QM_SPLATS", 1_000_000)), int(os.environ.get("QM_ITERS", 40))
tgt = np.random.default_rng(0).integers(0, 256, (S, S, 4), dtype=np.uint8)
box = [gs2d.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(box, src=0)
with gs2d.Fitter(S, S, N, device=local) as fit:
    fit.comm_init(rank, world, box[0])
    fit.set_target_u8(tgt)
    fit.restart()
    fit.step(5)
    if os.environ.get("QM_PROFILE"):
        fit.set_profiling(True)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    fit.step(1, want_mse=False)
    fit.step(K, want_mse=False)
    fit.sync()
    t = torch.tensor([fit.last_step_ms() / K], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if os.environ.get("QM_PROFILE"):
        prof, n = fit.get_profile()
        if rank == 0:
            print("groups (rank 0):", {k: round(v, 4) for k, v in prof.items()})
    if rank == 0:
        print(f"N={world} {S}x{S} {N}: {float(t[0]):.4f} ms/it = {1e3 / float(t[0]):.1f} it/s  [{os.environ.get('QM_TAG', '')}]", flush=True)
dist.destroy_process_group()
