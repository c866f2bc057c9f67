"""Band-sharded multi-GPU fit == single-GPU fit (needs >= 2 B200s on the box; skipped otherwise)."""
import os
import subprocess
import sys

import pytest

from gs2d_loader import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("peer,sparse,iters", [("1", "1", "12"), ("1", "1", "60"), ("1", "0", "12"), ("0", "1", "12")])
def test_two_ranks_match_single_gpu(peer, sparse, iters):
    """peer=1: fused reduce + Adam + broadcast over NVLink peer memory - sparse (only what a rank can have touched /
    will look at; also over 60 iterations, while Gaussians drift across the band boundary) or dense;
    peer=0: NCCL all-reduce of the gradients. Replayed as a CUDA graph (peer path) like a user's run."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "mgpu_check.py")]
    env = dict(os.environ, GS2D_PEER_ADAM=peer, GS2D_PEER_SPARSE=sparse, MGPU_ITERS=iters)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "replicated parameters identical on all ranks: True" in r.stdout
    if peer == "0":
        assert "peer_memory) = (2, False)" in r.stdout
