"""Band-sharded multi-GPU fit == single-GPU fit (needs >= 2 B200s on the box; skipped otherwise)."""
import os
import subprocess
import sys

import pytest

from gs2d_loader import ROOT

pytestmark = pytest.mark.gpu


def test_two_ranks_match_single_gpu():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tools", "mgpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "replicated parameters identical on all ranks: True" in r.stdout
