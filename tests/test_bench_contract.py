"""CPU test of bench.py's reference arm (the driver runs `bench.py --impl reference --gpus N --steps K --warmup W` on the
GPU box's host cores): one JSON line with the contract's keys, steps / warmup honoured literally (every step a real
iteration of the named problem - here shrunk through the development overrides --size / --splats), and under a
multi-rank launch only rank 0 works."""
import json
import os
import subprocess
import sys

from gs2d_loader import ROOT


def run(extra, env=None):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--size", "96", "--splats",
                        "300"] + extra, capture_output=True, text=True, timeout=300, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout


def test_reference_arm_line():
    out = run(["--steps", "3", "--warmup", "2"])
    lines = [l for l in out.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "fit_iters_per_sec" and d["unit"] == "it/s"
    assert d["steps"] == 3 and d["warmup"] == 2 and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["value"] > 0 and abs(d["value"] - 1e3 / d["ms_per_step"]) < 1e-9 * d["value"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == 1 and cb["value"] == d["value"] and "real iteration" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["image"] == [96, 96] and d["config"]["n_gaussians"] == 300
    # the timed region fits inside the process's own wall time (no extrapolation)
    assert d["ms_per_step"] * 1e-3 * d["steps"] <= d["wall_s"]


def test_reference_arm_other_ranks_exit_quietly():
    out = run(["--steps", "1", "--warmup", "0", "--gpus", "2"], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert out.strip() == ""


def test_c5_reference_arm_scales_one_image_to_the_batch():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c5", "--size", "64",
                        "--splats", "200", "--steps", "2", "--warmup", "0"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip())
    assert d["config"]["n_images"] == 256 and d["scaling"] == "weak" and "ONE of the 256" in d["cpu_baseline"]["sample"]
