"""GPU parity tests: the CUDA path, called through the C ABI (include/gs2d.h via the ctypes host layer), against the
CPU oracle (oracle/gs2d_oracle.c, proven bit-identical to the reference's main.cpp in test_oracle.py) and against the
golden fixtures produced by the unmodified reference.

Tolerances (north star: "within a stated FP32 relative tolerance, blend order preserved"):
  * initial state, footprint decisions, transmittance chain, colour accumulation, Adam: the kernels use the
    reference's operand order without FMA contraction => the rendered image is BIT-IDENTICAL wherever cosf/sinf of
    the device (double sincos rounded once) equal glibc's; we require >= 99.9 % of pixels bit-identical and the rest
    within IMG_ATOL.
  * gradients: same per-pair terms, but summed over pixels by warp shuffles + atomics instead of sequentially =>
    |g_gpu - g_cpu| <= GRAD_RTOL * sum|terms| ~ GRAD_RTOL * (|g_cpu| + mean|g|).
  * loss curve: free-running, chaotic (Adam's first step is +-lr); relative 1e-4 over the first 50 iterations.
"""
import os

import numpy as np
import pytest

import oracle_binding as ob
from gs2d_loader import gs2d

pytestmark = pytest.mark.gpu

IMG_ATOL = 1.0 / 256.0 * 1.01  # one flipped decision changes a pixel by at most T*alpha <= MIN_THROUGHPUT-ish
GRAD_RTOL = 2e-4
# Relaxed arithmetic (GS2D_EXACT=0, the library's default; gs2d_forward.cu): alpha = ( s (1 - d2/16) )^8 with FMA
# contraction and s = opacity^(1/8) folded into the inverse covariance: ~3 ulp on the argument, x8 through the power
# => alpha within ~2e-6 relative; colour = sum of <= ~100 such terms in [0, 1].
# Measured on a B200 (gpurun_out/r2a_pytest.log): image q99.99 <= 4.5e-7 on every configuration but the edge cases,
# where a 1024 x 1 needle (d2 = a difference of terms of size 1e3) reaches 2e-5; gradients <= 3e-5 relative on the
# reference's configurations, 8e-4 on 50 Gaussians of 100..1024 px; loss curve of the 268 x 213 fit 7e-10 at iteration
# 0, 1.1e-6 over 50 iterations (exact mode: 6.5e-7).
RELAXED_IMG_TOL = 2e-6       # 99.99 % of the pixels; the rest is bounded by one flipped T < 1/256 decision (IMG_ATOL)
RELAXED_IMG_TOL_NEEDLES = 1e-4  # ill-conditioned Gaussians (aspect ratio ~1000)
RELAXED_GRAD_RTOL = 1e-3     # |dg| <= rtol * (|g| + mean|g|)
MODES = ["exact", "relaxed"]


@pytest.fixture(params=MODES)
def mode(request):
    """Run the test once per arithmetic mode of the blend kernels; the default (relaxed) is restored afterwards."""
    gs2d.set_tuning("GS2D_EXACT", 1 if request.param == "exact" else 0)
    yield request.param
    gs2d.set_tuning("GS2D_EXACT", 0)


@pytest.fixture
def exact():
    gs2d.set_tuning("GS2D_EXACT", 1)
    yield
    gs2d.set_tuning("GS2D_EXACT", 0)


def grad_rtol(mode, exact_rtol=GRAD_RTOL):
    return exact_rtol if mode == "exact" else max(exact_rtol, RELAXED_GRAD_RTOL)


_libm = None


def safe_rotations(rng, n, lo=-7.0, hi=7.0):
    """Rotations whose glibc sinf / cosf (what the reference and the oracle call) are the correctly rounded values.
    The device evaluates sincos in double and rounds once, i.e. it returns the correctly rounded float; this
    container's glibc differs from that in the last bit for ~2.5 % of the arguments (sin or cos), and every such
    Gaussian perturbs the alpha of all its pixels by ~1e-7 relative - THE reason the bit-identical fraction of an
    image falls with the number of Gaussians per pixel (0.98 at 17 live per pixel, 0.90 at 80). With these rotations
    the device and the oracle start from identical sin / cos and the image must be bit-identical almost everywhere."""
    import ctypes

    global _libm
    if _libm is None:
        _libm = ctypes.CDLL("libm.so.6")
        for f in (_libm.sinf, _libm.cosf):
            f.restype, f.argtypes = ctypes.c_float, [ctypes.c_float]
    out = np.empty(0, np.float32)
    while len(out) < n:
        x = rng.uniform(lo, hi, 2 * n).astype(np.float32)
        s = np.array([_libm.sinf(float(v)) for v in x], np.float32)
        c = np.array([_libm.cosf(float(v)) for v in x], np.float32)
        ok = (s == np.sin(x.astype(np.float64)).astype(np.float32)) & (c == np.cos(x.astype(np.float64)).astype(np.float32))
        out = np.concatenate([out, x[ok]])
    return out[:n]


def load_target(golden_dir, cfg):
    rgb = np.load(os.path.join(golden_dir, f"{cfg}_target_rgb8.npz"))["rgb"]
    h, w, _ = rgb.shape
    rgba8 = np.concatenate([rgb, np.full((h, w, 1), 255, np.uint8)], axis=2)
    return w, h, rgba8


def random_splats(rng, W, H, n, smin=2.0, smax=24.0, opacity=(0.1, 1.0), safe_rot=False):
    s = np.zeros(n, ob.SPLAT_DTYPE)
    s["pos"][:, 0] = rng.uniform(-8, W + 8, n)
    s["pos"][:, 1] = rng.uniform(-8, H + 8, n)
    s["sx"] = rng.uniform(smin, smax, n)
    s["sy"] = rng.uniform(smin, smax, n)
    s["rot"] = safe_rotations(rng, n) if safe_rot else rng.uniform(-7, 7, n)
    s["color"] = rng.uniform(0, 1, (n, 3))
    s["opacity"] = rng.uniform(*opacity, n)
    return s


def assert_image_close(img_gpu, img_cpu, min_exact=0.98, mode="exact", relaxed_tol=RELAXED_IMG_TOL):
    """exact mode: >= min_exact of the pixels bit-identical, >= 99.99 % within 4 ulp of 1.0 (a last-bit difference of
    cosf/sinf on one of the covering splats), and nothing beyond one flipped coverage / liveness decision (IMG_ATOL).
    relaxed mode: >= 99.99 % within RELAXED_IMG_TOL, nothing beyond IMG_ATOL."""
    assert img_gpu.shape == img_cpu.shape
    assert (img_gpu[..., 3] == 1.0).all()  # w reset to 1, main.cpp:543-546
    same = (img_gpu[..., :3] == img_cpu[..., :3]).all(axis=2)
    diff = np.abs(img_gpu[..., :3] - img_cpu[..., :3]).max(axis=2)
    d = diff.max()
    print(f"[{mode}] image: bit-identical {same.mean():.5f}, q99.99 {np.quantile(diff, 0.9999):.3e}, max {d:.3e}")
    if mode == "exact":
        assert same.mean() >= min_exact, f"only {same.mean():.5f} of the pixels are bit-identical (max diff {d})"
        assert (diff <= 5e-7).mean() >= 0.9999, (diff <= 5e-7).mean()
    else:
        assert (diff <= relaxed_tol).mean() >= 0.9999, ((diff <= relaxed_tol).mean(), np.quantile(diff, 0.9999))
    assert d <= IMG_ATOL, d
    return same.mean(), d


def assert_grads_close(g_gpu, g_cpu, rtol=GRAD_RTOL):
    worst = 0.0
    for f in ("pos", "sx", "sy", "rot", "color", "opacity"):
        a, b = g_gpu[f].astype(np.float64), g_cpu[f].astype(np.float64)
        scale = np.abs(b).mean() + 1e-30
        err = np.abs(a - b) / (np.abs(b) + scale)
        worst = max(worst, err.max())
        print(f"grad {f}: max rel err {err.max():.3e} (rtol {rtol:.1e})")
        assert err.max() <= rtol, f"{f}: max rel err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"
    return worst


def test_init_matches_reference_init():
    for W, H, n in [(268, 213, 1024), (535, 426, 1024), (4096, 4096, 5000)]:
        with gs2d.Fitter(W, H, n) as fit:
            fit.restart()
            s, a, b1, b2, it = fit.download()
        ref = ob.init_splats(W, H, n)
        assert s.tobytes() == ref.tobytes()
        assert (b1, b2, it) == (1.0, 1.0, 0) and not a.view(np.float32).any()


@pytest.mark.parametrize("cfg", ["c1", "c2"])
def test_forward_backward_reference_jpegs(golden_dir, cfg, mode):
    W, H, rgba8 = load_target(golden_dir, cfg)
    target = ob.u8_to_target(rgba8)
    with gs2d.Fitter(W, H, 1024) as fit:
        fit.set_target_u8(rgba8)  # u8/255 conversion on the device, prp.cpp:796-807
        fit.restart()
        fit.forward()
        img = fit.image()
        mse = fit.last_mse()[0]
        stats = fit.stats()
        fit.backward()
        g = fit.grads()
        assert fit.stats()["live_pairs_backward"] == stats["live_pairs"]  # both passes walk the same live set
    splats = ob.init_splats(W, H, 1024)
    img_cpu, visited, live = ob.forward(W, H, splats)
    exact, d = assert_image_close(img, img_cpu, mode=mode)
    if cfg == "c1":  # the committed render of the unmodified reference
        ref0 = np.load(os.path.join(golden_dir, "c1_image_itr0.npz"))["rgb"]
        if mode == "exact":
            assert (img[..., :3] == ref0).all(axis=2).mean() >= 0.98
        else:
            assert np.abs(img[..., :3] - ref0).max() <= RELAXED_IMG_TOL
    mse_cpu = ob.mse(W, H, img_cpu, target)
    if mode == "exact":
        assert abs(mse - mse_cpu) <= 1e-9 * mse + 1e-6 * (1 - exact) * 65025
    else:
        assert abs(mse - mse_cpu) <= 2e-6 * mse_cpu, (mse, mse_cpu)
    assert abs(stats["live_pairs"] - live) <= 1e-4 * live  # same live set => same pair count
    g_cpu = ob.backward(W, H, splats, img_cpu, target)
    assert_grads_close(g, g_cpu, rtol=grad_rtol(mode))


def test_loss_curve_matches_reference_stdout(golden_dir, mode):
    """Free-running fit of the reference's default config against the mse lines the unmodified main.cpp printed."""
    W, H, rgba8 = load_target(golden_dir, "c1")
    ref = [float(l.split("mse")[1]) for l in open(os.path.join(golden_dir, "c1_mse.txt")).read().splitlines()]
    with gs2d.Fitter(W, H, 1024) as fit:
        fit.set_target_u8(rgba8)
        fit.restart()
        mse = fit.step(100)[:, 0]
        _, _, b1, b2, it = fit.download()
    assert it == 100
    assert b1 == pytest.approx(0.9 ** 100, rel=1e-5) and b2 == pytest.approx(0.99 ** 100, rel=1e-5)
    rel = np.abs(mse - np.array(ref[:100])) / np.array(ref[:100])
    print(f"[{mode}] loss curve: rel[0] {rel[0]:.2e}, max over 50 {rel[:50].max():.2e}, over 100 {rel.max():.2e}")
    if mode == "exact":
        assert "%d itr, mse %.4f" % (0, mse[0]) == "0 itr, mse %.4f" % ref[0]
    else:
        assert rel[0] <= 2e-6, rel[0]
    assert rel[:50].max() <= 1e-4, rel[:50].max()
    assert rel.max() <= 2e-3, rel.max()


def test_single_step_from_identical_state(mode):
    """One whole iteration (forward, backward, Adam, clamps) from the same uploaded state, incl. Adam moments."""
    rng = np.random.default_rng(3)
    W, H, n = 200, 150, 700
    target = gs2d.synth_target(W, H, seed=5)
    st = ob.State(W, H, splats=random_splats(rng, W, H, n))
    for _ in range(3):  # give the Adam moments some history on the oracle side
        st.step(target)
    p = ob.default_params(optimize_opacity=True)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target(target)
        fit.optimize_opacity = True
        fit.upload(st.splats, st.adam, float(st.beta[0]), float(st.beta[1]), st.iteration)
        mse_gpu = fit.step(1)[0, 0]
        s, a, b1, b2, it = fit.download()
    mse_cpu = st.step(target, p)
    assert abs(mse_gpu - mse_cpu) <= (1e-7 if mode == "exact" else 2e-6) * mse_cpu
    assert (b1, b2, it) == (float(st.beta[0]), float(st.beta[1]), st.iteration)
    # Adam turns gradient rounding noise into parameter noise only where |g| is tiny; everywhere else the update
    # agrees to float precision.
    for f, tol in (("pos", 2e-4), ("sx", 2e-4), ("sy", 2e-4), ("rot", 2e-4), ("color", 2e-4), ("opacity", 2e-4)):
        d = np.abs(s[f].astype(np.float64) - st.splats[f])
        assert np.quantile(d, 0.99) <= tol, (f, np.quantile(d, 0.99))
        assert d.max() <= 2.0 * 0.05 + 1e-6, (f, d.max())  # worst case: a sign flip of a +-lr first step
    dm = np.abs(a.view(np.float32) - st.adam.view(np.float32))
    assert np.quantile(dm, 0.99) <= 1e-3 * np.abs(st.adam.view(np.float32)).max()


@pytest.mark.parametrize("W,H,n,smin,smax", [(97, 61, 300, 1.0, 12.0), (320, 240, 2000, 3.0, 40.0), (64, 64, 50, 100.0, 1024.0)])
def test_random_states(W, H, n, smin, smax, mode):
    """Ragged image sizes, splats hanging over the border, wide scale range, random opacity / rotation.
    Rotations are drawn from the arguments for which glibc's sinf / cosf are correctly rounded (safe_rotations), so the
    device and the oracle work from identical sin / cos and the exact mode must be bit-identical on (nearly) every
    pixel - the 0.90 this test used to accept was the last-bit sinf / cosf noise of ~2.5 % of the Gaussians spread
    over the up to 80 Gaussians that cover a pixel here."""
    rng = np.random.default_rng(W * 7 + n)
    target = gs2d.synth_target(W, H, seed=2)
    splats = random_splats(rng, W, H, n, smin, smax, safe_rot=True)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target(target)
        fit.upload(splats)
        fit.forward()
        img = fit.image()
        fit.backward()
        g = fit.grads()
    img_cpu, _, _ = ob.forward(W, H, splats)
    assert_image_close(img, img_cpu, min_exact=0.999, mode=mode)
    g_cpu = ob.backward(W, H, splats, img_cpu, target)
    assert_grads_close(g, g_cpu, rtol=grad_rtol(mode, 5e-4) * (3 if smax > 100 else 1))


def test_random_states_libm_rotations(exact):
    """The same with arbitrary rotations: what glibc's last-bit sinf / cosf differences cost in bit-identity."""
    W, H, n = 320, 240, 2000
    rng = np.random.default_rng(W * 7 + n)
    splats = random_splats(rng, W, H, n, 3.0, 40.0)
    with gs2d.Fitter(W, H, n) as fit:
        fit.upload(splats)
        fit.forward()
        img = fit.image()
    img_cpu, _, _ = ob.forward(W, H, splats)
    assert_image_close(img, img_cpu, min_exact=0.90)


def test_edge_cases(mode):
    """Empty splat set; splats entirely off-image; degenerate covariance (NaN bounds); exact alpha = 1."""
    W, H = 80, 48
    target = gs2d.synth_target(W, H, seed=9)
    with gs2d.Fitter(W, H, 0) as fit:
        fit.set_target(target)
        fit.restart()
        mse = fit.step(2)
        img = fit.image()
    assert not img[..., :3].any() and (img[..., 3] == 1).all()
    assert mse[0, 0] == pytest.approx(ob.mse(W, H, np.concatenate([np.zeros((H, W, 3), np.float32), np.ones((H, W, 1), np.float32)], 2), target), rel=1e-12)

    s = np.zeros(6, ob.SPLAT_DTYPE)
    s["pos"] = [[-500, -500], [20, 15], [20.5, 15.5], [79.5, 47.5], [40, 24], [1000, 20]]
    s["sx"] = [8, 0, 5, 1024, 3, 8]
    s["sy"] = [8, 0, 5, 1, 3, 8]
    s["rot"] = [safe_rotations(np.random.default_rng(1), 1, 0.2, 0.4)[0], 0.0, 0.0,
                safe_rotations(np.random.default_rng(2), 1, 0.9, 1.1)[0], 0.0, 0.0]
    s["color"] = 0.7
    s["opacity"] = 1.0
    with gs2d.Fitter(W, H, len(s)) as fit:
        fit.set_target(target)
        fit.upload(s)
        fit.forward()
        img = fit.image()
        fit.backward()
        g = fit.grads()
    img_cpu, _, _ = ob.forward(W, H, s)
    assert_image_close(img, img_cpu, min_exact=0.999, mode=mode, relaxed_tol=RELAXED_IMG_TOL_NEEDLES)
    g_cpu = ob.backward(W, H, s, img_cpu, target)
    ok = np.isfinite(g_cpu.view(np.float32)).reshape(len(s), 9).all(axis=1)
    assert_grads_close(g[ok], g_cpu[ok], rtol=grad_rtol(mode, 1e-3))


def test_ellipse_culling_is_a_superset_of_the_footprint():
    """Binning with the per-tile-row ellipse spans must render bit-identical images to binning with the conservative
    rectangle (any dropped (tile, splat) pair that covers a pixel changes that pixel), with fewer list entries."""
    W, H, n = 700, 520, 6000
    rng = np.random.default_rng(11)
    cases = [random_splats(rng, W, H, n, smin=2, smax=14), random_splats(rng, W, H, 400, smin=20, smax=160),
             random_splats(rng, W, H, 3000, smin=1, smax=40)]
    cases[2]["sy"][:] = np.random.default_rng(5).uniform(1, 2.5, 3000)  # needles at every angle
    for splats in cases:
        out = []
        for cull in (0, 1):
            gs2d.set_tuning("GS2D_CULL", cull)
            try:
                with gs2d.Fitter(W, H, len(splats)) as fit:
                    fit.set_target(gs2d.synth_target(W, H, seed=3))
                    fit.upload(splats)
                    fit.forward()
                    out.append((fit.image(), fit.stats()))
            finally:
                gs2d.set_tuning("GS2D_CULL", 1)
        (img0, st0), (img1, st1) = out
        assert img0.tobytes() == img1.tobytes()
        assert st0["live_pairs"] == st1["live_pairs"]
        assert st1["keys"] < st0["keys"]


def test_slot_binning_equals_the_triple_path():
    """The three binning paths must build the same ordered lists: GS2D_SLOTS=0 (every pair a triple through k_place),
    1 (the first 8 arrivals of a (tile, slice) bucket stored into the bucket's slot row) and 2 (the default: one key
    per (Gaussian, group of 4 tiles) in a 16-slot row, sorted once per group, gs2d_bins.cuh) - bit-identical images,
    same live pairs and list entries, the same loss after three iterations - with a few entries per bucket, with
    hundreds (every row overflows), with an image width whose tile rows straddle the groups of four tiles, and on a
    rank's band of tile rows."""
    rng = np.random.default_rng(21)
    cases = [(700, 520, random_splats(rng, 700, 520, 6000, smin=2, smax=14), None),
             (700, 520, random_splats(rng, 700, 520, 1500, smin=100, smax=900, opacity=(0.1, 0.3)), None),
             (330, 250, random_splats(rng, 330, 250, 5000, smin=1, smax=30), None),   # 21 tiles per tile row
             (40, 400, random_splats(rng, 40, 400, 3000, smin=1, smax=20), None),     # 3 tiles per tile row
             (700, 520, random_splats(rng, 700, 520, 6000, smin=2, smax=30), (1, 3)),
             (330, 250, random_splats(rng, 330, 250, 5000, smin=1, smax=30), (1, 3))]  # band starts inside a group (tile 126)
    for W, H, splats, band in cases:
        out = []
        for slots in (0, 1, 2):
            gs2d.set_tuning("GS2D_SLOTS", slots)
            try:
                with gs2d.Fitter(W, H, len(splats)) as fit:
                    fit.set_target(gs2d.synth_target(W, H, seed=3))
                    fit.upload(splats)
                    if band:
                        fit.set_band(*band)
                    fit.forward()
                    img, st = fit.image(), fit.stats()
                    mse = fit.step(3)[:, 0]
                    out.append((img, st, mse))
            finally:
                gs2d.set_tuning("GS2D_SLOTS", 2)
        img0, st0, mse0 = out[0]
        assert st0["keys"] > 0
        for img, st, mse in out[1:]:
            assert img0.tobytes() == img.tobytes()
            assert st0["live_pairs"] == st["live_pairs"] and st0["keys"] == st["keys"]
            assert mse0[0] == mse[0] and np.allclose(mse0, mse, rtol=1e-6)


def test_row_table_equals_the_in_kernel_solve():
    """k_row_table solves every (Gaussian, image row) once with hand-scheduled IEEE division / square root sequences
    (row_span_dense); the blend's own in-kernel solve (row_mask, the path of Gaussians taller than 128 rows and of
    GS2D_ROWTAB=0) uses div.rn / sqrt.rn. Both must give the same coverage words: bit-identical images, same live
    pairs - round Gaussians, needles at every angle, tiny ones, ones hanging over the border, mixed with tall ones."""
    W, H = 700, 520
    rng = np.random.default_rng(31)
    cases = [random_splats(rng, W, H, 6000, smin=2, smax=14), random_splats(rng, W, H, 3000, smin=1, smax=40),
             random_splats(rng, W, H, 800, smin=1, smax=60)]
    cases[1]["sy"][:] = rng.uniform(1, 2.5, 3000)  # needles
    cases[2]["sx"][::7] = rng.uniform(100, 900, len(cases[2]["sx"][::7]))  # some far too tall for the table
    for splats in cases:
        out = []
        for rowtab in (0, 1):
            gs2d.set_tuning("GS2D_ROWTAB", rowtab)
            try:
                with gs2d.Fitter(W, H, len(splats)) as fit:
                    fit.set_target(gs2d.synth_target(W, H, seed=3))
                    fit.upload(splats)
                    fit.forward()
                    img, st = fit.image(), fit.stats()
                    mse = fit.step(3)[:, 0]
                    out.append((img, st, mse))
            finally:
                gs2d.set_tuning("GS2D_ROWTAB", 1)
        (img0, st0, mse0), (img1, st1, mse1) = out
        assert img0.tobytes() == img1.tobytes()
        assert st0["live_pairs"] == st1["live_pairs"] and st0["keys"] == st1["keys"]
        assert mse0[0] == mse1[0] and np.allclose(mse0, mse1, rtol=1e-6)


def test_binning_switches_paths_when_slot_rows_overflow():
    """A context whose Gaussians overflow nearly every slot row switches to the triple path at its first
    synchronisation point (one atomic per overflow arrival on a single counter is the slow case of slot binning) and
    must give the same loss curve as GS2D_SLOTS=0; a normal fit stays on the slot path. tools/adapt_check.py."""
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "adapt_check.py")], capture_output=True, text=True,
                       timeout=300, env=dict(os.environ, GS2D_DEBUG_BINNING="1"))
    assert r.returncode == 0 and "adapt_check OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
    assert "slot binning off at iteration 1" in r.stderr, r.stderr[-2000:]


def test_forward_backward_overlap_equals_the_serialised_launch():
    """gs2d_step launches the backward grid inside the forward grid's tail (programmatic dependent launch + per-tile
    flags, gs2d_blend.cuh). The fit must not depend on it: first loss bit-identical, the curve within the noise of the
    unordered gradient atomics, with the work counters on and off, replayed as a graph and as plain launches."""
    W, H, n = 1024, 768, 60000
    target = gs2d.synth_target(W, H, seed=5)
    runs = {}
    try:
        for pdl in (0, 1):
            for stats in (True, False):
                for graph in (1, 0):
                    gs2d.set_tuning("GS2D_PDL", pdl)
                    gs2d.set_tuning("GS2D_GRAPH", graph)
                    with gs2d.Fitter(W, H, n) as fit:
                        fit.set_target(target)
                        fit.set_stats(stats)
                        fit.restart()
                        mse = fit.step(12)[:, 0]
                        runs[(pdl, stats, graph)] = (mse, fit.image())
    finally:
        gs2d.set_tuning("GS2D_PDL", 1)
        gs2d.set_tuning("GS2D_GRAPH", 1)
    ref_mse, _ = runs[(0, True, 0)]
    for key, (mse, img) in runs.items():
        assert mse[0] == ref_mse[0], key
        assert np.max(np.abs(mse - ref_mse) / ref_mse) < 1e-6, key
        assert np.isfinite(img).all() and (img[..., 3] == 1.0).all(), key


def test_overlap_with_mostly_empty_tiles():
    """empty tiles' backward warps leave without waiting for their tile flag, so the ordering of whatever
    follows the backward grid (k_finalize_loss sums the loss partials of ALL forward warps) rests on the
    griddepcontrol.wait every backward warp ends with. 60 Gaussians on 1024 x 1024 (4096 tiles, > 95 % of them
    empty): the loss of every iteration with the overlap on must equal the serialised launch's."""
    W = H = 1024
    n = 60
    rng = np.random.default_rng(8)
    splats = random_splats(rng, W, H, n, smin=3, smax=12)
    target = gs2d.synth_target(W, H, seed=6)
    runs = {}
    try:
        for pdl in (0, 1):
            for graph in (1, 0):
                gs2d.set_tuning("GS2D_PDL", pdl)
                gs2d.set_tuning("GS2D_GRAPH", graph)
                with gs2d.Fitter(W, H, n) as fit:
                    fit.set_target(target)
                    fit.upload(splats)
                    runs[(pdl, graph)] = fit.step(40)[:, 0]
    finally:
        gs2d.set_tuning("GS2D_PDL", 1)
        gs2d.set_tuning("GS2D_GRAPH", 1)
    ref = runs[(0, 0)]
    assert ref[0] > 0
    for key, mse in runs.items():
        assert mse[0] == ref[0], key
        assert np.max(np.abs(mse - ref) / ref) < 1e-6, key


def test_nonfinite_freezes_the_fit_like_the_reference_abort():
    """The reference abort()s in the iteration whose Adam step produced a non-finite parameter, after the whole update
    loop and before that iteration's printf (main.cpp:752-785, :807). The device mirrors it: the iteration counter stays
    at the failing iteration's index + 1 and nothing enqueued behind it changes the state."""
    W, H, n = 64, 64, 16
    s = ob.init_splats(W, H, n)
    target = gs2d.synth_target(W, H)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target(target)
        fit.upload(s)
        fit.step(3)
        good, _, _, _, it = fit.download()
        assert it == 3
        bad = good.copy()
        bad["rot"][5] = np.nan  # no footprint; Adam leaves the NaN in place: flagged by the next step (main.cpp:776-779)
        fit.upload(bad, iteration=3)
        with pytest.raises(gs2d.Gs2dError) as e:
            fit.step(10)  # ten iterations enqueued at once: only the first may take effect
        assert e.value.code == gs2d.GS2D_ERR_NONFINITE
        s1, _, b1, b2, it1 = fit.download()
        assert it1 == 4  # frozen at the failing iteration's index + 1
        with pytest.raises(gs2d.Gs2dError):
            fit.step(5)
        s2, _, _, _, it2 = fit.download()
        assert it2 == 4 and s2.tobytes() == s1.tobytes()
        # the parameters are the ones one Adam step from `bad` gives (the reference holds them when it aborts)
        o = ob.State(W, H, splats=bad)
        o.step(target)
        ok = np.isfinite(o.splats.view(np.float32)).reshape(n, 9).all(axis=1)
        assert np.quantile(np.abs(s1[ok].view(np.float32) - o.splats[ok].view(np.float32)), 0.99) <= 2e-4
        # a fresh state un-freezes the context
        fit.upload(good, iteration=3)
        assert fit.step(2).shape == (2, 1) and fit.download()[4] == 5


def test_nonfinite_is_reported_not_aborted():
    W, H = 64, 64
    s = ob.init_splats(W, H, 8)
    s["rot"][3] = np.nan  # the reference would abort() at main.cpp:776-779
    with gs2d.Fitter(W, H, 8) as fit:
        fit.set_target(gs2d.synth_target(W, H))
        fit.upload(s)
        with pytest.raises(gs2d.Gs2dError) as e:
            fit.step(1)
        assert e.value.code == gs2d.GS2D_ERR_NONFINITE


def test_capacity_growth_with_huge_splats():
    """Splats at the s_max clamp cover every tile: the key buffers must grow and the iteration be redone."""
    W, H, n = 512, 512, 2000
    rng = np.random.default_rng(1)
    splats = random_splats(rng, W, H, n, 200.0, 1024.0, opacity=(0.1, 0.3))
    target = gs2d.synth_target(W, H, seed=4)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target(target)
        fit.upload(splats)
        cap0 = fit.stats()["capacity"]
        mse = fit.step(2)[:, 0]
        st = fit.stats()
        assert st["capacity"] > cap0 and st["keys"] > cap0
        s1, _, _, _, it = fit.download()
    assert it == 2
    o = ob.State(W, H, splats=splats)
    m0 = o.step(target)
    assert abs(mse[0] - m0) <= 1e-6 * m0


def test_band_sharding_sums_to_the_whole():
    """Tile-row bands (the multi-GPU partition) rendered one after the other on one GPU add up to the full result."""
    W, H, n = 300, 260, 1500
    target = gs2d.synth_target(W, H, seed=6)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target(target)
        fit.restart()
        fit.forward()
        img_full = fit.image()
        mse_full = fit.last_mse()[0]
        fit.backward()
        g_full = fit.grads()
        img_sum = np.zeros_like(img_full)
        g_sum = np.zeros(n * 9, np.float64)
        mse_sum = 0.0
        for r in range(3):
            fit.set_band(r, 3)
            fit.forward()
            mse_sum += fit.last_mse()[0]
            rows = slice(min(H, r * 6 * 16), min(H, (r + 1) * 6 * 16))  # ceil(17 tile rows / 3) = 6
            img_sum[rows] = fit.image()[rows]
            fit.backward()
            g_sum += fit.grads().view(np.float32).astype(np.float64)
    assert np.array_equal(img_sum, img_full)
    assert mse_sum == pytest.approx(mse_full, rel=1e-12)
    gf = g_full.view(np.float32).astype(np.float64)
    assert np.abs(g_sum - gf).max() <= 1e-4 * np.abs(gf).max()


def test_independent_images_batch():
    """BASELINE config 5 in miniature: several images fitted concurrently == each fitted alone."""
    W, H, n, B = 96, 80, 400, 3
    targets = [gs2d.synth_target(W, H, seed=1 + i) for i in range(B)]
    with gs2d.Fitter(W, H, n, n_images=B) as fit:
        for i, t in enumerate(targets):
            fit.set_target(t, image=i)
        fit.restart()
        mse_b = fit.step(5)
        imgs = [fit.image(i) for i in range(B)]
        states = [fit.download(i)[0] for i in range(B)]
    for i, t in enumerate(targets):
        with gs2d.Fitter(W, H, n) as fit:
            fit.set_target(t)
            fit.restart()
            mse_1 = fit.step(5)[:, 0]
            img1 = fit.image()
            s1 = fit.download()[0]
        np.testing.assert_allclose(mse_b[:, i], mse_1, rtol=1e-5)
        assert np.abs(imgs[i] - img1).max() <= 1e-3
        assert np.quantile(np.abs(states[i].view(np.float32) - s1.view(np.float32)), 0.99) <= 1e-3


def tiled_target_u8(W, H, tile=1024, seed=1):
    """RGBA8 target of a large image from one synthetic tile repeated (generating 67 M pixels of pcg3d noise on the
    host would take longer than the test); only the window the oracle looks at has to be known exactly."""
    t = np.rint(gs2d.synth_target(tile, tile, seed=seed) * 255.0).astype(np.uint8)
    return np.ascontiguousarray(np.tile(t, ((H + tile - 1) // tile, (W + tile - 1) // tile, 1))[:H, :W])


def crop_check(fit, splats, rgba8, window, mode, label):
    """Forward AND backward of a full-size problem against the oracle on a window. Pixels are independent given the
    splats, so the oracle renders the window alone (splats shifted into window coordinates; the window origin is a
    multiple of 16 and the coordinates are below 2^13, so the shift is exact in float), and the gradients of the
    Gaussians whose whole footprint lies inside the window's interior are complete sums over pixels the oracle has:
    they must agree like on a small image. What differs: pos.x + u of the footprint solve (main.cpp:508-509) is
    rounded at a different magnitude after the shift, so a span end that sits within ~1e-3 px of an integer may move by
    one pixel - a pixel with alpha ~1e-3 at the 3-sigma rim; hence quantiles, not maxima, carry the tight bounds."""
    x0, y0, w, h = window
    W, H = fit.W, fit.H
    fit.forward()
    img = fit.image()
    st = fit.stats()
    fit.backward()
    g = fit.grads()
    assert fit.stats()["live_pairs_backward"] == st["live_pairs"]
    s = splats.copy()
    s["pos"][:, 0] -= x0
    s["pos"][:, 1] -= y0
    crop_cpu, _, _ = ob.forward(w, h, s)
    crop_gpu = img[y0 : y0 + h, x0 : x0 + w]
    # interior only: the crop border clips footprints differently (x < 0 rule, main.cpp:513)
    d = np.abs(crop_gpu[8:-8, 8:-8, :3] - crop_cpu[8:-8, 8:-8, :3])
    q = np.quantile(d, 0.999)
    print(f"[{mode}] {label}: crop image q99.9 {q:.3e} max {d.max():.3e}; keys {st['keys']}, live/px {st['live_pairs'] / (W * H):.2f}")
    assert q <= 1e-5 and d.max() <= IMG_ATOL, (q, d.max())
    # gradients: Gaussians whose footprint (+ margin) lies inside the interior of the window
    target_crop = ob.u8_to_target(np.ascontiguousarray(rgba8[y0 : y0 + h, x0 : x0 + w]))
    g_cpu = ob.backward(w, h, s, crop_cpu, target_crop)
    r = 3.0 * np.maximum(s["sx"], s["sy"]) + 3.0
    inside = ((s["pos"][:, 0] - r >= 10) & (s["pos"][:, 0] + r <= w - 10) & (s["pos"][:, 1] - r >= 10) &
              (s["pos"][:, 1] + r <= h - 10))
    n_in = int(inside.sum())
    assert n_in >= 100, n_in
    rtol = grad_rtol(mode)
    worst = 0.0
    for f in ("pos", "sx", "sy", "rot", "color"):
        a, b = g[f][inside].astype(np.float64), g_cpu[f][inside].astype(np.float64)
        err = np.abs(a - b) / (np.abs(b) + np.abs(b).mean() + 1e-30)
        worst = max(worst, err.max())
        assert np.quantile(err, 0.99) <= rtol, (f, np.quantile(err, 0.99))
        assert err.max() <= 20 * rtol, (f, err.max())
    print(f"[{mode}] {label}: gradients of {n_in} Gaussians inside the window, worst rel err {worst:.3e}")
    return img, g, st


def test_target_formats_can_be_mixed():
    """RGBA8 targets are read by the blend kernels as they are (4 bytes per pixel, converted through a 256-entry table
    of c / 255.0f, prp.cpp:796-807) as long as no image of the context has a float32 target; the first float32 upload
    converts the 8-bit images once and the context carries on in float32. All routes must give the same fit."""
    W, H, n = 150, 110, 300
    rng = np.random.default_rng(4)
    t8 = [rng.integers(0, 256, (H, W, 4), dtype=np.uint8) for _ in range(2)]
    tf = [ob.u8_to_target(t) for t in t8]
    curves = []
    for route in ("u8 u8", "f32 f32", "u8 then f32", "f32 then u8", "u8, step, then f32 for the other image"):
        with gs2d.Fitter(W, H, n, n_images=2) as fit:
            if route == "u8 u8":
                fit.set_target_u8(t8[0], 0); fit.set_target_u8(t8[1], 1)
            elif route == "f32 f32":
                fit.set_target(tf[0], 0); fit.set_target(tf[1], 1)
            elif route == "u8 then f32":
                fit.set_target_u8(t8[0], 0); fit.set_target(tf[1], 1)
            elif route == "f32 then u8":
                fit.set_target(tf[0], 0); fit.set_target_u8(t8[1], 1)
            else:
                fit.set_target_u8(t8[0], 0); fit.set_target_u8(t8[0], 1)
                fit.restart(); fit.step(2)
                fit.set_target(tf[1], 1)
            fit.restart()
            curves.append(fit.step(6))
    for c in curves[1:]:
        assert c[0, 0] == curves[0][0, 0] and c[0, 1] == curves[0][0, 1]
        assert np.allclose(c, curves[0], rtol=1e-7, atol=0)


def test_full_size_crop_against_oracle(mode):
    """BASELINE config 3 size (4096^2, 1 M Gaussians): forward and backward against the oracle on a 256x192 window."""
    W = H = 4096
    n = 1_000_000
    rgba8 = tiled_target_u8(W, H)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target_u8(rgba8)
        fit.restart()
        splats = fit.download(with_adam=False)[0]
        img, g, st = crop_check(fit, splats, rgba8, (1936, 2064, 256, 192), mode, "C3")
        fit.forward()
        assert np.array_equal(img, fit.image())  # idempotent / deterministic
    assert 10e6 < st["keys"] < 40e6 and 15 * W * H < st["live_pairs"] < 22 * W * H  # ~18.5 live / px


def test_c4_size_crop_and_huge_splats_against_oracle(mode):
    """BASELINE config 4 size on one GPU (8192^2 = 262 144 tiles, 4 M Gaussians, 128 index slices): forward and
    backward against the oracle on a window, first with the reference's init(), then with a dozen Gaussians at the
    1024 scale clamp (main.cpp:744-745) on top - each covers 147 000 tiles, so they go through the overflow path of
    the slot binning, the in-kernel footprint solve (too tall for the row table) and, with 6 M extra list entries in
    a few index slices, the 16-bit bucket counters and 32-bit list offsets at their largest."""
    W = H = 8192
    n = 4_000_000
    rgba8 = tiled_target_u8(W, H)
    window = (4096 + 304, 4096 - 400, 256, 192)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target_u8(rgba8)
        fit.restart()
        splats = fit.download(with_adam=False)[0]
        _, _, st0 = crop_check(fit, splats, rgba8, window, mode, "C4")
        assert 40e6 < st0["keys"] < 160e6
        rng = np.random.default_rng(4)
        huge = np.concatenate([[0, 1, 2, 3], rng.choice(n, 8, replace=False)])
        splats["pos"][huge] = rng.uniform(3500, 4700, (len(huge), 2)).astype(np.float32)
        splats["sx"][huge] = 1024.0
        splats["sy"][huge] = rng.uniform(700, 1024, len(huge)).astype(np.float32)
        splats["rot"][huge] = safe_rotations(rng, len(huge))
        splats["opacity"][huge] = 0.1
        splats["color"][huge] = rng.uniform(0, 1, (len(huge), 3)).astype(np.float32)
        fit.upload(splats)
        _, _, st1 = crop_check(fit, splats, rgba8, window, mode, "C4 + 12 Gaussians at the scale clamp")
        assert st1["keys"] > st0["keys"] + 1_000_000


def test_eight_bands_of_the_headline_config_on_one_gpu():
    """The 8-rank partition of BASELINE config 3 (tile-row bands, gs2d_set_band) rendered band after band on ONE GPU:
    images, loss and gradients must add up to the un-sharded result. Keeps the band logic of k_preprocess / the row
    table / the scan window under test on a 1-GPU box."""
    W = H = 4096
    n = 1_000_000
    rgba8 = tiled_target_u8(W, H)
    with gs2d.Fitter(W, H, n) as fit:
        fit.set_target_u8(rgba8)
        fit.restart()
        fit.forward()
        img_full = fit.image()
        mse_full = fit.last_mse()[0]
        live_full = fit.stats()["live_pairs"]
        fit.backward()
        g_full = fit.grads().view(np.float32).astype(np.float64)
        img_sum = np.zeros_like(img_full)
        g_sum = np.zeros_like(g_full)
        mse_sum, live_sum = 0.0, 0
        for r in range(8):
            fit.set_band(r, 8)
            fit.forward()
            mse_sum += fit.last_mse()[0]
            live_sum += fit.stats()["live_pairs"]
            rows = slice(r * 512, (r + 1) * 512)  # 256 tile rows / 8
            img_sum[rows] = fit.image()[rows]
            fit.backward()
            g_sum += fit.grads().view(np.float32).astype(np.float64)
    assert np.array_equal(img_sum, img_full)
    assert live_sum == live_full
    assert mse_sum == pytest.approx(mse_full, rel=1e-12)
    scale = np.abs(g_full).mean()
    assert np.abs(g_sum - g_full).max() <= 2e-4 * (np.abs(g_full).max() + scale)
