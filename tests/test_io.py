"""SURVEY.md §8(f) rows either side of the hot path, host code only (no GPU): target ingest (f2), image dump (f1),
overlay geometry (f4). The checkpoint round trip (f3) needs a context and lives in test_gpu_parity.py."""
import os

import numpy as np
import pytest

from gs2d_loader import ROOT, gs2d

GOLD = os.path.join(ROOT, "tests", "golden")
REF_BIN = "/root/reference/bin"


def _cases():
    z = np.load(os.path.join(GOLD, "jpeg", "expected.npz"))
    return sorted(z.files)


@pytest.mark.parametrize("name", _cases())
def test_jpeg_decoder_equals_stb(name):
    """Every coding mode (baseline/progressive, 4:4:4 / 4:2:2 / 4:2:0 / gray, restart intervals, 1-pixel-wide chroma):
    bit-identical to the reference's vendored stb_image v2.23 (goldens made by tests/golden/make_jpeg_golden.py)."""
    want = np.load(os.path.join(GOLD, "jpeg", "expected.npz"))[name]
    got = gs2d.load_image(os.path.join(GOLD, "jpeg", name + ".jpg"))
    assert got.shape == want.shape
    assert np.array_equal(got, want), f"{(got != want).sum()} differing bytes"
    data = open(os.path.join(GOLD, "jpeg", name + ".jpg"), "rb").read()
    assert np.array_equal(gs2d.decode_jpeg(data), want)


@pytest.mark.parametrize("cfg,jpeg", [("c1", "squirrel_cls_mini.jpg"), ("c2", "squirrel_cls.jpg")])
def test_reference_targets_decode_bit_identically(cfg, jpeg):
    """BASELINE configs[0] / [1]: the two JPEGs the reference fits, against the target the unmodified main.cpp saw."""
    path = os.path.join(REF_BIN, jpeg)
    if not os.path.exists(path):
        pytest.skip("/root/reference is not present on this box")
    want = np.load(os.path.join(GOLD, f"{cfg}_target_rgb8.npz"))["rgb"]
    got = gs2d.load_image(path)
    assert np.array_equal(got[..., :3], want) and (got[..., 3] == 255).all()


def test_truncated_and_garbage_files_are_errors(tmp_path):
    data = open(os.path.join(GOLD, "jpeg", "base_420.jpg"), "rb").read()
    with pytest.raises(gs2d.Gs2dError) as e:
        gs2d.decode_jpeg(b"\x00" * 64)
    assert e.value.code == -9
    with pytest.raises(gs2d.Gs2dError):
        gs2d.decode_jpeg(data[:40])  # ends inside the tables
    with pytest.raises(gs2d.Gs2dError) as e:
        gs2d.load_image(str(tmp_path / "missing.jpg"))
    assert e.value.code == -8
    # a scan cut short still yields a picture of the right size (stb's behaviour: missing data decodes as zeros)
    cut = gs2d.decode_jpeg(data[: len(data) - 200])
    assert cut.shape == (29, 37, 4)


def test_ppm_round_trip(tmp_path):
    rng = np.random.default_rng(3)
    img = rng.random((19, 23, 4), dtype=np.float32)
    img[0, 0] = (-0.5, 1.5, 0.25, 1.0)
    p = str(tmp_path / "a.ppm")
    gs2d.write_ppm(p, img)
    back = gs2d.load_image(p)
    want = np.rint(np.clip(img[..., :3], 0, 1) * np.float32(255.0)).astype(np.uint8)
    assert np.array_equal(back[..., :3], want) and (back[..., 3] == 255).all()


def test_overlay_geometry_matches_reference_formulas():
    """main.cpp:187-234, :421-465 restated in numpy float32: eigen axes, ellipse recurrence, box half-extents."""
    f = np.float32
    s = np.zeros(3, gs2d.SPLAT_DTYPE)
    s["pos"] = [(10.5, 20.25), (3.0, 4.0), (100.0, 50.0)]
    s["sx"] = [6.0, 10.0, 2.5]
    s["sy"] = [9.0, 3.0, 2.5]
    s["rot"] = [0.3, 2.0, 1.0]
    rec, ell = gs2d.overlay_geometry(s, nvtx=16)
    for i in range(3):
        c, sn = f(np.cos(f(s["rot"][i]))), f(np.sin(f(s["rot"][i])))
        l0, l1 = f(s["sx"][i] * s["sx"][i]), f(s["sy"][i] * s["sy"][i])
        s11 = f(f(f(l0 * c) * c) + f(f(l1 * sn) * sn))
        s12 = f(f(f(l0 - l1) * sn) * c)
        s22 = f(f(l0 + l1) - s11)
        cov = np.array([[s11, s12], [s12, s22]], np.float64)
        w, v = np.linalg.eigh(cov)
        a0, a1 = rec["axis0"][i].astype(np.float64), rec["axis1"][i].astype(np.float64)
        # axis0 is the eigenvector of the larger eigenvalue scaled by its square root; axis1 is perpendicular
        assert np.isclose(np.linalg.norm(a0), np.sqrt(w[1]), rtol=1e-4)
        assert np.isclose(np.linalg.norm(a1), np.sqrt(w[0]), rtol=1e-4, atol=1e-4)
        assert abs(a0 @ a1) < 1e-3 * np.linalg.norm(a0) * max(np.linalg.norm(a1), 1e-6)
        if w[1] - w[0] > 1e-3:
            assert abs(abs(a0 @ v[:, 1]) - np.linalg.norm(a0)) < 1e-3 * np.linalg.norm(a0)
        # half-extent of the 1-sigma box = sqrt of the covariance diagonal
        assert np.allclose(rec["half_extent"][i], [np.sqrt(s11), np.sqrt(s22)], rtol=1e-5)
        # the polyline starts and ends on center + axis1 (sin 0 = 0, cos 0 = 1) and stays on the 1-sigma ellipse
        assert np.allclose(ell[i, 0], s["pos"][i] + rec["axis1"][i], atol=1e-5)
        assert np.allclose(ell[i, 16], ell[i, 0], atol=1e-3)
        d = ell[i].astype(np.float64) - s["pos"][i]
        m = np.einsum("ki,ij,kj->k", d, np.linalg.inv(cov), d)
        assert np.allclose(m, 1.0, atol=2e-3)


def test_overlay_geometry_matches_the_reference_draw_calls(golden_dir):
    """gs2d_overlay_geometry against the PrimVertex calls of the UNMODIFIED reference (main.cpp:442-476), recorded by
    the level-0 shim in frames 0 and 6 of the default configuration (tests/golden/make_overlay_golden.py). The
    Gaussians of frame 6 come from six iterations of the level-1 oracle (bit-identical to level 0, test_oracle.py)."""
    import oracle_binding as ob

    g = np.load(os.path.join(golden_dir, "c1_overlay.npz"))
    rgb = np.load(os.path.join(golden_dir, "c1_target_rgb8.npz"))["rgb"]
    h, w, _ = rgb.shape
    target = ob.u8_to_target(np.concatenate([rgb, np.full((h, w, 1), 255, np.uint8)], axis=2))
    st = ob.State(w, h, 1024)
    done = 0
    for k, frame in enumerate(g["frames"]):
        while done < frame:
            st.step(target)
            done += 1
        s = st.splats
        rec, ell = gs2d.overlay_geometry(s, nvtx=16)
        xy = g["xy"][k].astype(np.float64)            # [1024, 46, 2], y flipped for GL (main.cpp:447)
        ref = xy * np.array([1.0, -1.0])              # back to image coordinates
        c = s["pos"].astype(np.float64)
        tol = dict(rtol=0, atol=2e-5)                 # coordinates of a few hundred: ~1 ulp
        assert np.allclose(ref[:, 0], c, **tol) and np.allclose(ref[:, 2], c, **tol)
        assert np.allclose(ref[:, 1], c + rec["axis0"], **tol)          # eigen axis of the larger eigenvalue
        assert np.allclose(ref[:, 3], c + rec["axis1"], **tol)
        # ellipse: vertices 4 + 2 i are the segment starts E_i, i = 0..16 (CircleGenerator recurrence, prp.hpp:276-313)
        assert np.allclose(ref[:, 4:38:2], ell, **tol)
        # segment ends: E_{i+1}
        assert np.allclose(ref[:, 5:37:2], ell[:, 1:], **tol)
        # 1-sigma box: (-hx,-hy) (hx,-hy) (hx,hy) (-hx,hy) in GL coordinates around (x, -y)
        hx, hy = rec["half_extent"][:, 0].astype(np.float64), rec["half_extent"][:, 1].astype(np.float64)
        gl = xy[:, 38:46:2] - xy[:, 0][:, None, :]
        want = np.stack([np.stack([-hx, -hy], 1), np.stack([hx, -hy], 1), np.stack([hx, hy], 1), np.stack([-hx, hy], 1)], 1)
        assert np.allclose(gl, want, **tol)
        # colours: white / 230 axes, splat colour * 255 truncated, 128 box (main.cpp:446-470)
        col = g["rgb"][k]
        assert (col[:, 0:3] == 255).all() and (col[:, 3] == 230).all() and (col[:, 38:] == 128).all()
        assert np.array_equal(col[:, 4], (s["color"] * np.float32(255.0)).astype(np.uint32).astype(np.uint8))
