"""CPU tests of the product's footprint arithmetic (no GPU): 2dgaussiansplatting_b200/csrc/gs2d_footprint.cuh — the very
header the kernels include — is compiled for the host by tests/footprint_host.cpp (rounding intrinsics mapped to plain
IEEE float operations, -ffp-contract=off) and checked against the CPU oracle:

  * splat_geometry + row_mask, assembled tile row by tile row as k_forward does, must cover EXACTLY the pixels the
    reference's scan-line loop visits (main.cpp:489-514) - the footprint is a step function of rounded values, so
    this is a bit-for-bit statement about the operand order, the citardauq root form and the x86 truncation;
  * tile_span (the per-tile-row ellipse spans of the binning) must emit a SUPERSET of the tiles that hold a covered
    pixel, and fewer tiles than the conservative rectangle.

The header's functions are the device code textually; what the host build cannot show is the device's code
generation (FMA contraction is excluded there by the _rn intrinsics) - that is what the -m gpu parity tests are for."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_binding as ob

HERE = os.path.dirname(os.path.abspath(__file__))
T = 16


@pytest.fixture(scope="module")
def fp(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("fp") / "libfootprint_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC",
                    os.path.join(HERE, "footprint_host.cpp"), "-o", so, "-lm"], check=True)
    L = C.CDLL(so)
    f = C.c_float
    L.fp_coverage.argtypes = [C.c_int, C.c_int, f, f, f, f, f, C.c_void_p]
    L.fp_coverage.restype = C.c_longlong
    L.fp_binned_tiles.argtypes = [C.c_int, C.c_int, f, f, f, f, f, C.c_int, C.c_void_p, C.c_void_p]
    L.fp_binned_tiles.restype = C.c_int
    return L


def random_cases(rng, W, H, n):
    """Positions inside, on the edges of and outside the image; round, needle-like and huge Gaussians; any angle."""
    px = rng.uniform(-20, W + 20, n).astype(np.float32)
    py = rng.uniform(-20, H + 20, n).astype(np.float32)
    sx = np.exp(rng.uniform(0, np.log(80.0), n)).astype(np.float32)   # 1 .. 80 (the clamp keeps scales >= 1)
    sy = np.exp(rng.uniform(0, np.log(80.0), n)).astype(np.float32)
    rot = rng.uniform(-7, 7, n).astype(np.float32)
    # the truncation quirk of main.cpp:508 needs chords ending just left of / above the image
    px[: n // 8] = rng.uniform(-4, 3, n // 8)
    py[n // 8: n // 4] = rng.uniform(-4, 3, n // 4 - n // 8)
    return px, py, sx, sy, rot


def oracle_coverage(W, H, px, py, sx, sy, rot):
    """The pixels the reference's loop visits for ONE Gaussian: colour 1, opacity 1 -> every visited pixel whose
    falloff is not exactly 0 gets a non-zero value; counts[0] is the number of visited pixels."""
    s = np.zeros(1, ob.SPLAT_DTYPE)
    s["pos"][0] = (px, py)
    s["sx"], s["sy"], s["rot"], s["color"], s["opacity"] = sx, sy, rot, 1.0, 1.0
    img, visited, _ = ob.forward(W, H, s)
    return img[..., 0] != 0.0, visited


def host_coverage(fp, W, H, px, py, sx, sy, rot):
    out = np.zeros((H + 1) * W, np.uint8)  # one guard row: a bit outside the image would land there
    n = fp.fp_coverage(W, H, px, py, sx, sy, rot, out.ctypes.data_as(C.c_void_p))
    assert not out[H * W:].any()
    return out[: H * W].reshape(H, W).astype(bool), n


@pytest.mark.parametrize("W,H", [(150, 110), (64, 48), (33, 17)])
def test_row_mask_covers_exactly_the_reference_footprint(fp, W, H):
    rng = np.random.default_rng(W * 1000 + H)
    n = 400
    checked = 0
    for px, py, sx, sy, rot in zip(*random_cases(rng, W, H, n)):
        nz, visited = oracle_coverage(W, H, px, py, sx, sy, rot)
        cov, n_cov = host_coverage(fp, W, H, px, py, sx, sy, rot)
        assert n_cov == cov.sum()
        # same number of pixels, and every pixel the oracle blended something into is covered: the sets are equal
        # (a visited pixel can stay 0 only where exp_approx is exactly 0, i.e. > 4 sigma: the column-0 quirk)
        assert n_cov == visited, (px, py, sx, sy, rot, n_cov, visited)
        assert not (nz & ~cov).any(), (px, py, sx, sy, rot)
        checked += visited
    assert checked > 50 * n  # the cases do cover pixels


def test_tile_spans_are_a_superset_of_the_footprint_and_smaller_than_the_rectangle(fp):
    W, H = 700, 520
    tx_n, ty_n = (W + T - 1) // T, (H + T - 1) // T
    rng = np.random.default_rng(7)
    n = 500
    pairs_span = pairs_rect = pairs_exact = 0
    for px, py, sx, sy, rot in zip(*random_cases(rng, W, H, n)):
        cov, _ = host_coverage(fp, W, H, px, py, sx, sy, rot)
        pad = np.zeros((ty_n * T, tx_n * T), bool)
        pad[:H, :W] = cov
        touched = pad.reshape(ty_n, T, tx_n, T).any(axis=(1, 3))
        tiles = np.zeros((ty_n, tx_n), np.uint8)
        counts = np.zeros(2, np.int64)
        fp.fp_binned_tiles(W, H, px, py, sx, sy, rot, 1, tiles.ctypes.data_as(C.c_void_p), counts.ctypes.data_as(C.c_void_p))
        assert not (touched & ~tiles.astype(bool)).any(), (px, py, sx, sy, rot)
        assert counts[0] == tiles.sum() and counts[0] <= counts[1]
        # culling off (NaN slack): exactly the rectangle
        tiles0 = np.zeros((ty_n, tx_n), np.uint8)
        counts0 = np.zeros(2, np.int64)
        fp.fp_binned_tiles(W, H, px, py, sx, sy, rot, 0, tiles0.ctypes.data_as(C.c_void_p), counts0.ctypes.data_as(C.c_void_p))
        assert counts0[0] == counts0[1] == counts[1]
        assert not (tiles.astype(bool) & ~tiles0.astype(bool)).any()
        pairs_span += counts[0]
        pairs_rect += counts[1]
        pairs_exact += touched.sum()
    assert pairs_exact <= pairs_span < 0.9 * pairs_rect  # the spans do cull (-17 % at the benchmark's scales)


def test_degenerate_inputs_give_empty_or_finite_masks(fp):
    """NaN / zero-determinant covariances (main.cpp would hit UB in the (int) cast: treated as empty rows), Gaussians
    far outside, and the 1024 clamp: no crash, no bit outside the image."""
    W, H = 96, 80
    cases = [(10.0, 10.0, 0.0, 0.0, 0.0), (10.0, 10.0, np.nan, 3.0, 0.0), (np.nan, 10.0, 3.0, 3.0, 0.0),
             (1e9, 1e9, 5.0, 5.0, 1.0), (-1e9, 40.0, 5.0, 5.0, 1.0), (48.0, 40.0, 1024.0, 1024.0, 0.3),
             (48.0, 40.0, 1024.0, 1.0, 0.77), (48.0, 40.0, 1e-20, 1e-20, 0.0), (48.0, 40.0, np.inf, 2.0, 0.0)]
    for px, py, sx, sy, rot in cases:
        cov, n_cov = host_coverage(fp, W, H, *map(np.float32, (px, py, sx, sy, rot)))
        assert n_cov == cov.sum()
        if all(np.isfinite([px, py, sx, sy])) and sx >= 1 and sy >= 1:
            nz, visited = oracle_coverage(W, H, px, py, sx, sy, rot)
            assert n_cov == visited and not (nz & ~cov).any()
