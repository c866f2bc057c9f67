"""CPU tests of the product's state arithmetic (no GPU): 2dgaussiansplatting_b200/csrc/gs2d_update.cuh — the header
k_adam, k_adam_peer and k_init_reference include — compiled for the host by tests/update_host.cpp and compared BIT FOR
BIT with the CPU oracle (which is pinned to the unmodified main.cpp by test_oracle.py):

  * init_splat / pcg3d  vs  oracle_init_splats (main.cpp:17-24, :280-305), several image sizes and counts;
  * adam_one + clamp_glm in the loop of k_advance / k_adam  vs  oracle_adam (main.cpp:144-156, :714-785): many steps
    from random states and gradients, zero gradients (m = v = 0 -> 0/(0 + 1e-15) = 0), values at the clamps, opacity
    optimised or not, NaN / inf gradients (the finite check covers colour, sx, sy, rot, pos.x only)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_binding as ob

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def up(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("up") / "libupdate_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC",
                    os.path.join(HERE, "update_host.cpp"), "-o", so, "-lm"], check=True)
    L = C.CDLL(so)
    vp = C.c_void_p
    L.up_init.argtypes = [C.c_int, C.c_int, C.c_int, vp]
    L.up_adam.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.POINTER(ob.Params)]
    L.up_adam.restype = C.c_int
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.mark.parametrize("W,H,n", [(268, 213, 1024), (4096, 4096, 50000), (8192, 8192, 3000), (17, 5, 7), (1, 1, 3)])
def test_init_is_bit_identical_to_the_reference_init(up, W, H, n):
    ref = ob.init_splats(W, H, n)
    mine = np.zeros(n, ob.SPLAT_DTYPE)
    up.up_init(W, H, n, _p(mine))
    assert mine.tobytes() == ref.tobytes()


def random_state(rng, W, H, n):
    s = np.zeros(n, ob.SPLAT_DTYPE)
    s["pos"][:, 0] = rng.uniform(0, W - 1, n)
    s["pos"][:, 1] = rng.uniform(0, H - 1, n)
    s["sx"], s["sy"] = rng.uniform(1, 40, n), rng.uniform(1, 40, n)
    s["rot"] = rng.uniform(-7, 7, n)
    s["color"] = rng.uniform(0, 1, (n, 3))
    s["opacity"] = rng.uniform(0.1, 1, n)
    # some values exactly at / next to the clamps
    s["sx"][: n // 10] = 1.0
    s["sy"][n // 10: n // 5] = 1024.0
    s["color"][: n // 10] = 0.0
    s["color"][n // 10: n // 5] = 1.0
    s["pos"][: n // 10, 0] = 0.0
    s["pos"][n // 10: n // 5, 1] = H - 1
    return s


def random_grads(rng, n, scale):
    g = np.zeros(n, ob.SPLAT_DTYPE)
    v = g.view(np.float32).reshape(n, 9)
    v[:] = rng.standard_normal((n, 9)) * np.exp(rng.uniform(np.log(1e-12), np.log(scale), (n, 9)))
    v[rng.uniform(size=(n, 9)) < 0.2] = 0.0  # occluded splats: zero gradient
    return g


@pytest.mark.parametrize("optimize_opacity", [False, True])
def test_adam_and_clamps_are_bit_identical_to_the_reference(up, optimize_opacity):
    W, H, n = 300, 200, 4000
    rng = np.random.default_rng(3 + optimize_opacity)
    params = ob.default_params(optimize_opacity)
    st = ob.State(W, H, splats=random_state(rng, W, H, n))
    mine_s, mine_a, mine_b = st.splats.copy(), st.adam.copy(), st.beta.copy()
    for it in range(40):
        g = random_grads(rng, n, 1e4 if it % 3 else 1e-3)
        if it == 20:
            g.view(np.float32)[:] = 0.0  # a whole step without gradient
        bad_ref = st.adam_only(g, params)
        bad = up.up_adam(W, H, n, _p(mine_s), _p(mine_a), _p(g), _p(mine_b[0:1]), _p(mine_b[1:2]), C.byref(params))
        assert bad == bad_ref == 0
        assert mine_b.tobytes() == st.beta.tobytes(), it
        assert mine_a.tobytes() == st.adam.tobytes(), it
        assert mine_s.tobytes() == st.splats.tobytes(), it
    if not optimize_opacity:
        assert (st.adam["opacity"] == 0).all()  # its moments stay 0 (main.cpp:733)


def test_nonfinite_gradients_are_flagged_like_the_reference_aborts(up):
    W, H, n = 64, 48, 16
    params = ob.default_params()
    fields = [("pos", 0), ("pos", 1), ("sx", None), ("sy", None), ("rot", None), ("color", 1), ("opacity", None)]
    for bad_value in (np.nan, np.inf, -np.inf):
        for name, sub in fields:
            rng = np.random.default_rng(11)
            st = ob.State(W, H, splats=random_state(rng, W, H, n))
            mine_s, mine_a, mine_b = st.splats.copy(), st.adam.copy(), st.beta.copy()
            g = random_grads(rng, n, 1.0)
            if sub is None:
                g[name][5] = bad_value
            else:
                g[name][5, sub] = bad_value
            bad_ref = st.adam_only(g, params)
            bad = up.up_adam(W, H, n, _p(mine_s), _p(mine_a), _p(g), _p(mine_b[0:1]), _p(mine_b[1:2]), C.byref(params))
            assert bad == bad_ref, (name, sub, bad_value)
            assert mine_s.tobytes() == st.splats.tobytes() and mine_a.tobytes() == st.adam.tobytes()
            # pos.y and (unoptimised) opacity are not part of the reference's finite check (main.cpp:752-785)
            if (name, sub) in (("pos", 1), ("opacity", None)):
                assert bad == 0


@pytest.mark.parametrize("n_ranks,H,lr", [(8, 4096, 0.05), (2, 520, 0.05), (4, 1024, 2.0)])
def test_sparse_parameter_broadcast_keeps_every_rank_exact_where_it_looks(up, n_ranks, H, lr):
    """The sparse exchange of the multi-GPU path, simulated on the CPU with the product's own two band
    tests (band_rejects = what k_preprocess of a rank applies, reach_mask = what the owner applies; gs2d_footprint.cuh
    compiled for the host): the owner sends a Gaussian's new parameters to the ranks in reach_now | reach_prev only.
    Invariant, at EVERY iteration and for EVERY rank: the rank's own copy equals the true parameters for every
    Gaussian its band test does not reject. Gaussians random-walk across the band boundaries with Adam-sized steps
    (|d py|, |d sx|, |d sy| <= lr per iteration), some in big jumps at the start of 'uploads' (all copies refreshed)."""
    up.up_band_tests.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p,
                                 C.c_void_p]
    rng = np.random.default_rng(n_ranks)
    n = 6000
    tiles_y = (H + 15) // 16
    band_px = ((tiles_y + n_ranks - 1) // n_ranks) * 16
    slack = np.float32(1.0 + 3.0 * lr * 1.001 + lr)  # gs2d_api.cu, setup_peer_update

    def tests(p):
        rej = np.zeros((n_ranks, n), np.uint8)
        masks = np.zeros(n, np.uint32)
        up.up_band_tests(n, _p(p[0]), _p(p[1]), _p(p[2]), band_px, n_ranks, slack, _p(rej), _p(masks))
        return rej.astype(bool), masks

    true = np.stack([rng.uniform(-20, H + 20, n), rng.uniform(1, 40, n), rng.uniform(1, 40, n)]).astype(np.float32)
    true[1, ::50] = 1024.0  # a few that reach every band
    true[0, 7] = np.nan     # NaN reaches everyone and is never rejected
    copies = [true.copy() for _ in range(n_ranks)]
    prev = np.full(n, (1 << n_ranks) - 1, np.uint32)  # after an upload: everything is sent once
    stale_seen = 0
    for it in range(300):
        _, now = tests(true)                           # k_preprocess of the owner, parameters BEFORE the step
        step = rng.uniform(-lr, lr, true.shape).astype(np.float32)
        step[:, rng.random(n) < 0.3] = 0               # occluded Gaussians do not move
        new = true + step
        new[1:] = np.clip(new[1:], 1.0, 1024.0)        # main.cpp:744-745
        send = now | prev
        for r in range(n_ranks):
            sel = (send >> r) & 1 == 1
            copies[r][:, sel] = new[:, sel]
        prev, true = now, new
        rej, _ = tests(true)                           # k_preprocess of every rank at the NEXT iteration
        for r in range(n_ranks):
            look = ~rej[r]
            same = (copies[r].view(np.uint32) == true.view(np.uint32)).all(axis=0)
            assert same[look].all(), (it, r, np.flatnonzero(look & ~same)[:5])
            # ... and a rank that does not look must also reject its STALE copy (else it would blend stale data)
            rej_stale = np.zeros((n_ranks, n), np.uint8)
            m2 = np.zeros(n, np.uint32)
            c = copies[r]
            up.up_band_tests(n, _p(np.ascontiguousarray(c[0])), _p(np.ascontiguousarray(c[1])), _p(np.ascontiguousarray(c[2])),
                             band_px, n_ranks, slack, _p(rej_stale), _p(m2))
            assert (rej_stale[r].astype(bool) | same).all(), (it, r)
            stale_seen += int((~same).sum())
    assert stale_seen > 0  # the test did exercise stale copies
