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
