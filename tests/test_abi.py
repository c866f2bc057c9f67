"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/gs2d.h declares,
its struct layouts equal the reference's AoS types, and the product path fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from gs2d_loader import ROOT, gs2d


def header_symbols():
    text = open(os.path.join(ROOT, "include", "gs2d.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gs2d_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = gs2d.load_library()
    declared = header_symbols()
    assert len(declared) >= 25
    for sym in declared:
        assert hasattr(lib, sym), f"libgs2d.so does not export {sym}"
    assert sorted(gs2d.ABI_SYMBOLS) == declared


def test_struct_layouts_match_reference():
    # struct Splat = 9 floats (main.cpp:85-93), struct SplatAdam = 9 x {m,v} (main.cpp:158-166)
    assert gs2d.SPLAT_DTYPE.itemsize == 36 and gs2d.ADAM_DTYPE.itemsize == 72
    assert gs2d.SPLAT_DTYPE.fields["rot"][1] == 16 and gs2d.SPLAT_DTYPE.fields["opacity"][1] == 32
    assert gs2d.ADAM_DTYPE.fields["color"][1] == 40
    assert C.sizeof(gs2d.Params) == 44


def test_default_params_are_the_reference_literals():
    p = gs2d.default_params()
    assert p.lr == np.float32(0.05) and p.beta1 == np.float32(0.9) and p.beta2 == np.float32(0.99)
    assert p.eps == np.float32(1e-15) and p.bounds == 3.0 and p.min_throughput == 1.0 / 256.0
    assert (p.s_min, p.s_max, p.o_min, p.o_max) == (1.0, 1024.0, np.float32(0.1), 1.0)
    assert p.optimize_opacity == 0
    assert gs2d.load_library().gs2d_version() == 100


def test_no_cpu_fallback():
    """Without a B200 the product refuses to run (GS2D_ERR_NO_DEVICE) instead of computing on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(gs2d.Gs2dError) as e:
        gs2d.Fitter(64, 64, 16)
    assert e.value.code == -2


def test_invalid_arguments():
    lib = gs2d.load_library()
    ctx = C.c_void_p()
    assert lib.gs2d_create(C.byref(ctx), 0, 10, 1, 1, 0, None) == -1
    assert lib.gs2d_create(C.byref(ctx), 10, 10, -1, 1, 0, None) == -1
    assert lib.gs2d_create(None, 10, 10, 1, 1, 0, None) == -1
    assert lib.gs2d_step(None, 1, None) == -1
    assert lib.gs2d_strerror(-4).decode().startswith("non-finite")


def test_synth_target_is_deterministic():
    a = gs2d.synth_target(96, 64, seed=3)
    b = gs2d.synth_target(96, 64, seed=3)
    c = gs2d.synth_target(96, 64, seed=4)
    assert a.dtype == np.float32 and a.shape == (64, 96, 4)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert (a[..., 3] == 1).all() and a[..., :3].min() >= 0 and a[..., :3].max() <= 1
    # pcg3d of the host layer == the oracle's (main.cpp:17-24)
    import oracle_binding as ob

    v = np.array([5, 1, 0xFFFFFFFF], np.uint32)
    ob.lib().oracle_pcg3d(v.ctypes.data_as(C.c_void_p))
    x, y, z = gs2d.pcg3d([5], [1], [0xFFFFFFFF])
    assert (int(x[0]), int(y[0]), int(z[0])) == tuple(int(t) for t in v)
