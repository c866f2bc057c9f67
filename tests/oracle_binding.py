"""ctypes binding of the CPU oracle (oracle/libgs2d_oracle.so) — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module; the
product package never does (it must fail loudly without its CUDA library instead).
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "libgs2d_oracle.so")
REF_MAIN = os.path.join(ORACLE_DIR, "_ref", "ref_main")

# numpy views of the reference's AoS types (main.cpp:85-93, :158-166)
SPLAT_DTYPE = np.dtype(
    [("pos", "<f4", (2,)), ("sx", "<f4"), ("sy", "<f4"), ("rot", "<f4"), ("color", "<f4", (3,)), ("opacity", "<f4")]
)
ADAM_DTYPE = np.dtype(
    [
        ("pos", "<f4", (2, 2)),
        ("sx", "<f4", (2,)),
        ("sy", "<f4", (2,)),
        ("rot", "<f4", (2,)),
        ("color", "<f4", (3, 2)),
        ("opacity", "<f4", (2,)),
    ]
)
assert SPLAT_DTYPE.itemsize == 36 and ADAM_DTYPE.itemsize == 72


class Params(C.Structure):
    _fields_ = [
        ("lr", C.c_float),
        ("beta1", C.c_float),
        ("beta2", C.c_float),
        ("eps", C.c_float),
        ("bounds", C.c_float),
        ("min_throughput", C.c_float),
        ("s_min", C.c_float),
        ("s_max", C.c_float),
        ("o_min", C.c_float),
        ("o_max", C.c_float),
        ("optimize_opacity", C.c_int),
    ]


def default_params(optimize_opacity=False):
    """The reference's literals (main.cpp:7-8, :136-137, :154, :317, :715, :744-749)."""
    return Params(0.05, 0.9, 0.99, 1.0e-15, 3.0, 1.0 / 256.0, 1.0, 1024.0, 0.1, 1.0, int(optimize_opacity))


def build():
    subprocess.run(["make", "-C", ORACLE_DIR, "port", "ref"], check=True, stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.oracle_pcg3d.argtypes = [vp]
        L.oracle_init_splats.argtypes = [C.c_int, C.c_int, C.c_int, vp]
        L.oracle_forward.argtypes = [C.c_int, C.c_int, C.c_int, vp, C.POINTER(Params), vp, vp]
        L.oracle_backward.argtypes = [C.c_int, C.c_int, C.c_int, vp, C.POINTER(Params), vp, vp, vp, vp]
        L.oracle_adam.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.POINTER(Params)]
        L.oracle_adam.restype = C.c_int
        L.oracle_mse.argtypes = [C.c_int, C.c_int, vp, vp]
        L.oracle_mse.restype = C.c_double
        L.oracle_step.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.POINTER(Params), vp, vp, vp, vp, vp]
        L.oracle_step.restype = C.c_double
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def init_splats(W, H, n):
    s = np.zeros(n, SPLAT_DTYPE)
    lib().oracle_init_splats(W, H, n, _p(s))
    return s


def forward(W, H, splats, params=None):
    params = params or default_params()
    img = np.empty((H, W, 4), np.float32)
    counts = np.zeros(2, np.uint64)
    lib().oracle_forward(W, H, len(splats), _p(splats), C.byref(params), _p(img), _p(counts))
    return img, int(counts[0]), int(counts[1])


def backward(W, H, splats, image, target, params=None):
    params = params or default_params()
    scratch = np.empty((H, W, 4), np.float32)
    ds = np.zeros(len(splats), SPLAT_DTYPE)
    lib().oracle_backward(W, H, len(splats), _p(splats), C.byref(params), _p(image), _p(target), _p(scratch), _p(ds))
    return ds


def mse(W, H, image, target):
    return lib().oracle_mse(W, H, _p(image), _p(target))


class State:
    """Loop-carried state of the reference (main.cpp:271-278)."""

    def __init__(self, W, H, n=None, splats=None):
        self.W, self.H = W, H
        self.splats = init_splats(W, H, n) if splats is None else splats.copy()
        self.adam = np.zeros(len(self.splats), ADAM_DTYPE)
        self.beta = np.ones(2, np.float32)
        self.iteration = 0
        self.image = np.empty((H, W, 4), np.float32)
        self.scratch = np.empty((H, W, 4), np.float32)
        self.dsplats = np.zeros(len(self.splats), SPLAT_DTYPE)
        self.nonfinite = 0
        self.counts = np.zeros(2, np.uint64)

    def step(self, target, params=None):
        """One loop body (main.cpp:414-809); returns the printed mse."""
        params = params or default_params()
        bad = C.c_int(0)
        b = self.beta
        m = lib().oracle_step(
            self.W, self.H, len(self.splats), _p(self.splats), _p(self.adam), _p(b[0:1]), _p(b[1:2]), _p(target),
            C.byref(params), _p(self.image), _p(self.scratch), _p(self.dsplats), C.byref(bad), _p(self.counts),
        )
        self.nonfinite = bad.value
        self.iteration += 1
        return m

    def adam_only(self, dsplats, params=None):
        params = params or default_params()
        b = self.beta
        return lib().oracle_adam(
            self.W, self.H, len(self.splats), _p(self.splats), _p(self.adam), _p(dsplats), _p(b[0:1]), _p(b[1:2]),
            C.byref(params),
        )


def u8_to_target(rgba8):
    """Image2DRGBA8_to_Image2DRGBA32 (prp.cpp:796-807): vec4(u8) / vec4(255.0f), float division."""
    return (rgba8.astype(np.float32) / np.float32(255.0)).astype(np.float32)


def run_level0(data_dir, iters, image=None, dump_dir=None, target_path=None):
    """Run the UNMODIFIED reference main.cpp (oracle/_ref/ref_main); returns its stdout lines."""
    env = dict(os.environ, PR_DATA=data_dir, PR_ITERS=str(iters))
    if image:
        env["PR_IMAGE"] = image
    if dump_dir:
        env["PR_DUMP"] = dump_dir
    if target_path:
        env["PR_TARGET"] = target_path
    out = subprocess.run([REF_MAIN], env=env, check=True, capture_output=True, text=True).stdout
    return out.splitlines()
