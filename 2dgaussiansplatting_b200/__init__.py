"""gs2d — B200-native 2D Gaussian-splatting image fitting (host-side mirror of the reference's fit loop).

The product is `libgs2d.so` (hand-written sm_100a CUDA behind the C ABI of include/gs2d.h). This module is the thin
ctypes host layer a Python caller uses; it mirrors the reference's driver semantics (main.cpp:236-334, :414-851):

    fit = Fitter(W, H, n_splats)          # splats(NSplat), splatAdams, beta1t = beta2t = 1        main.cpp:271-278
    fit.set_target_u8(rgba8)              # Image2DRGBA8::load + Image2DRGBA8_to_Image2DRGBA32       main.cpp:254-258
    fit.restart()                         # init() / the "Restart" button                            main.cpp:280-307, :828
    for mse in fit.step(100): ...         # 100 loop bodies; mse == what main.cpp:807 prints
    fit.optimize_opacity = True           # the ImGui checkbox                                       main.cpp:825

There is NO CPU fallback: importing works anywhere (so that the symbol table can be checked on a CPU-only box), but
every compute call needs a B200 and raises Gs2dError otherwise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GS2D_LIB") or os.path.join(_HERE, "libgs2d.so")  # GS2D_LIB: developer override

# the reference's AoS types (main.cpp:85-93, :158-166)
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

GS2D_OK = 0
GS2D_ERR_INVALID = -1
GS2D_ERR_NONFINITE = -4

# every symbol include/gs2d.h declares (tests check the library exports all of them)
ABI_SYMBOLS = [
    "gs2d_default_params", "gs2d_version", "gs2d_strerror", "gs2d_last_error", "gs2d_create", "gs2d_destroy",
    "gs2d_set_target_rgba32f", "gs2d_set_target_rgba8", "gs2d_init_reference", "gs2d_upload_state",
    "gs2d_download_state", "gs2d_set_optimize_opacity", "gs2d_step", "gs2d_forward", "gs2d_backward", "gs2d_adam",
    "gs2d_download_image_rgba32f", "gs2d_download_grads", "gs2d_last_mse", "gs2d_mse_ticket", "gs2d_mse_wait", "gs2d_get_stats", "gs2d_comm_unique_id",
    "gs2d_comm_init", "gs2d_set_band", "gs2d_last_step_ms", "gs2d_profile_step", "gs2d_launch_count", "gs2d_sync",
    "gs2d_set_profiling", "gs2d_get_profile", "gs2d_set_stats", "gs2d_set_tuning",
    "gs2d_microbench_fp32", "gs2d_microbench_hbm",
    "gs2d_get_optimize_opacity", "gs2d_get_dims", "gs2d_set_target_rgba8_async", "gs2d_comm_mode", "gs2d_comm_traffic", "gs2d_save_checkpoint", "gs2d_load_checkpoint",
    "gs2d_image_load_rgba8", "gs2d_image_decode_jpeg_rgba8", "gs2d_image_free", "gs2d_image_write_ppm",
    "gs2d_overlay_geometry", "gs2d_mse_static",
]


class Params(C.Structure):
    _fields_ = [
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float), ("bounds", C.c_float),
        ("min_throughput", C.c_float), ("s_min", C.c_float), ("s_max", C.c_float), ("o_min", C.c_float),
        ("o_max", C.c_float), ("optimize_opacity", C.c_int),
    ]


class Stats(C.Structure):
    _fields_ = [("keys", C.c_uint64), ("live_pairs", C.c_uint64), ("staged", C.c_uint64), ("capacity", C.c_uint64),
                ("live_pairs_sum", C.c_uint64), ("staged_sum", C.c_uint64), ("keys_sum", C.c_uint64),
                ("n_forwards", C.c_uint64), ("live_pairs_backward", C.c_uint64)]


class Gs2dError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"gs2d error {code}: {text}")
        self.code = code


_lib = None


def load_library():
    """dlopen libgs2d.so (built in-tree by __graft_entry__.build() / make -C 2dgaussiansplatting_b200/csrc)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Gs2dError(-2, f"{LIB_PATH} is missing: build it with `make -C {_HERE}/csrc` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i, f = C.c_void_p, C.c_int, C.c_float
    L.gs2d_default_params.argtypes = [C.POINTER(Params)]
    L.gs2d_default_params.restype = None
    L.gs2d_version.restype = i
    L.gs2d_strerror.argtypes = [i]
    L.gs2d_strerror.restype = C.c_char_p
    L.gs2d_last_error.argtypes = [vp]
    L.gs2d_last_error.restype = C.c_char_p
    L.gs2d_create.argtypes = [C.POINTER(vp), i, i, i, i, i, C.POINTER(Params)]
    L.gs2d_destroy.argtypes = [vp]
    L.gs2d_destroy.restype = None
    L.gs2d_set_target_rgba32f.argtypes = [vp, i, vp]
    L.gs2d_set_target_rgba8.argtypes = [vp, i, vp]
    L.gs2d_set_target_rgba8_async.argtypes = [vp, i, vp]
    L.gs2d_init_reference.argtypes = [vp]
    L.gs2d_upload_state.argtypes = [vp, i, vp, vp, f, f, i]
    L.gs2d_download_state.argtypes = [vp, i, vp, vp, C.POINTER(f), C.POINTER(f), C.POINTER(i)]
    L.gs2d_set_optimize_opacity.argtypes = [vp, i]
    L.gs2d_step.argtypes = [vp, i, vp]
    L.gs2d_forward.argtypes = [vp]
    L.gs2d_backward.argtypes = [vp]
    L.gs2d_adam.argtypes = [vp]
    L.gs2d_download_image_rgba32f.argtypes = [vp, i, vp]
    L.gs2d_download_grads.argtypes = [vp, i, vp]
    L.gs2d_last_mse.argtypes = [vp, vp]
    L.gs2d_mse_ticket.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.gs2d_mse_wait.argtypes = [vp, C.c_uint64, vp]
    L.gs2d_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.gs2d_comm_unique_id.argtypes = [vp]
    L.gs2d_comm_init.argtypes = [vp, i, i, vp]
    L.gs2d_set_band.argtypes = [vp, i, i]
    L.gs2d_comm_mode.argtypes = [vp, C.POINTER(i), C.POINTER(i)]
    L.gs2d_comm_traffic.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(i), i]
    L.gs2d_last_step_ms.argtypes = [vp, C.POINTER(f)]
    L.gs2d_profile_step.argtypes = [vp, C.POINTER(f * 6)]
    L.gs2d_set_profiling.argtypes = [vp, i]
    L.gs2d_set_stats.argtypes = [vp, i]
    L.gs2d_set_tuning.argtypes = [C.c_char_p, i]
    L.gs2d_get_profile.argtypes = [vp, C.POINTER(C.c_double * 6), C.POINTER(C.c_longlong)]
    L.gs2d_launch_count.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.gs2d_sync.argtypes = [vp]
    L.gs2d_microbench_fp32.argtypes = [i, C.POINTER(C.c_double)]
    L.gs2d_microbench_hbm.argtypes = [i, C.POINTER(C.c_double)]
    L.gs2d_mse_static.argtypes = [i, vp, vp, i, i, vp, vp, C.c_char_p, C.c_size_t]
    L.gs2d_get_optimize_opacity.argtypes = [vp, C.POINTER(i)]
    L.gs2d_get_dims.argtypes = [vp, C.POINTER(i), C.POINTER(i), C.POINTER(i), C.POINTER(i)]
    L.gs2d_save_checkpoint.argtypes = [vp, C.c_char_p]
    L.gs2d_load_checkpoint.argtypes = [vp, C.c_char_p]
    u8pp = C.POINTER(C.POINTER(C.c_uint8))
    L.gs2d_image_load_rgba8.argtypes = [C.c_char_p, u8pp, C.POINTER(i), C.POINTER(i)]
    L.gs2d_image_decode_jpeg_rgba8.argtypes = [vp, C.c_size_t, u8pp, C.POINTER(i), C.POINTER(i)]
    L.gs2d_image_free.argtypes = [C.POINTER(C.c_uint8)]
    L.gs2d_image_free.restype = None
    L.gs2d_image_write_ppm.argtypes = [C.c_char_p, vp, i, i]
    L.gs2d_overlay_geometry.argtypes = [vp, i, i, vp, vp]
    _lib = L
    return L


def default_params():
    p = Params()
    load_library().gs2d_default_params(C.byref(p))
    return p


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def pcg3d(x, y, z):
    """pcg3d of the reference (main.cpp:17-24), vectorised over uint32 arrays."""
    x = np.asarray(x, np.uint32).copy()
    y = np.asarray(y, np.uint32).copy()
    z = np.asarray(z, np.uint32).copy()
    with np.errstate(over="ignore"):
        x = x * np.uint32(1664525) + np.uint32(1013904223)
        y = y * np.uint32(1664525) + np.uint32(1013904223)
        z = z * np.uint32(1664525) + np.uint32(1013904223)
        x += y * z
        y += z * x
        z += x * y
        x ^= x >> np.uint32(16)
        y ^= y >> np.uint32(16)
        z ^= z >> np.uint32(16)
        x += y * z
        y += z * x
        z += x * y
    return x, y, z


def synth_target(W, H, seed=1, f1=3.0, f2=5.0):
    """Synthetic fit target of SURVEY.md §8(d): smooth sinusoid + pcg3d white noise, RGBA float32, A = 1.

    target(x,y,ch) = clamp(0.5 + 0.25 sin(2 pi (f1 x + f2 y)/W + phi_ch) + 0.25 (pcg3d(x,y,seed).ch / 2^32 - 0.5), 0, 1)
    with phi = (0, 2, 4). Generated on the host once; the CPU oracle and the GPU get the identical bytes.
    """
    out = np.empty((H, W, 4), np.float32)
    xs = np.arange(W, dtype=np.uint32)
    rows = 256
    for y0 in range(0, H, rows):
        ys = np.arange(y0, min(H, y0 + rows), dtype=np.uint32)
        X, Y = np.meshgrid(xs, ys)
        n = pcg3d(X, Y, np.full_like(X, seed))
        phase = 2.0 * np.pi * (f1 * X.astype(np.float64) + f2 * Y.astype(np.float64)) / W
        for ch, phi in enumerate((0.0, 2.0, 4.0)):
            v = 0.5 + 0.25 * np.sin(phase + phi) + 0.25 * (n[ch].astype(np.float64) / 4294967296.0 - 0.5)
            out[y0 : y0 + len(ys), :, ch] = np.clip(v, 0.0, 1.0).astype(np.float32)
    out[..., 3] = 1.0
    return out


class Fitter:
    """One fitting context == the loop-carried state of the reference's main() (main.cpp:271-278) on one B200."""

    def __init__(self, W, H, n_splats, n_images=1, device=0, params=None):
        self.lib = load_library()
        self.W, self.H, self.n_splats, self.n_images = int(W), int(H), int(n_splats), int(n_images)
        self._ctx = C.c_void_p()
        self._rank, self._n_ranks = 0, 1
        self.params = params if params is not None else default_params()
        rc = self.lib.gs2d_create(C.byref(self._ctx), self.W, self.H, self.n_splats, self.n_images, int(device),
                                  C.byref(self.params))
        if rc != GS2D_OK:
            self._ctx = C.c_void_p()
            raise Gs2dError(rc, self.lib.gs2d_strerror(rc).decode())

    # -- plumbing ------------------------------------------------------------------------------------------------
    def _ck(self, rc, allow_nonfinite=False):
        if rc == GS2D_OK or (allow_nonfinite and rc == GS2D_ERR_NONFINITE):
            return rc
        raise Gs2dError(rc, self.lib.gs2d_strerror(rc).decode() + " | " + self.lib.gs2d_last_error(self._ctx).decode())

    def close(self):
        if self._ctx:
            self.lib.gs2d_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- target (imageRef) ---------------------------------------------------------------------------------------
    def set_target(self, rgba32f, image=0):
        a = np.ascontiguousarray(rgba32f, np.float32)
        assert a.shape == (self.H, self.W, 4), a.shape
        self._ck(self.lib.gs2d_set_target_rgba32f(self._ctx, image, _ptr(a)))

    def set_target_u8(self, rgba8, image=0):
        a = np.ascontiguousarray(rgba8, np.uint8)
        assert a.shape == (self.H, self.W, 4), a.shape
        self._ck(self.lib.gs2d_set_target_rgba8(self._ctx, image, _ptr(a)))

    # -- state ---------------------------------------------------------------------------------------------------
    def restart(self):
        """init() of the reference, on the device (main.cpp:280-305)."""
        self._ck(self.lib.gs2d_init_reference(self._ctx))

    def upload(self, splats, adam=None, beta1t=1.0, beta2t=1.0, iteration=0, image=0):
        s = np.ascontiguousarray(splats, SPLAT_DTYPE)
        assert len(s) == self.n_splats
        a = np.ascontiguousarray(adam, ADAM_DTYPE) if adam is not None else None
        self._ck(self.lib.gs2d_upload_state(self._ctx, image, _ptr(s), _ptr(a), beta1t, beta2t, iteration))

    def download(self, image=0, with_adam=True):
        s = np.zeros(self.n_splats, SPLAT_DTYPE)
        a = np.zeros(self.n_splats, ADAM_DTYPE) if with_adam else None
        b1, b2, it = C.c_float(), C.c_float(), C.c_int()
        rc = self.lib.gs2d_download_state(self._ctx, image, _ptr(s), _ptr(a), C.byref(b1), C.byref(b2), C.byref(it))
        self._ck(rc, allow_nonfinite=True)
        return s, a, b1.value, b2.value, it.value

    @property
    def optimize_opacity(self):
        return bool(self.params.optimize_opacity)

    @optimize_opacity.setter
    def optimize_opacity(self, v):
        self.params.optimize_opacity = int(bool(v))
        self._ck(self.lib.gs2d_set_optimize_opacity(self._ctx, int(bool(v))))

    # -- the loop body -------------------------------------------------------------------------------------------
    def step(self, n_iters=1, want_mse=True):
        """n_iters loop bodies (main.cpp:414-809). Returns mse[n_iters, n_images] (what main.cpp:807 prints), or None
        when want_mse is False (fully asynchronous enqueue). Raises on GS2D_ERR_NONFINITE like the reference aborts."""
        if not want_mse:
            self._ck(self.lib.gs2d_step(self._ctx, n_iters, None))
            return None
        mse = np.zeros((n_iters, self.n_images), np.float64)
        self._ck(self.lib.gs2d_step(self._ctx, n_iters, _ptr(mse)))
        return mse

    def forward(self):
        self._ck(self.lib.gs2d_forward(self._ctx))

    def backward(self):
        self._ck(self.lib.gs2d_backward(self._ctx))

    def adam(self):
        return self._ck(self.lib.gs2d_adam(self._ctx), allow_nonfinite=True)

    def sync(self):
        return self._ck(self.lib.gs2d_sync(self._ctx), allow_nonfinite=True)

    # -- results -------------------------------------------------------------------------------------------------
    def image(self, image=0):
        out = np.empty((self.H, self.W, 4), np.float32)
        self._ck(self.lib.gs2d_download_image_rgba32f(self._ctx, image, _ptr(out)))
        return out

    def grads(self, image=0):
        out = np.zeros(self.n_splats, SPLAT_DTYPE)
        self._ck(self.lib.gs2d_download_grads(self._ctx, image, _ptr(out)))
        return out

    def last_mse(self):
        out = np.zeros(self.n_images, np.float64)
        self._ck(self.lib.gs2d_last_mse(self._ctx, _ptr(out)))
        return out

    def mse_ticket(self):
        """Enqueue the read-back of the newest enqueued iteration's loss; returns a ticket for mse_wait()."""
        t = C.c_uint64()
        self._ck(self.lib.gs2d_mse_ticket(self._ctx, C.byref(t)))
        return t.value

    def mse_wait(self, ticket):
        """Block until the ticket's copy has landed (later iterations keep running); mse[n_images]."""
        out = np.zeros(self.n_images, np.float64)
        self._ck(self.lib.gs2d_mse_wait(self._ctx, C.c_uint64(ticket), _ptr(out)))
        return out

    def stats(self):
        s = Stats()
        self._ck(self.lib.gs2d_get_stats(self._ctx, C.byref(s)))
        return {f: getattr(s, f) for f, _ in Stats._fields_}

    def last_step_ms(self):
        ms = C.c_float()
        self._ck(self.lib.gs2d_last_step_ms(self._ctx, C.byref(ms)))
        return ms.value

    def profile_step(self):
        out = (C.c_float * 6)()
        self._ck(self.lib.gs2d_profile_step(self._ctx, C.byref(out)), allow_nonfinite=True)
        names = ["preprocess_emit", "sort", "forward", "backward", "allreduce", "adam"]
        return dict(zip(names, [float(v) for v in out]))

    GROUPS = ["preprocess_emit", "sort", "forward", "backward", "allreduce", "adam"]

    def set_profiling(self, enable=True):
        """Bracket every iteration of step() with CUDA events on the library's stream (resets the sums)."""
        self._ck(self.lib.gs2d_set_profiling(self._ctx, int(bool(enable))))

    def set_stats(self, enable=True):
        """Work counters (live pairs, staged entries) of the forward blend launched by step(): off by default (the counting
        variant of the kernel is ~10 % slower); forward() / backward() always count."""
        self._ck(self.lib.gs2d_set_stats(self._ctx, int(bool(enable))))

    def get_profile(self):
        """-> ({group: mean ms per iteration}, iterations measured) since set_profiling()."""
        sums, n = (C.c_double * 6)(), C.c_longlong()
        self._ck(self.lib.gs2d_get_profile(self._ctx, C.byref(sums), C.byref(n)), allow_nonfinite=True)
        k = max(n.value, 1)
        return {g: sums[j] / k for j, g in enumerate(self.GROUPS)}, n.value

    def set_target_u8_ptr(self, ptr, image=0):
        """Target from a raw host pointer (e.g. a pinned torch tensor's data_ptr()): H*W*4 bytes RGBA8."""
        self._ck(self.lib.gs2d_set_target_rgba8(self._ctx, image, C.c_void_p(ptr)))

    def set_target_u8_ptr_async(self, ptr, image=0):
        """Non-blocking upload from PINNED host memory; takes effect at the next forward pass (double-buffered, so
        the copy of the next target overlaps the iterations running on the current one)."""
        self._ck(self.lib.gs2d_set_target_rgba8_async(self._ctx, image, C.c_void_p(ptr)))

    def target_bytes_per_upload(self):
        """Bytes one target upload moves on this context: the whole RGBA8 image, or this rank's band once sharded."""
        t0, t1 = (0, (self.H + 15) // 16) if self._n_ranks == 1 else band_rows((self.H + 15) // 16, self._rank, self._n_ranks)
        return (min(t1 * 16, self.H) - min(t0 * 16, self.H)) * self.W * 4

    def launch_count(self):
        n = C.c_uint64()
        self._ck(self.lib.gs2d_launch_count(self._ctx, C.byref(n)))
        return n.value

    # -- checkpoint ----------------------------------------------------------------------------
    def save_checkpoint(self, path):
        """{splats, splatAdams, beta1t, beta2t, iterations} of every image (main.cpp:271-278) -> file."""
        self._ck(self.lib.gs2d_save_checkpoint(self._ctx, os.fsencode(path)))

    def load_checkpoint(self, path):
        self._ck(self.lib.gs2d_load_checkpoint(self._ctx, os.fsencode(path)))

    # -- multi-GPU -----------------------------------------------------------------------------------------------
    def set_band(self, rank, n_ranks):
        self._ck(self.lib.gs2d_set_band(self._ctx, rank, n_ranks))

    def comm_init(self, rank, n_ranks, unique_id):
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        self._ck(self.lib.gs2d_comm_init(self._ctx, rank, n_ranks, buf))
        self._rank, self._n_ranks = rank, n_ranks

    def comm_mode(self):
        """-> (n_ranks, peer_memory): peer_memory is True when the gradient exchange is the fused reduce + Adam +
        broadcast kernel over NVLink peer memory, False when it is the NCCL all-reduce."""
        n, p = C.c_int(), C.c_int()
        self._ck(self.lib.gs2d_comm_mode(self._ctx, C.byref(n), C.byref(p)))
        return n.value, bool(p.value)


def _comm_traffic(self, reset=False):
    """-> (bytes read from peers, bytes written to peers, sparse): NVLink payload of the peer-memory exchange."""
    r, w, s = C.c_uint64(), C.c_uint64(), C.c_int()
    self._ck(self.lib.gs2d_comm_traffic(self._ctx, C.byref(r), C.byref(w), C.byref(s), int(bool(reset))))
    return r.value, w.value, bool(s.value)


Fitter.comm_traffic = _comm_traffic


def _take_image(rc, p, w, h):
    L = load_library()
    if rc != GS2D_OK:
        raise Gs2dError(rc, L.gs2d_strerror(rc).decode())
    out = np.ctypeslib.as_array(p, shape=(h.value, w.value, 4)).copy()
    L.gs2d_image_free(p)
    return out


def load_image(path):
    """Target ingest (Image2DRGBA8::load, prp.cpp:179-193): JPEG decoded with stb_image's arithmetic, or binary PPM.
    -> uint8 [H, W, 4], A = 255. Host code (gs2d_jpeg.cpp), no GPU needed."""
    p, w, h = C.POINTER(C.c_uint8)(), C.c_int(), C.c_int()
    rc = load_library().gs2d_image_load_rgba8(os.fsencode(path), C.byref(p), C.byref(w), C.byref(h))
    return _take_image(rc, p, w, h)


def decode_jpeg(data):
    """JPEG bytes -> uint8 [H, W, 4] (same decoder as load_image)."""
    buf = (C.c_uint8 * len(data)).from_buffer_copy(bytes(data))
    p, w, h = C.POINTER(C.c_uint8)(), C.c_int(), C.c_int()
    rc = load_library().gs2d_image_decode_jpeg_rgba8(buf, len(data), C.byref(p), C.byref(w), C.byref(h))
    return _take_image(rc, p, w, h)


def write_ppm(path, rgba32f):
    """Headless stand-in for tex0->upload( image0 ) (main.cpp:794): 8-bit binary PPM of an RGBA float image."""
    a = np.ascontiguousarray(rgba32f, np.float32)
    assert a.ndim == 3 and a.shape[2] == 4
    rc = load_library().gs2d_image_write_ppm(os.fsencode(path), _ptr(a), a.shape[1], a.shape[0])
    if rc != GS2D_OK:
        raise Gs2dError(rc, load_library().gs2d_strerror(rc).decode())


OVERLAY_DTYPE = np.dtype([("center", "<f4", (2,)), ("axis0", "<f4", (2,)), ("axis1", "<f4", (2,)),
                          ("half_extent", "<f4", (2,))])


def overlay_geometry(splats, nvtx=16):
    """The per-splat overlay main() draws (main.cpp:421-476) as data: (records [n] of OVERLAY_DTYPE, ellipse polylines
    [n, nvtx + 1, 2]); image coordinates, y down."""
    s = np.ascontiguousarray(splats, SPLAT_DTYPE)
    out = np.zeros(len(s), OVERLAY_DTYPE)
    ell = np.zeros((len(s), nvtx + 1, 2), np.float32)
    rc = load_library().gs2d_overlay_geometry(_ptr(s), len(s), nvtx, _ptr(out), _ptr(ell))
    if rc != GS2D_OK:
        raise Gs2dError(rc, load_library().gs2d_strerror(rc).decode())
    return out, ell


def set_tuning(name, value):
    """Developer knob of the native library (process-wide), e.g. set_tuning("GS2D_CULL", 0)."""
    rc = load_library().gs2d_set_tuning(name.encode(), int(value))
    if rc:
        raise Gs2dError(rc, "gs2d_set_tuning")


def band_rows(tiles_y, rank, n_ranks):
    """Tile rows [ty0, ty1) owned by `rank` — the same partition gs2d_comm_init / gs2d_set_band apply on the device."""
    per = (tiles_y + n_ranks - 1) // n_ranks
    return min(rank * per, tiles_y), min((rank + 1) * per, tiles_y)


def comm_unique_id():
    buf = (C.c_uint8 * 128)()
    rc = load_library().gs2d_comm_unique_id(buf)
    if rc != GS2D_OK:
        raise Gs2dError(rc, load_library().gs2d_strerror(rc).decode())
    return bytes(buf)


def microbench(device=0):
    """Measured roofline denominators on this device: (FP32 FMA TFLOP/s, HBM copy GB/s)."""
    L = load_library()
    t, g = C.c_double(), C.c_double()
    rc = L.gs2d_microbench_fp32(device, C.byref(t))
    if rc != GS2D_OK:
        raise Gs2dError(rc, L.gs2d_strerror(rc).decode())
    rc = L.gs2d_microbench_hbm(device, C.byref(g))
    if rc != GS2D_OK:
        raise Gs2dError(rc, L.gs2d_strerror(rc).decode())
    return t.value, g.value


def mse_static(image_ptr, target_ptr, W, H, mse_ptr, stream=None, device=0):
    """gs2d_mse_static: the loss of main.cpp:796-805 over DEVICE buffers of the caller, allocating nothing (scratch in a
    static __device__ array of the library, GS2D_MSE_STATIC_MAX_TILES tiles). Arguments are raw device addresses; the
    result lands in *mse_ptr when the stream reaches it. No counterpart in the reference."""
    L = load_library()
    err = C.create_string_buffer(256)
    rc = L.gs2d_mse_static(int(device), C.c_void_p(image_ptr), C.c_void_p(target_ptr), int(W), int(H),
                           C.c_void_p(mse_ptr), C.c_void_p(stream or 0), err, len(err))
    if rc != GS2D_OK:
        raise Gs2dError(rc, err.value.decode() or L.gs2d_strerror(rc).decode())


MSE_STATIC_MAX_TILES = 65536
