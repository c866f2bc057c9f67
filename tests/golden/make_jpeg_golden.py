#!/usr/bin/env python
"""Generate tests/golden/jpeg/*: small JPEG files covering the coding modes the ingest decoder supports, and the pixels
the REFERENCE's own decoder (oracle/_ref/stb_decode = the vendored stb_image.h v2.23, `make -C oracle ref`) produces
for them. Run in the build container (needs /root/reference and PIL):  python tests/golden/make_jpeg_golden.py

Every case is <name>.jpg + one entry `<name>` (H x W x 4 uint8, as stbi_load(..., 4)) in expected.npz.
"""
import io
import os
import subprocess
import tempfile

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "jpeg")
STB = os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle", "_ref", "stb_decode")


def picture(w, h, seed):
    """Smooth gradients + a sharp edge + noise: exercises every IDCT coefficient and the chroma filters."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    r = 127 + 120 * np.sin(x / 5.0 + seed) * np.cos(y / 7.0)
    g = 255 * (x + y) / max(w + h - 2, 1)
    b = np.where((x // 8 + y // 8) % 2 == 0, 230.0, 20.0)
    img = np.stack([r, g, b], -1) + rng.normal(0, 12, (h, w, 3))
    return np.clip(img, 0, 255).astype(np.uint8)


# name: (w, h, mode, PIL save kwargs)
CASES = {
    "base_444": (37, 29, "RGB", dict(quality=90, subsampling=0)),
    "base_422": (37, 29, "RGB", dict(quality=85, subsampling=1)),
    "base_420": (37, 29, "RGB", dict(quality=75, subsampling=2)),
    "base_420_w1": (1, 9, "RGB", dict(quality=75, subsampling=2)),
    "base_420_even": (64, 48, "RGB", dict(quality=95, subsampling=2)),
    "base_gray": (23, 31, "L", dict(quality=80)),
    "prog_444": (37, 29, "RGB", dict(quality=90, subsampling=0, progressive=True)),
    "prog_420": (45, 33, "RGB", dict(quality=70, subsampling=2, progressive=True)),
    "prog_gray": (23, 31, "L", dict(quality=60, progressive=True)),
    "base_420_rst": (70, 50, "RGB", dict(quality=80, subsampling=2, restart_marker_blocks=3)),
    "prog_420_rst": (70, 50, "RGB", dict(quality=80, subsampling=2, progressive=True, restart_marker_blocks=2)),
    "base_444_q25_opt": (40, 40, "RGB", dict(quality=25, subsampling=0, optimize=True)),
    "base_420_q100": (33, 17, "RGB", dict(quality=100, subsampling=2)),
}


def main():
    os.makedirs(OUT, exist_ok=True)
    expected = {}
    for seed, (name, (w, h, mode, kw)) in enumerate(CASES.items()):
        img = picture(w, h, seed)
        pil = Image.fromarray(img if mode == "RGB" else img[..., 1], mode)
        buf = io.BytesIO()
        pil.save(buf, "JPEG", **kw)
        path = os.path.join(OUT, name + ".jpg")
        open(path, "wb").write(buf.getvalue())
        with tempfile.NamedTemporaryFile(suffix=".bin") as t:
            subprocess.run([STB, path, t.name], check=True)
            raw = np.fromfile(t.name, np.uint8)
        ww, hh = np.frombuffer(raw[:8].tobytes(), np.int32)
        assert (ww, hh) == (w, h)
        expected[name] = raw[8:].reshape(h, w, 4)
        print(f"{name}: {w}x{h} {len(buf.getvalue())} bytes")
    np.savez_compressed(os.path.join(OUT, "expected.npz"), **expected)


if __name__ == "__main__":
    main()
