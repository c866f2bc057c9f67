#!/usr/bin/env python
"""Generate tests/golden/* from ORACLE LEVEL 0 — the reference's own, unmodified main.cpp
(oracle/_ref/ref_main, built by `make -C oracle ref` from /root/reference/main.cpp).

Run in the build container (needs /root/reference for the two JPEGs):
    python tests/golden/make_golden.py

Writes, per config (C1 = squirrel_cls_mini.jpg 268x213, C2 = squirrel_cls.jpg 535x426; N=1024 both, main.cpp:271):
    <cfg>_target_rgb8.npz   the target as decoded by the reference's stb_image (alpha is 255 everywhere)
    <cfg>_mse.txt           the reference's stdout, one "%d itr, mse %.4f" line per iteration   (main.cpp:807)
    <cfg>_image_sha256.txt  sha256 of the raw RGBA32F bytes of the image handed to tex0->upload (main.cpp:794),
                            one line per iteration -> pins the render bit-exactly at every iteration, which from
                            iteration 1 on transitively pins gradients + Adam + clamps
    c1_image_itr0.npz       the iteration-0 render itself (RGB float32), C1 only
The values depend on this image's glibc sinf/cosf; the GPU box runs the same image.
"""
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_binding as ob  # noqa: E402

REF_BIN = "/root/reference/bin"
CONFIGS = {"c1": ("squirrel_cls_mini.jpg", 300), "c2": ("squirrel_cls.jpg", 60)}


def main():
    ob.build()
    for cfg, (jpeg, iters) in CONFIGS.items():
        with tempfile.TemporaryDirectory() as d:
            tpath = os.path.join(d, "target.bin")
            lines = ob.run_level0(REF_BIN, iters, image=jpeg, dump_dir=d, target_path=tpath)
            raw = np.fromfile(tpath, np.uint8)
            w, h = np.frombuffer(raw[:8].tobytes(), np.int32)
            rgba = raw[8:].reshape(h, w, 4)
            assert (rgba[..., 3] == 255).all()
            np.savez_compressed(os.path.join(HERE, f"{cfg}_target_rgb8.npz"), rgb=rgba[..., :3])
            with open(os.path.join(HERE, f"{cfg}_mse.txt"), "w") as f:
                f.write("\n".join(lines) + "\n")
            with open(os.path.join(HERE, f"{cfg}_image_sha256.txt"), "w") as f:
                for it in range(iters):
                    b = open(os.path.join(d, f"itr{it:05d}.rgba32f"), "rb").read()
                    assert len(b) == w * h * 16
                    f.write(hashlib.sha256(b).hexdigest() + "\n")
            if cfg == "c1":
                img0 = np.fromfile(os.path.join(d, "itr00000.rgba32f"), np.float32).reshape(h, w, 4)
                np.savez_compressed(os.path.join(HERE, f"{cfg}_image_itr0.npz"), rgb=img0[..., :3])
            print(cfg, jpeg, f"{w}x{h}", iters, "iterations;", lines[0], "...", lines[-1])


if __name__ == "__main__":
    main()
