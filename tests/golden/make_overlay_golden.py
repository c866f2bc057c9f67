#!/usr/bin/env python
"""Generate tests/golden/c1_overlay.npz from ORACLE LEVEL 0 - the PrimVertex calls of the reference's own, unmodified
main.cpp (main.cpp:442-476: eigen axes, 16-segment ellipse, 1-sigma bounding box of every splat), recorded by the
headless shim (oracle/shim/shim.cpp, PR_PRIMS) in frames 0 and 6 of the default configuration (squirrel_cls_mini.jpg,
1024 splats). Pins gs2d_overlay_geometry to the reference: tests/test_io.py.

Run in the build container (needs /root/reference):   python tests/golden/make_overlay_golden.py
Stored: xy[frame][splat][46][2] float32 (z is 0 everywhere) and rgb[frame][splat][46][3] uint8, in the reference's
draw order per splat: 4 axis vertices, 17 x 2 ellipse segment ends, 4 x 2 box edge ends.
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_binding as ob  # noqa: E402

FRAMES = (0, 6)


def main():
    ob.build()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "prims.bin")
        env = dict(os.environ, PR_DATA="/root/reference/bin", PR_ITERS=str(max(FRAMES) + 1), PR_PRIMS=path,
                   PR_PRIMS_FRAMES=",".join(map(str, FRAMES)))
        import subprocess

        subprocess.run([ob.REF_MAIN], env=env, check=True, capture_output=True)
        rec = np.fromfile(path, dtype=[("f", "<i4"), ("p", "<f4", 3), ("c", "u1", 4)])
    assert len(rec) == len(FRAMES) * 1024 * 46, len(rec)
    assert (rec["p"][:, 2] == 0).all()
    xy = np.stack([rec["p"][rec["f"] == f][:, :2].reshape(1024, 46, 2) for f in FRAMES])
    rgb = np.stack([rec["c"][rec["f"] == f][:, :3].reshape(1024, 46, 3) for f in FRAMES])
    np.savez_compressed(os.path.join(HERE, "c1_overlay.npz"), frames=np.array(FRAMES), xy=xy, rgb=rgb)
    print("c1_overlay.npz:", xy.shape, os.path.getsize(os.path.join(HERE, "c1_overlay.npz")), "bytes")


if __name__ == "__main__":
    main()
