"""gs2d_mse_static (include/gs2d.h): the loss of main.cpp:796-805 computed without any allocation, its scratch in a
static __device__ array of the library. Checked through the C ABI against the existing entry points
(gs2d_forward + gs2d_last_mse) on the image and target those used, on two calls in a row that share the scratch.

Tolerance: the new kernels repeat the summation order of the product path, so the values are required to be
BIT-IDENTICAL (the integer/byte bar of the other parity tests; no floating-point slack is needed).

The file is named so that pytest collects it last.
"""
import numpy as np
import pytest

from gs2d_loader import gs2d

pytestmark = pytest.mark.gpu


def forward_of_a_fit(W, H, n, seed, iters):
    """(image, target, mse) of the existing entry points after `iters` iterations of a small fit."""
    target = gs2d.synth_target(W, H, seed=seed)
    with gs2d.Fitter(W, H, n) as f:
        f.set_target(target)
        f.restart()
        if iters:
            f.step(iters)
        f.forward()
        mse = float(f.last_mse()[0])
        image = f.image()
    return image, target, mse


def test_static_scratch_mse_equals_the_existing_entry_point_twice_in_a_row():
    import torch

    dev = torch.device("cuda:0")
    # the larger image first: the second call finds the first call's partial sums beyond its own range in the scratch
    cases = [(333, 270, 700, 3, 2), (100, 37, 64, 5, 0)]
    expected = [forward_of_a_fit(*c) for c in cases]
    out = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
    held = []
    stream = torch.cuda.current_stream(dev).cuda_stream
    for k, ((W, H, _, _, _), (image, target, _)) in enumerate(zip(cases, expected)):
        img_d = torch.from_numpy(image).to(dev)
        tgt_d = torch.from_numpy(target).to(dev)
        held += [img_d, tgt_d]
        gs2d.mse_static(img_d.data_ptr(), tgt_d.data_ptr(), W, H, out.data_ptr() + 8 * k, stream=stream)
    torch.cuda.synchronize(dev)
    got = out.cpu().numpy()
    for k, (_, _, mse) in enumerate(expected):
        assert mse > 0.0
        assert got[k] == mse, (k, got[k], mse)
    # and against the formula itself in float64 (the device sums float32 squares: 1e-6 relative is generous)
    for k, (image, target, _) in enumerate(expected):
        d = (image[..., :3].astype(np.float64) - target[..., :3].astype(np.float64)) * 255.0
        assert abs(got[k] - np.mean(d * d)) <= 1e-6 * np.mean(d * d)

    # the same inputs again, in the other order, on the default stream: same bits
    out2 = torch.full((2,), -1.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    for k in (1, 0):
        W, H = cases[k][0], cases[k][1]
        gs2d.mse_static(held[2 * k].data_ptr(), held[2 * k + 1].data_ptr(), W, H, out2.data_ptr() + 8 * k)
    torch.cuda.synchronize(dev)
    assert np.array_equal(out2.cpu().numpy(), got)


def test_static_scratch_mse_rejects_what_does_not_fit():
    import torch

    dev = torch.device("cuda:0")
    buf = torch.zeros(64, dtype=torch.float32, device=dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    # 4112 x 4096 pixels = 257 x 256 tiles > GS2D_MSE_STATIC_MAX_TILES: refused before anything is launched or read
    with pytest.raises(gs2d.Gs2dError) as e:
        gs2d.mse_static(buf.data_ptr(), buf.data_ptr(), 4112, 4096, out.data_ptr())
    assert e.value.code == gs2d.GS2D_ERR_INVALID and "too large" in str(e.value) and "65536" in str(e.value)
    with pytest.raises(gs2d.Gs2dError):
        gs2d.mse_static(0, buf.data_ptr(), 16, 16, out.data_ptr())
    with pytest.raises(gs2d.Gs2dError):
        gs2d.mse_static(buf.data_ptr() + 4, buf.data_ptr(), 2, 2, out.data_ptr())
    torch.cuda.synchronize(dev)
    assert float(out.cpu()[0]) == 0.0
