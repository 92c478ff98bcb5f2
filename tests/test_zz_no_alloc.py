"""-m gpu, collected last: b200rt_format_ppm_noalloc (static device scratch, no cudaMalloc) against b200rt_format_ppm.

The entry point has no counterpart in the reference (include/b200rt.h says why it exists); its output must be the file image the
existing entry point produces, byte for byte (the tolerance every output-stage test here uses), on calls in a row that share the
scratch array: the second image is smaller than the first, so stale words of the first call lie behind it."""
import numpy as np
import pytest
from conftest import abi, golden, pkg

pytestmark = pytest.mark.gpu


def frame(W, H, seed):
    rng = np.random.RandomState(seed)
    s = rng.rand(H, W, 3) * 10 ** rng.uniform(-3, 3, size=(H, W, 3))
    s.reshape(-1)[:3] = [np.nan, -3.0, 1e300]  # INT_MIN, 0 and a value far above 255: what operator<<(int) prints for them
    s.reshape(-1)[-1] = 0.0
    return s


def test_format_ppm_noalloc_matches_format_ppm_on_calls_in_a_row():
    import torch

    g = golden("cornell")
    ctx = pkg.backend.Context(0)
    try:
        for k, (W, H) in enumerate(((1200, 675), (640, 360), (37, 29), (1024, 1024))):  # headline frame, smaller ones behind it, the limit
            cam = g.cam_small.copy()
            cam.image_width, cam.image_height = W, H
            ctx.set_camera(cam)
            s = frame(W, H, 100 + k)
            want = ctx.format_ppm(s, 16)  # the existing entry point: host buffers, cudaMalloc'd scratch
            assert want.startswith(f"P3\n{W} {H}\n255\n".encode())
            d_sum = torch.from_numpy(s).to("cuda:0")
            cap = len(want) + 17
            d_out = torch.full((cap,), 0xAB, dtype=torch.uint8, device="cuda:0")
            torch.cuda.synchronize()
            n = ctx.format_ppm_noalloc(d_sum.data_ptr(), 16, d_out.data_ptr(), cap)
            out = d_out.cpu().numpy()
            assert n == len(want)
            assert out[:n].tobytes() == want
            assert (out[n:] == 0xAB).all()  # nothing written past the file
            # exact capacity passes, one byte less is refused with the size needed
            d_out.fill_(0)
            torch.cuda.synchronize()
            assert ctx.format_ppm_noalloc(d_sum.data_ptr(), 16, d_out.data_ptr(), len(want)) == len(want)
            assert d_out[: len(want)].cpu().numpy().tobytes() == want
            with pytest.raises(pkg.backend.B200RTError, match="needs"):
                ctx.format_ppm_noalloc(d_sum.data_ptr(), 16, d_out.data_ptr(), len(want) - 1)
    finally:
        ctx.close()


def test_format_ppm_noalloc_rejects_frames_beyond_the_static_scratch():
    import torch

    g = golden("cornell")
    ctx = pkg.backend.Context(0)
    try:
        cam = g.cam_small.copy()
        cam.image_width, cam.image_height = 1025, 1024
        assert cam.image_width * cam.image_height > abi.NOALLOC_MAX_PIXELS
        ctx.set_camera(cam)
        d_sum = torch.zeros(1025 * 1024 * 3, dtype=torch.float64, device="cuda:0")
        d_out = torch.zeros(64, dtype=torch.uint8, device="cuda:0")
        torch.cuda.synchronize()
        with pytest.raises(pkg.backend.B200RTError, match="B200RT_NOALLOC_MAX_PIXELS"):
            ctx.format_ppm_noalloc(d_sum.data_ptr(), 16, d_out.data_ptr(), 64)
        with pytest.raises(pkg.backend.B200RTError):
            ctx.format_ppm_noalloc(0, 16, d_out.data_ptr(), 64)
    finally:
        ctx.close()
