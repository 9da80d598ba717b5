"""rt_encode_p3_static -- the encoder for callers that may not allocate (scratch in a static __device__ array inside the
library; no counterpart in the reference).  Checked through the C ABI against
the existing entry point rt_encode_p3 (byte work: the bar is byte-identical), on calls in a row that share the scratch.
Collected last on purpose (the file name sorts after every other test file)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PIXELS = 1 << 24          # RT_STATIC_MAX_PIXELS, include/rt_b200.h


def _encode_static(ctx, torch, rgb8):
    """one call of the new entry point on device copies of rgb8 -> the text as bytes"""
    h, w, _ = rgb8.shape
    cap = 32 + 12 * w * h
    d_in = torch.from_numpy(np.ascontiguousarray(rgb8).reshape(-1)).cuda()
    d_out = torch.full((cap,), 0x55, dtype=torch.uint8, device="cuda")      # not zero: stale bytes must not leak into the text
    d_len = torch.zeros(1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()                                                 # the library works on its own stream
    ctx.encode_p3_static(d_in.data_ptr(), w, h, d_out.data_ptr(), cap, d_len.data_ptr())
    ctx.synchronize()
    n = int(d_len.item())
    assert 0 < n <= cap
    return d_out[:n].cpu().numpy().tobytes()


def test_static_scratch_encoder_matches_the_allocating_one_on_calls_in_a_row(rt):
    import torch
    ctx = rt.Context(0)
    try:
        rng = np.random.default_rng(7)
        # two different frames back to back (the second finds the first one's scratch contents), the first again, then a
        # ragged size (pixel count not a multiple of 4 nor of the 1024-pixel tile) and a single pixel
        a = rng.integers(0, 256, size=(270, 480, 3), dtype=np.uint8)
        b = rng.integers(0, 256, size=(1080, 1920, 3), dtype=np.uint8)
        b[:200] = rng.integers(0, 10, size=(200, 1920, 3), dtype=np.uint8)     # short lines: offsets differ from frame to frame
        c = rng.integers(0, 256, size=(37, 53, 3), dtype=np.uint8)
        d = np.array([[[0, 9, 255]]], dtype=np.uint8)
        for img in (a, b, a, c, d, b):
            want = ctx.encode_p3(img)                  # existing entry point (allocates its own device buffers)
            got = _encode_static(ctx, torch, img)
            assert got == want, f"{img.shape}: static-scratch text differs from rt_encode_p3"
            assert got.startswith(b"P3\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        # an unaligned input address takes the byte-load path
        flat = torch.from_numpy(np.concatenate([np.zeros(1, np.uint8), c.reshape(-1)])).cuda()
        cap = 32 + 12 * c.shape[0] * c.shape[1]
        out = torch.zeros(cap, dtype=torch.uint8, device="cuda")
        ln = torch.zeros(1, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        ctx.encode_p3_static(flat.data_ptr() + 1, c.shape[1], c.shape[0], out.data_ptr(), cap, ln.data_ptr())
        ctx.synchronize()
        assert out[:int(ln.item())].cpu().numpy().tobytes() == ctx.encode_p3(c)
    finally:
        ctx.close()


def test_static_scratch_encoder_refuses_what_it_cannot_serve(rt):
    import torch
    ctx = rt.Context(0)
    try:
        buf = torch.zeros(4096, dtype=torch.uint8, device="cuda")
        ln = torch.zeros(1, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        with pytest.raises(RuntimeError, match="exceed the static scratch"):          # one pixel row too many: 4096 x 4097
            ctx.encode_p3_static(buf.data_ptr(), 4096, MAX_PIXELS // 4096 + 1, buf.data_ptr(), 1 << 40, ln.data_ptr())
        with pytest.raises(RuntimeError, match="below the worst case"):
            ctx.encode_p3_static(buf.data_ptr(), 16, 16, buf.data_ptr(), 16 * 16 * 12, ln.data_ptr())
        with pytest.raises(RuntimeError):
            ctx.encode_p3_static(0, 16, 16, buf.data_ptr(), 4096, ln.data_ptr())
        ctx.synchronize()
    finally:
        ctx.close()
