"""The float fast paths of the framebuffer passes (rt_frame.cu resolve_bytes_f32_kernel, with or without the character count: float sums,
sqrt.approx, the exact write_color arithmetic of src/color.h:14-35 only near a byte boundary) must give the SAME bytes and the
same P3 text as the general kernels, which follow the reference's double arithmetic on every channel.  Byte work: bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both_ways(ctx, spp_total):
    """(rgb8 by the bytes-only fast kernel, rgb8 + text by the counting resolve and the text writer, rgb8 + text by the general kernels)"""
    a = ctx.readback(spp_total, want_sum=False)                       # resolve_bytes_f32_kernel
    b = ctx.readback(spp_total, want_sum=False, want_ppm=True)        # counting resolve + LUT text writer (two passes)
    c = ctx.readback(spp_total, want_sum=True)                        # a sum is wanted: resolve_kernel (double sums, every channel)
    return a["rgb8"], b["rgb8"], b["ppm"], c["rgb8"], c["sum"]


@pytest.mark.parametrize("slices", [1, 3, 8])
def test_fast_resolve_of_a_rendered_frame_equals_the_general_kernels(ctx_scene, rt, slices):
    ctx = ctx_scene
    spp = 16
    cam = rt.camera_rtiow(333, spp, 12)                              # 333 x 187: the pixel count is not a multiple of 4
    ctx.render(cam, precision=rt.F32, seed=3, variant=rt.WAVEFRONT, slices=slices)
    assert ctx.stats()["slices"] == slices and ctx.accum_buffer()[1] == slices * ctx.width * ctx.height * 4
    for total in (spp, 0):                                           # 0: divide by the samples in the buffer
        fast, fused, text, exact, sums = _both_ways(ctx, total)
        assert np.array_equal(fast, exact)
        assert np.array_equal(fused, exact)
        assert text == ctx.encode_p3(exact)
        # and the general kernel's bytes are write_color of its sums (rt_write_color: the double path on its own)
        assert np.array_equal(exact.reshape(-1, 3), ctx.write_color(sums.reshape(-1, 3), spp))


@pytest.mark.parametrize("slices", [1, 4])
def test_fast_resolve_on_byte_boundaries(ctx_scene, rt, slices):
    """accumulation buffers written by the test: sums that sit exactly on a byte boundary of write_color ((k / 256)^2 x spp), one
    float step to either side, zeros, values above 1, tiny values, NaN"""
    import torch
    from raytracingproject_b200 import sharding
    ctx = ctx_scene
    spp = 8
    W = 256
    cam = rt.camera_rtiow(W, spp, 2)
    ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT, slices=slices)
    ctx.synchronize()
    H = ctx.height
    acc = sharding.alias_accum_tensor(ctx).view(slices, H, W, 4)
    k = np.arange(256, dtype=np.float64)
    edge = ((k / 256.0) ** 2 * spp).astype(np.float32)               # exact in float (k^2 / 2^13)
    rows = [edge, np.nextafter(edge, np.float32(np.inf)), np.nextafter(edge, np.float32(-np.inf)), np.zeros(256, np.float32),
            np.linspace(0.9, 3.0, 256).astype(np.float32) * spp, np.full(256, 1e-30, np.float32), np.full(256, np.nan, np.float32)]
    rng = np.random.default_rng(5)
    img = rng.random((H, W, 3), dtype=np.float32) * spp
    for r, row in enumerate(rows):
        for c in range(3):
            img[3 * r + c, :, :] = rng.random((W, 3), dtype=np.float32) * spp
            img[3 * r + c, :, c] = np.maximum(row, 0) if not np.isnan(row).any() else row
    # split the sums over the layers so that the float additions are exact for the boundary rows (halves / quarters)
    parts = np.zeros((slices, H, W, 4), np.float32)
    for s in range(slices):
        parts[s, :, :, :3] = img / slices
        parts[s, :, :, 3] = spp / slices
    acc.copy_(torch.from_numpy(parts).cuda())
    torch.cuda.synchronize()
    fast, fused, text, exact, sums = _both_ways(ctx, spp)
    assert np.array_equal(fast, exact)
    assert np.array_equal(fused, exact)
    assert text == ctx.encode_p3(exact)
    good = ~np.isnan(sums).any(-1)
    # the boundary row itself: byte k exactly (int(256 * sqrt((k/256)^2)) = k), 255 from 0.999 on (src/color.h:31)
    assert np.array_equal(exact[0, :, 0], np.minimum(np.arange(256), 255).astype(np.uint8))
    assert good.sum() >= (H - 3) * W

