"""rt_readback_begin / rt_readback_end: the read-back in two halves, so that a caller rendering frame after frame overlaps the
copies of frame k with the render of frame k + 1.  Byte work: what arrives must be byte-identical to rt_readback's output of the
same frame, also when the next frame has been rendered (and resolved) in between."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pinned(torch, n):
    return torch.empty(n, dtype=torch.uint8, pin_memory=True).numpy()


@pytest.mark.parametrize("precision", ["F32", "F64"])
def test_split_readback_equals_readback_with_the_next_frame_in_flight(ctx_scene, rt, precision):
    import torch
    ctx = ctx_scene
    prec = getattr(rt, precision)
    spp = 6
    cam = rt.camera_rtiow(333, spp, 10)
    # every expectation is rt_readback of the SAME accumulation buffer (fast mode sums a pixel's samples in no fixed order, so
    # two renders of one seed may differ in a last bit)
    ctx.render(cam, precision=prec, seed=1)
    want1 = ctx.readback(spp, want_sum=False, want_ppm=True)
    H, W = ctx.height, ctx.width
    rgb_a, ppm_a = _pinned(torch, H * W * 3), _pinned(torch, H * W * 12 + 64)
    rgb_b, ppm_b = _pinned(torch, H * W * 3), _pinned(torch, H * W * 12 + 64)
    ctx.readback_begin(spp, rgb_a, ppm_a)                     # frame 1 on its way
    with pytest.raises(RuntimeError, match="in flight"):      # one at a time
        ctx.readback_begin(spp, rgb_b, ppm_b)
    ctx.render(cam, precision=prec, seed=2)                   # frame 2 rendered while frame 1 is copied
    n1 = ctx.readback_end()
    assert n1 == len(want1["ppm"]) and ppm_a[:n1].tobytes() == want1["ppm"]
    assert np.array_equal(rgb_a.reshape(H, W, 3), want1["rgb8"])
    ctx.readback_begin(spp, rgb_b, ppm_b)                     # frame 2; frame 1's buffers stay as they are
    n2 = ctx.readback_end()
    want2 = ctx.readback(spp, want_sum=False, want_ppm=True)
    assert want1["ppm"] != want2["ppm"]
    assert n2 == len(want2["ppm"]) and ppm_b[:n2].tobytes() == want2["ppm"]
    assert np.array_equal(rgb_b.reshape(H, W, 3), want2["rgb8"])
    assert ppm_a[:n1].tobytes() == want1["ppm"]

    # a plain rt_readback between begin and end waits for the copies on the device before it overwrites the frame's bytes
    ctx.render(cam, precision=prec, seed=1)
    want1 = ctx.readback(spp, want_sum=False, want_ppm=True)
    ctx.readback_begin(spp, rgb_a, ppm_a)
    ctx.render(cam, precision=prec, seed=2)
    between = ctx.readback(spp, want_sum=False, want_ppm=True)
    n1 = ctx.readback_end()
    assert n1 == len(want1["ppm"]) and ppm_a[:n1].tobytes() == want1["ppm"]
    assert between["ppm"] == ctx.readback(spp, want_sum=False, want_ppm=True)["ppm"] != want1["ppm"]

    # only one of the two outputs; pageable host memory works as well (the copy is then staged by the driver)
    only_rgb = np.zeros(H * W * 3, np.uint8)
    ctx.readback_begin(spp, only_rgb, None)
    assert ctx.readback_end() == 0
    assert np.array_equal(only_rgb.reshape(H, W, 3), between["rgb8"])


def test_split_readback_refuses_what_it_cannot_serve(ctx_scene, rt):
    import torch
    ctx = ctx_scene
    with pytest.raises(RuntimeError, match="no split read-back"):
        ctx.readback_end()
    cam = rt.camera_rtiow(64, 2, 4)
    ctx.render(cam, precision=rt.F32, seed=0)
    with pytest.raises(RuntimeError, match="neither"):
        ctx.readback_begin(2, None, None)
    small = _pinned(torch, 100)
    with pytest.raises(RuntimeError, match="too small"):
        ctx.readback_begin(2, None, small)
    with pytest.raises(RuntimeError, match="no split read-back"):        # a refused begin leaves nothing in flight
        ctx.readback_end()
    full = _pinned(torch, ctx.height * ctx.width * 12 + 64)
    ctx.readback_begin(2, None, full)
    n = ctx.readback_end()
    assert full[:n].tobytes() == ctx.readback(2, want_sum=False, want_ppm=True)["ppm"]
