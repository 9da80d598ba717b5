"""P6 / PFM next to the reference's P3 text, progressive accumulation with a headless
rgba8 preview buffer (the compute-side stand-in for start_interactive_loop, src/graphical_environment.h:31).
Bytes are compared with the oracle's write_color / sums (oracle/oracle.py: ppm_p3, ppm_p6, pfm)."""
import numpy as np
import pytest


def test_abi_declares_the_format_entry_points(rt):
    for name in ("rt_readback_image", "rt_preview"):
        assert name in rt.SYMBOLS and hasattr(rt.lib(), name)
    assert (rt.IMG_P3, rt.IMG_P6, rt.IMG_PFM) == (0, 1, 2)


def test_oracle_p6_and_pfm_layout(O):
    rg = np.random.default_rng(0)
    rgb8 = rg.integers(0, 256, (3, 5, 3), np.uint8)
    p6 = O.ppm_p6(rgb8)
    assert p6.startswith(b"P6\n5 3\n255\n") and len(p6) == 11 + 45 and p6[11:] == rgb8.tobytes()
    sums = rg.uniform(0, 4, (3, 5, 3))
    f = O.pfm(sums, 4)
    assert f.startswith(b"PF\n5 3\n-1.0\n")
    body = np.frombuffer(f[len(b"PF\n5 3\n-1.0\n"):], "<f4").reshape(3, 5, 3)
    assert np.array_equal(body[::-1], (sums * 0.25).astype("f4"))       # bottom row first


@pytest.mark.gpu
@pytest.mark.parametrize("width,spp", [(160, 4), (97, 3), (1, 2)])
def test_gpu_formats_match_oracle_bytes(ctx_scene, rt, O, scene, width, spp):
    s, m = scene
    cam = rt.camera_rtiow(width, spp, 50)
    ctx_scene.render(cam, precision=rt.F64, seed=21)
    ref = O.render(s, m, O.camera_rtiow(width, spp, 50), rng_mode=O.RNG_PHILOX, seed=21, use_bvh=True)
    assert ctx_scene.readback_image(rt.IMG_P3, spp) == O.ppm_p3(ref["rgb8"])
    assert ctx_scene.readback_image(rt.IMG_P6, spp) == O.ppm_p6(ref["rgb8"])
    got = ctx_scene.readback_image(rt.IMG_PFM, spp)
    want = O.pfm(ref["sum"], spp)
    hl = want.index(b"-1.0\n") + 5
    assert got[:hl] == want[:hl] and len(got) == len(want)
    a, b = np.frombuffer(got[hl:], "<f4"), np.frombuffer(want[hl:], "<f4")
    assert np.allclose(a, b, rtol=3e-7, atol=1e-12)          # sums agree to 1e-13 rel (parity mode); one float rounding on top
    assert (a == b).mean() > 0.999
    dev, img = ctx_scene.preview(spp)
    assert dev and np.array_equal(img[..., :3], ref["rgb8"]) and (img[..., 3] == 255).all()


@pytest.mark.gpu
def test_gpu_small_buffer_and_unknown_format_are_errors(ctx_scene, rt):
    import ctypes as C
    cam = rt.camera_rtiow(64, 1, 5)
    ctx_scene.render(cam, precision=rt.F32, seed=0)
    with pytest.raises(rt.RtError):
        ctx_scene.readback_image(7, 1)
    n = C.c_size_t(10)
    buf = np.zeros(10, np.uint8)
    rc = rt.lib().rt_readback_image(ctx_scene._h, 1, rt.IMG_P6, buf.ctypes.data_as(C.c_void_p), C.byref(n))
    assert rc != 0 and n.value == len(b"P6\n64 36\n255\n") + 64 * 36 * 3


@pytest.mark.gpu
@pytest.mark.parametrize("precision", ["f64", "f32"])
def test_gpu_progressive_passes_preview_every_prefix(ctx_scene, rt, O, scene, precision):
    """Render [0,2) [2,3) [3,8) into the same buffer; after every pass the preview is the image of the samples so
    far (divisor = samples in the buffer), and the last one is the one-shot 8 spp render."""
    s, m = scene
    prec = rt.F64 if precision == "f64" else rt.F32
    cam = rt.camera_rtiow(120, 8, 50)
    done = 0
    for k, count in enumerate((2, 1, 5)):
        ctx_scene.render(cam, precision=prec, seed=6, sample_begin=done, sample_count=count, accumulate=k > 0)
        done += count
        _, img = ctx_scene.preview(0)
        ref = O.render(s, m, O.camera_rtiow(120, 8, 50), rng_mode=O.RNG_PHILOX, seed=6, use_bvh=True, sample0=0, nsamples=done)
        d = np.abs(img[..., :3].astype(int) - ref["rgb8"].astype(int))
        if precision == "f64":
            assert d.max() <= 1 and (d > 0).any(-1).mean() < 1e-3       # pass-wise sums re-associate the last bit only
        else:
            assert (d <= 1).all(-1).mean() >= 0.999
    ctx_scene.render(cam, precision=prec, seed=6)
    _, one_shot = ctx_scene.preview(8)
    d = np.abs(one_shot.astype(int) - img.astype(int))
    assert d.max() <= 1 and (d > 0).any(-1).mean() < 1e-3
