"""the windowed demo's compute stage (reference assets/shaders/ray_tracing.comp) as a CUDA
kernel behind rt_demo_compute.  Oracle = oracle/rt_oracle.c:o_demo_compute, pinned here against the shader
ITSELF compiled as C++ (oracle/glsl_harness.cpp -> fixtures in tests/golden/ref_demo_shader.npz, made by
tools/make_golden.py; live too where oracle/_ref/libref_glsl.so exists).  Integer / byte output: bit-exact."""
import hashlib

import numpy as np
import pytest

from conftest import golden


def demo_scene(O, n=20, seed=3):
    """n spheres in front of the demo's camera (it looks down +z from the origin); sphere_count 20 is
    src/common_objects.h:10."""
    rg = np.random.default_rng(seed)
    s = np.zeros(n, O.DEMO_SPHERE)
    s["center"] = np.stack([rg.uniform(-2, 2, n), rg.uniform(-1.2, 1.2, n), rg.uniform(2, 8, n)], -1)
    s["radius"] = rg.uniform(0.1, 0.8, n)
    s["color"] = np.concatenate([rg.uniform(0, 1, (n, 3)), np.ones((n, 1))], -1)
    return s


# ------------------------------------------------------------------ oracle vs the shader (CPU) --
def test_oracle_matches_shader_fixture(O):
    g = golden("ref_demo_shader.npz")
    assert np.array_equal(O.demo_compute(160, 120), g["small"])
    full = O.demo_compute(1024, 768)
    assert hashlib.sha256(full.tobytes()).digest() == g["full_sha256"].tobytes()
    assert int((full[..., 0] == 255).sum()) == int(g["full_hit_pixels"])
    # the background the shader stores: (0.5, 0, 0.25, 1) as rgba8
    assert tuple(full[0, 0]) == (128, 0, 64, 255)


def test_oracle_hit_matches_shader_fixture(O):
    g = golden("ref_demo_shader.npz")
    got = O.demo_hit(g["origins"], g["directions"], g["centers"], g["radii"])
    assert np.array_equal(got, g["hit"])
    assert 100 < int(got.sum()) < len(got) - 100          # both outcomes are exercised


def test_oracle_matches_shader_live(O):
    if not O.have_ref_glsl():
        pytest.skip("oracle/_ref/libref_glsl.so not built (reference tree absent)")
    for w, h in ((1, 1), (33, 7), (160, 120), (640, 480)):
        assert np.array_equal(O.demo_compute(w, h), O.demo_compute(w, h, impl="ref"))
    rg = np.random.default_rng(5)
    n = 3000
    o = rg.standard_normal((n, 3)).astype("f4"); d = rg.standard_normal((n, 3)).astype("f4")
    c = (2 * rg.standard_normal((n, 3))).astype("f4"); r = rg.uniform(0.1, 3, n).astype("f4")
    assert np.array_equal(O.demo_hit(o, d, c, r), O.demo_hit(o, d, c, r, impl="ref"))


def test_oracle_sphere_list_is_the_hit_statement_per_sphere(O):
    """The list form is the shader's `if (hit) pixel_color = sphere.color` once per sphere: each pixel carries
    the colour of the last sphere whose hit() is true for that pixel's ray."""
    W, H = 96, 64
    s = demo_scene(O, 7)
    img = O.demo_compute(W, H, s)
    ys, xs = np.mgrid[0:H, 0:W]
    hc = ((xs.astype("f4") * np.float32(2) - np.float32(W)) / np.float32(W)).astype("f4")
    vc = ((ys.astype("f4") * np.float32(2) - np.float32(H)) / np.float32(W)).astype("f4")
    d = np.stack([hc, vc, np.ones_like(hc)], -1).reshape(-1, 3)
    want = np.tile(np.array([128, 0, 64, 255], np.uint8), (W * H, 1))
    for k in range(len(s)):
        hit = O.demo_hit(np.zeros_like(d), d, np.tile(s["center"][k], (len(d), 1)), np.full(len(d), s["radius"][k], "f4"))
        want[hit] = np.rint(np.clip(s["color"][k], 0, 1) * np.float32(255)).astype(np.uint8)
    assert np.array_equal(img.reshape(-1, 4), want)
    assert len(np.unique(img.reshape(-1, 4), axis=0)) > 3


def test_abi_declares_demo_compute(rt):
    assert "rt_demo_compute" in rt.SYMBOLS
    assert rt.DEMO_SPHERE.itemsize == 32
    assert hasattr(rt.lib(), "rt_demo_compute")


# ------------------------------------------------------------------------ CUDA vs oracle (GPU) --
@pytest.mark.gpu
@pytest.mark.parametrize("size", [(1, 1), (33, 7), (160, 120), (1024, 768), (3840, 2160)])
def test_gpu_demo_compute_hard_coded_sphere(ctx, O, size):
    w, h = size
    got = ctx.demo_compute(w, h)
    assert np.array_equal(got, O.demo_compute(w, h))
    if size == (160, 120):
        assert np.array_equal(got, golden("ref_demo_shader.npz")["small"])
    if size == (1024, 768):
        assert hashlib.sha256(got.tobytes()).digest() == golden("ref_demo_shader.npz")["full_sha256"].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 20, 500])
def test_gpu_demo_compute_sphere_list(ctx, O, rt, n):
    s = demo_scene(O, n) if n else np.zeros(0, O.DEMO_SPHERE)
    got = ctx.demo_compute(1024, 768, s.view(rt.DEMO_SPHERE))
    want = O.demo_compute(1024, 768, s)
    assert np.array_equal(got, want)
    if n >= 20:
        assert (got[..., :3] != np.array([128, 0, 64], np.uint8)).any()


@pytest.mark.gpu
def test_gpu_demo_compute_rejects_bad_sizes(ctx, rt):
    with pytest.raises(rt.RtError):
        ctx.demo_compute(0, 10)
