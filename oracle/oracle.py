"""ctypes/numpy bindings for the CPU oracle (TEST INFRASTRUCTURE, not product code).

Two libraries are wrapped with the same array conventions:

* ``liboracle.so``           -- rt_oracle.c, the plain-C restatement (always buildable)
* ``_ref/libref_harness.so`` -- the reference itself compiled in place from
  /root/reference/src (built where the reference exists; the .so travels)

Only tests/, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_ORACLE = os.path.join(HERE, "liboracle.so")
LIB_REF = os.path.join(HERE, "_ref", "libref_harness.so")
LIB_REF_GLSL = os.path.join(HERE, "_ref", "libref_glsl.so")
DEMO_SPHERE = np.dtype([("center", "f4", 3), ("radius", "f4"), ("color", "f4", 4)])    # src/common_objects.h:17-22

RNG_MT19937, RNG_PHILOX = 0, 1
LAMBERTIAN, METAL, DIELECTRIC = 0, 1, 2

VEC3 = np.dtype([("x", "f8"), ("y", "f8"), ("z", "f8")])
RAY = np.dtype([("orig", VEC3), ("dir", VEC3), ("tm", "f8")])
MATERIAL = np.dtype([("type", "i4"), ("pad", "i4"), ("albedo", VEC3), ("param", "f8")])
SPHERE = np.dtype([("center1", VEC3), ("center_vec", VEC3), ("radius", "f8"),
                   ("is_moving", "i4"), ("mat", "i4")])
HIT = np.dtype([("p", VEC3), ("normal", VEC3), ("t", "f8"), ("front_face", "i4"), ("prim", "i4")])
CAMERA = np.dtype([("aspect_ratio", "f8"), ("image_width", "i4"), ("samples_per_pixel", "i4"),
                   ("max_depth", "i4"), ("pad0", "i4"), ("vfov", "f8"),
                   ("lookfrom", VEC3), ("lookat", VEC3), ("vup", VEC3),
                   ("defocus_angle", "f8"), ("focus_dist", "f8"),
                   ("image_height", "i4"), ("pad1", "i4"),
                   ("center", VEC3), ("pixel00_loc", VEC3), ("pixel_delta_u", VEC3),
                   ("pixel_delta_v", VEC3), ("u", VEC3), ("v", VEC3), ("w", VEC3),
                   ("defocus_disk_u", VEC3), ("defocus_disk_v", VEC3)])
RENDER_OPTS = np.dtype([("rng_mode", "i4"), ("nthreads", "i4"), ("seed", "u8"),
                        ("sample0", "i4"), ("nsamples", "i4"), ("row0", "i4"), ("row1", "i4"),
                        ("mt_skip", "u8")])


class _World(C.Structure):
    _fields_ = [("spheres", C.c_void_p), ("materials", C.c_void_p),
                ("n_spheres", C.c_size_t), ("bvh", C.c_void_p)]


def build(force=False):
    """Compile liboracle.so (and _ref/ when /root/reference is present)."""
    need = force or not os.path.exists(LIB_ORACLE) or (
        os.path.getmtime(LIB_ORACLE) < max(os.path.getmtime(os.path.join(HERE, f))
                                           for f in ("rt_oracle.c", "rt_oracle.h")))
    need_ref = os.path.isdir("/root/reference/src") and (
        force or not os.path.exists(LIB_REF) or
        os.path.getmtime(LIB_REF) < os.path.getmtime(os.path.join(HERE, "ref_harness.cpp")) or
        not os.path.exists(LIB_REF_GLSL) or
        os.path.getmtime(LIB_REF_GLSL) < os.path.getmtime(os.path.join(HERE, "glsl_harness.cpp")))
    if need or need_ref:
        subprocess.run(["make", "-C", HERE, "all"], check=True, capture_output=True)


def vec3s(a):
    """structured VEC3 array -> (..., 3) float64 view/copy"""
    return np.stack([a["x"], a["y"], a["z"]], axis=-1)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        build()
        L = C.CDLL(LIB_ORACLE)
        L.o_scene_rtiow.restype = C.c_size_t
        L.o_scene_rtiow.argtypes = [C.c_int, C.c_double, C.c_double, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.o_camera_rtiow.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.o_camera_initialize.argtypes = [C.c_void_p]
        L.o_bvh_build.restype = C.c_void_p
        L.o_bvh_build.argtypes = [C.c_void_p, C.c_size_t]
        L.o_bvh_free.argtypes = [C.c_void_p]
        L.o_render.argtypes = [C.c_void_p] * 6
        L.o_ppm_p3.restype = C.c_size_t
        L.o_ppm_p3.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t]
        L.o_random_double.restype = C.c_double
        L.o_random_double.argtypes = [C.c_void_p]
        L.o_rng_init_mt.argtypes = [C.c_void_p, C.c_uint32]
        L.o_rng_init_philox.argtypes = [C.c_void_p, C.c_uint64]
        L.o_rng_set_stream.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        L.o_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ob_closest_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_void_p]
        L.ob_scatter.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ob_get_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p]
        L.ob_pixel_match.argtypes = [C.c_void_p]
        L.ob_write_color.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.o_demo_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
        L.o_demo_compute.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        _oracle = L
    return _oracle


def have_ref():
    build()
    return os.path.exists(LIB_REF)


def ref():
    global _ref
    if _ref is None:
        build()
        L = C.CDLL(LIB_REF)
        L.ref_scene_main.restype = C.c_size_t
        L.ref_scene_main.argtypes = [C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ref_main_ppm.restype = C.c_size_t
        L.ref_main_ppm.argtypes = [C.c_void_p, C.c_size_t]
        L.ref_world_create.restype = C.c_void_p
        L.ref_world_create.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        L.ref_world_destroy.argtypes = [C.c_void_p]
        L.ref_closest_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_double, C.c_int, C.c_void_p]
        L.ref_scatter.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_get_rays.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p]
        L.ref_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_pixel_match.argtypes = [C.c_void_p]
        _ref = L
    return _ref


# ------------------------------------------------------------------ scenes --

def scene_rtiow(g=11, p_diffuse=0.8, p_metal=0.95, mt_seed=5489, with_draws=False):
    """src/main.cpp:11-53 restated (oracle).  Returns (spheres, materials[, mt draws used])."""
    cap = 4 * g * g + 8
    s = np.zeros(cap, SPHERE)
    m = np.zeros(cap, MATERIAL)
    draws = C.c_uint64(0)
    n = oracle().o_scene_rtiow(g, p_diffuse, p_metal, mt_seed, _p(s), _p(m), cap, C.addressof(draws))
    if with_draws:
        return s[:n].copy(), m[:n].copy(), draws.value
    return s[:n].copy(), m[:n].copy()


def ref_scene_main(mt_seed=5489, with_draws=False):
    """The world the reference's own main() builds."""
    s = np.zeros(1024, SPHERE)
    m = np.zeros(1024, MATERIAL)
    draws = C.c_uint64(0)
    n = ref().ref_scene_main(mt_seed, _p(s), _p(m), 1024, C.addressof(draws))
    if with_draws:
        return s[:n].copy(), m[:n].copy(), draws.value
    return s[:n].copy(), m[:n].copy()


def camera_rtiow(image_width=400, spp=30, max_depth=50):
    c = np.zeros(1, CAMERA)
    oracle().o_camera_rtiow(_p(c), image_width, spp, max_depth)
    oracle().o_camera_initialize(_p(c))
    return c


def camera_initialize(c):
    oracle().o_camera_initialize(_p(c))
    return c


# ------------------------------------------------------------- batch calls --

def _opts(rng_mode, nthreads, seed, sample0, nsamples, row0=0, row1=0, mt_skip=0):
    o = np.zeros(1, RENDER_OPTS)
    o["mt_skip"] = mt_skip
    o["rng_mode"], o["nthreads"], o["seed"] = rng_mode, nthreads, seed
    o["sample0"], o["nsamples"], o["row0"], o["row1"] = sample0, nsamples, row0, row1
    return o


def render(spheres, materials, cam, rng_mode=RNG_PHILOX, seed=0, nthreads=None, use_bvh=False,
           sample0=0, nsamples=None, impl="oracle", row0=0, row1=0, mt_skip=0):
    """Full render.  Returns dict(sum=(H,W,3) f8, rgb8=(H,W,3) u8, rays=int)."""
    cam = cam.copy()
    camera_initialize(cam)
    W, H = int(cam["image_width"][0]), int(cam["image_height"][0])
    nsamples = int(cam["samples_per_pixel"][0]) if nsamples is None else nsamples
    nthreads = nthreads or os.cpu_count() or 1
    spheres = np.ascontiguousarray(spheres)
    materials = np.ascontiguousarray(materials)
    out_sum = np.zeros((H, W, 3), "f8")
    rgb8 = np.zeros((H, W, 3), "u1")
    rays = C.c_uint64(0)
    o = _opts(rng_mode, nthreads, seed, sample0, nsamples, row0, row1, mt_skip)
    if impl == "oracle":
        L = oracle()
        bvh = L.o_bvh_build(_p(spheres), len(spheres)) if use_bvh else None
        w = _World(spheres.ctypes.data, materials.ctypes.data, len(spheres), bvh)
        L.o_render(C.addressof(w), _p(cam), _p(o), _p(out_sum), _p(rgb8), C.addressof(rays))
        if bvh:
            L.o_bvh_free(bvh)
    else:
        L = ref()
        h = L.ref_world_create(_p(spheres), _p(materials), len(spheres), int(use_bvh))
        L.ref_render(h, _p(cam), _p(o), int(use_bvh), _p(out_sum), _p(rgb8), C.addressof(rays))
        L.ref_world_destroy(h)
    return dict(sum=out_sum, rgb8=rgb8, rays=rays.value, width=W, height=H)


def ppm_p3(rgb8):
    H, W, _ = rgb8.shape
    rgb8 = np.ascontiguousarray(rgb8)
    n = oracle().o_ppm_p3(_p(rgb8), W, H, None, 0)
    buf = C.create_string_buffer(n + 1)
    oracle().o_ppm_p3(_p(rgb8), W, H, buf, n + 1)
    return buf.raw[:n]


def ppm_p6(rgb8):
    """binary PPM of the write_color bytes"""
    h, w = rgb8.shape[:2]
    return b"P6\n%d %d\n255\n" % (w, h) + np.ascontiguousarray(rgb8, np.uint8).tobytes()


def pfm(sums, spp):
    """PFM of the mean radiance pixel_color * (1.0 / spp) (src/color.h:20-23, before gamma): float32 little
    endian, rows bottom to top"""
    h, w = sums.shape[:2]
    mean = (np.asarray(sums, np.float64) * (1.0 / spp)).astype("<f4")
    return b"PF\n%d %d\n-1.0\n" % (w, h) + np.ascontiguousarray(mean[::-1]).tobytes()


def closest_hit(spheres, rays, tmin=0.001, tmax=np.inf, impl="oracle", use_bvh=False):
    spheres = np.ascontiguousarray(spheres)
    rays = np.ascontiguousarray(rays)
    out = np.zeros(len(rays), HIT)
    if impl == "oracle":
        L = oracle()
        bvh = L.o_bvh_build(_p(spheres), len(spheres)) if use_bvh else None
        w = _World(spheres.ctypes.data, None, len(spheres), bvh)
        L.ob_closest_hit(C.addressof(w), _p(rays), len(rays), tmin, tmax, _p(out))
        if bvh:
            L.o_bvh_free(bvh)
    else:
        L = ref()
        mats = np.zeros(len(spheres), MATERIAL)
        sp = spheres.copy()
        sp["mat"] = np.arange(len(sp))
        h = L.ref_world_create(_p(sp), _p(mats), len(sp), int(use_bvh))
        L.ref_closest_hit(h, _p(rays), len(rays), tmin, tmax, int(use_bvh), _p(out))
        L.ref_world_destroy(h)
    return out


def scatter(materials, rays_in, hits, seed=0, impl="oracle"):
    """Per-element scatter; element i draws from stream (pixel=i, sample=0, stage=1)."""
    n = len(rays_in)
    materials = np.ascontiguousarray(materials)
    rays_in = np.ascontiguousarray(rays_in)
    hits = np.ascontiguousarray(hits)
    flag = np.zeros(n, "i4")
    att = np.zeros(n, VEC3)
    out = np.zeros(n, RAY)
    fn = oracle().ob_scatter if impl == "oracle" else ref().ref_scatter
    fn(_p(materials), _p(rays_in), _p(hits), n, seed, _p(flag), _p(att), _p(out))
    return flag, att, out


def get_rays(cam, px_i, px_j, sample, seed=0, impl="oracle"):
    cam = cam.copy()
    px_i = np.ascontiguousarray(px_i, "i4")
    px_j = np.ascontiguousarray(px_j, "i4")
    sample = np.ascontiguousarray(sample, "u4")
    out = np.zeros(len(px_i), RAY)
    fn = oracle().ob_get_rays if impl == "oracle" else ref().ref_get_rays
    fn(_p(cam), _p(px_i), _p(px_j), _p(sample), len(px_i), seed, _p(out))
    return out, cam


def pixel_match(impl="oracle"):
    out = np.zeros(3, "f8")
    (oracle().ob_pixel_match if impl == "oracle" else ref().ref_pixel_match)(_p(out))
    return out


def write_color(sums, spp):
    """(N,3) f8 per-pixel sums -> (N,3) u8, src/color.h:14-35."""
    sums = np.ascontiguousarray(sums, "f8").reshape(-1, 3)
    out = np.zeros(sums.shape, "u1")
    oracle().ob_write_color(_p(sums), len(sums), spp, _p(out))
    return out


def ref_main_ppm():
    """stdout of the reference's unmodified main() (shipped settings, ~35 s)."""
    n = 4 << 20
    buf = C.create_string_buffer(n)
    m = ref().ref_main_ppm(buf, n)
    return buf.raw[:m]


def philox(ctr, key):
    ctr = np.ascontiguousarray(ctr, "u4")
    key = np.ascontiguousarray(key, "u4")
    out = np.zeros(4, "u4")
    oracle().o_philox4x32_10(_p(ctr), _p(key), _p(out))
    return out


# ---------------------------------------------- 8f-3: the demo's compute shader --
_ref_glsl = None


def have_ref_glsl():
    build()
    return os.path.exists(LIB_REF_GLSL)


def ref_glsl():
    """assets/shaders/ray_tracing.comp compiled as C++ (oracle/glsl_harness.cpp)."""
    global _ref_glsl
    if _ref_glsl is None:
        build()
        L = C.CDLL(LIB_REF_GLSL)
        L.ref_demo_compute.argtypes = [C.c_int, C.c_int, C.c_void_p]
        L.ref_demo_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float]
        _ref_glsl = L
    return _ref_glsl


def demo_compute(width, height, spheres=None, impl="oracle"):
    """rgba8 image (height, width, 4) of the demo's compute stage; impl="ref" runs the shader itself
    (hard-coded sphere only)."""
    img = np.zeros((height, width, 4), np.uint8)
    if impl == "ref":
        assert spheres is None
        ref_glsl().ref_demo_compute(width, height, _p(img))
    else:
        s = None if spheres is None else np.ascontiguousarray(spheres, DEMO_SPHERE)
        if s is not None and len(s) == 0:
            s = np.zeros(1, DEMO_SPHERE)[:0]
        keep = np.zeros(1, DEMO_SPHERE) if s is not None and len(s) == 0 else s     # a non-NULL pointer for an empty list
        oracle().o_demo_compute(None if s is None else _p(keep), 0 if s is None else len(s), width, height, _p(img))
    return img


def demo_hit(origins, directions, centers, radii, impl="oracle"):
    """hit() of the shader (:65-73) on n rays x their spheres -> bool array"""
    o = np.ascontiguousarray(origins, "f4"); d = np.ascontiguousarray(directions, "f4"); c = np.ascontiguousarray(centers, "f4")
    fn = ref_glsl().ref_demo_hit if impl == "ref" else oracle().o_demo_hit
    return np.array([bool(fn(_p(o[k]), _p(d[k]), _p(c[k]), float(radii[k]))) for k in range(len(o))])
