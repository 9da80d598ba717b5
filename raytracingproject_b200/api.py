"""Host-side mirror of the reference's camera / hittable_list / sphere / material interface
(reference src/camera.h, src/hittable_list.h, src/sphere.h, src/material.h) over the C ABI of
librt_b200.so.  `cam.render(world)` keeps the reference's meaning -- P3 text on stdout
(src/camera.h:32-50) -- and runs on the GPU.  Python is only the test/bench driver; the C++
drop-in headers are under include/rtp/.
"""
import ctypes as C
import sys

import numpy as np

from . import _lib as L


class RtError(RuntimeError):
    pass


# ------------------------------------------------------------------- materials / spheres --
class Lambertian:                      # src/material.h:15-29
    type = L.LAMBERTIAN

    def __init__(self, albedo):
        self.albedo = tuple(float(x) for x in albedo)
        self.param = 0.0


class Metal:                           # src/material.h:31-46 (fuzz clamped to <= 1, :33)
    type = L.METAL

    def __init__(self, albedo, fuzz):
        self.albedo = tuple(float(x) for x in albedo)
        self.param = float(fuzz) if fuzz < 1 else 1.0


class Dielectric:                      # src/material.h:48-82
    type = L.DIELECTRIC

    def __init__(self, index_of_refraction):
        self.albedo = (0.0, 0.0, 0.0)
        self.param = float(index_of_refraction)


class Sphere:
    """sphere(center, radius, mat) or sphere(center1, center2, radius, mat) -- src/sphere.h:9-28."""

    def __init__(self, *args):
        if len(args) == 3:
            c1, radius, mat = args
            c2, self.is_moving = c1, False
        elif len(args) == 4:
            c1, c2, radius, mat = args
            self.is_moving = True
        else:
            raise TypeError("Sphere(center, radius, material) or Sphere(center1, center2, radius, material)")
        self.center1 = tuple(float(x) for x in c1)
        # center_vec = center2 - center1 (src/sphere.h:27)
        self.center_vec = tuple(float(b) - float(a) for a, b in zip(c1, c2)) if self.is_moving else (0.0, 0.0, 0.0)
        self.radius = float(radius)
        self.mat = mat


class HittableList:                    # src/hittable_list.h:11-45
    def __init__(self, obj=None):
        self.objects = []
        if obj is not None:
            self.add(obj)

    def clear(self):
        self.objects.clear()

    def add(self, obj):
        self.objects.append(obj)

    def to_arrays(self):
        """-> (rt_sphere[n], rt_material[n]) structured arrays, one material per sphere."""
        n = len(self.objects)
        s = np.zeros(n, L.SPHERE)
        m = np.zeros(n, L.MATERIAL)
        for k, o in enumerate(self.objects):
            if not isinstance(o, Sphere):
                raise RtError("B200 path renders hittable_lists of spheres only")
            s[k]["center1"] = o.center1
            s[k]["center_vec"] = o.center_vec
            s[k]["radius"] = o.radius
            s[k]["is_moving"] = int(o.is_moving)
            s[k]["material"] = k
            m[k]["type"] = o.mat.type
            m[k]["albedo"] = o.mat.albedo
            m[k]["param"] = o.mat.param
        return s, m


# --------------------------------------------------------------------------- context --
class Context:
    """One device context (rt_ctx).  Not re-entrant, like the reference."""

    def __init__(self, device=0):
        self._lib = L.lib()
        h = C.c_void_p()
        rc = self._lib.rt_create(C.byref(h), device)
        if rc != 0:
            raise RtError(self._lib.rt_last_error(None).decode())
        self._h = h
        self.width = self.height = 0

    def close(self):
        if getattr(self, "_h", None):
            self._lib.rt_destroy(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise RtError(self._lib.rt_last_error(self._h).decode())

    def upload_scene(self, spheres, materials, builder=L.BVH_SAH_HOST, max_leaf=0, big_radius=0.0, wide=0):
        spheres = np.ascontiguousarray(spheres)
        materials = np.ascontiguousarray(materials)
        assert spheres.dtype.itemsize == 64 and materials.dtype.itemsize == 40
        o = L.BvhOpts(L.ABI, builder, max_leaf, wide, big_radius)
        self._ck(self._lib.rt_upload_scene(self._h, L.ptr(spheres), len(spheres), L.ptr(materials), len(materials), C.byref(o)))

    def update_spheres(self, spheres, materials):
        """rt_update_spheres: the uploaded world with its spheres moved (same count and order); the tree is refitted on the device."""
        spheres = np.ascontiguousarray(spheres)
        materials = np.ascontiguousarray(materials)
        assert spheres.dtype.itemsize == 64 and materials.dtype.itemsize == 40
        self._ck(self._lib.rt_update_spheres(self._h, L.ptr(spheres), len(spheres), L.ptr(materials), len(materials)))

    def render(self, cam, precision=L.F32, seed=0, variant=L.WAVEFRONT, sample_begin=0, sample_count=0,
               accumulate=False, slices=0, sched=0, tile_rank=0, tile_world=0, top_cap=0, shared_jobs=False):
        o = L.RenderOpts(L.ABI, precision, variant, sample_begin, sample_count, int(accumulate), seed, slices, sched,
                         tile_rank, tile_world, top_cap, int(shared_jobs))
        self._ck(self._lib.rt_render(self._h, C.byref(cam), C.byref(o)))
        w, h = C.c_int(), C.c_int()
        self._lib.rt_image_size(C.byref(cam), C.byref(w), C.byref(h))
        self.width, self.height = w.value, h.value

    def readback(self, spp_total=0, want_sum=True, want_rgb8=True, want_ppm=False):
        H, W = self.height, self.width
        s = np.zeros((H, W, 3), "f8") if want_sum else None
        b = np.zeros((H, W, 3), "u1") if want_rgb8 else None
        text = None
        if want_ppm:
            cap = C.c_size_t(H * W * 12 + 64)
            buf = C.create_string_buffer(cap.value)
            self._ck(self._lib.rt_readback(self._h, spp_total, L.ptr(s), L.ptr(b), buf, C.byref(cap)))
            text = buf.raw[:cap.value]
        else:
            self._ck(self._lib.rt_readback(self._h, spp_total, L.ptr(s), L.ptr(b), None, None))
        return dict(sum=s, rgb8=b, ppm=text)

    def readback_into(self, spp_total, rgb8=None, ppm=None):
        """rt_readback into caller-owned buffers (numpy uint8 arrays, e.g. over pinned host memory): no
        allocation or extra copy per call.  Returns the length of the P3 text written to `ppm` (0 if None)."""
        n = C.c_size_t(ppm.nbytes if ppm is not None else 0)
        self._ck(self._lib.rt_readback(self._h, spp_total, None, L.ptr(rgb8), L.ptr(ppm), C.byref(n) if ppm is not None else None))
        return n.value

    def readback_begin(self, spp_total, rgb8=None, ppm=None):
        """rt_readback_begin: resolve (+ P3 text) on the device, then start the copies into the caller's buffers (numpy uint8
        arrays, pinned for a truly asynchronous copy) and return; render the next frame, then readback_end()."""
        self._ck(self._lib.rt_readback_begin(self._h, spp_total, L.ptr(rgb8), L.ptr(ppm), ppm.nbytes if ppm is not None else 0))

    def readback_end(self):
        """rt_readback_end: wait for the copies begun by readback_begin; returns the length of the P3 text (0 if none)."""
        n = C.c_size_t(0)
        self._ck(self._lib.rt_readback_end(self._h, C.byref(n)))
        return n.value

    def readback_image(self, fmt, spp_total=0):
        """The frame as file bytes: IMG_P3 (the reference's stdout), IMG_P6 or IMG_PFM (rt_readback_image)."""
        n = C.c_size_t(0)
        self._ck(self._lib.rt_readback_image(self._h, spp_total, fmt, None, C.byref(n)))
        buf = np.empty(n.value, np.uint8)
        self._ck(self._lib.rt_readback_image(self._h, spp_total, fmt, L.ptr(buf), C.byref(n)))
        return buf[:n.value].tobytes()

    def preview(self, spp_total=0, copy=True):
        """rt_preview: (device pointer, (H, W, 4) rgba8 host copy or None)."""
        p = C.c_void_p()
        img = np.empty((self.height, self.width, 4), np.uint8) if copy else None
        self._ck(self._lib.rt_preview(self._h, spp_total, C.byref(p), L.ptr(img)))
        return p.value, img

    def demo_compute(self, width=1024, height=768, spheres=None):
        """The windowed demo's compute stage (reference assets/shaders/ray_tracing.comp:35-73; default size
        src/common_objects.h:8-9) into a headless (height, width, 4) rgba8 image.  spheres: DEMO_SPHERE
        array, or None for the sphere the shader hard-codes."""
        img = np.zeros((height, width, 4), np.uint8)
        s = None if spheres is None else np.ascontiguousarray(spheres, L.DEMO_SPHERE)
        self._ck(self._lib.rt_demo_compute(self._h, L.ptr(s), 0 if s is None else len(s), width, height, L.ptr(img)))
        return img

    def synchronize(self):
        self._ck(self._lib.rt_synchronize(self._h))

    def stats(self):
        st = L.Stats()
        self._ck(self._lib.rt_get_stats(self._h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in st._fields_ if k != "pad"}

    def debug_counters(self):
        out = np.zeros(32, np.uint64)
        self._ck(self._lib.rt_debug_counters(self._h, L.ptr(out)))
        return out

    def debug_warp_times(self, n_warps=148 * 32):
        out = np.zeros((n_warps, 8), np.uint64)
        self._ck(self._lib.rt_debug_warp_times(self._h, L.ptr(out), out.size))
        return out

    def accum_buffer(self):
        p, n, e = C.c_void_p(), C.c_size_t(), C.c_int()
        self._ck(self._lib.rt_accum_buffer(self._h, C.byref(p), C.byref(n), C.byref(e)))
        return p.value, n.value, e.value

    def fold_slices(self):
        """Sum the sample layers into layer 0 (asynchronous); the combine then moves one layer per rank."""
        self._ck(self._lib.rt_fold_slices(self._h))

    def stream(self):
        p = C.c_void_p()
        self._ck(self._lib.rt_stream(self._h, C.byref(p)))
        return p.value or 0

    def ipc_export(self):
        buf = (C.c_ubyte * L.IPC_HANDLE_BYTES)()
        self._ck(self._lib.rt_ipc_export(self._h, buf))
        return bytes(buf)

    def ipc_open_peers(self, handles):
        blob = b"".join(handles)
        self._ck(self._lib.rt_ipc_open_peers(self._h, blob, len(handles)))

    def ipc_close_peers(self):
        self._ck(self._lib.rt_ipc_close_peers(self._h))

    # ---- tile-sharded frames: bytes written into the root's image over NVLink (rt_b200.h) ----
    # dynamic sharding: one job counter per frame for all GPUs (rt_jobs_*)
    def jobs_export(self):
        buf = (C.c_ubyte * L.IPC_HANDLE_BYTES)()
        self._ck(self._lib.rt_jobs_export(self._h, buf))
        return bytes(buf)

    def jobs_open(self, blob):
        self._ck(self._lib.rt_jobs_open(self._h, blob))

    def jobs_share(self, root):
        self._ck(self._lib.rt_jobs_share(self._h, root._h))

    def jobs_close(self):
        self._ck(self._lib.rt_jobs_close(self._h))

    def frame_export(self):
        buf = (C.c_ubyte * L.IPC_HANDLE_BYTES)()
        self._ck(self._lib.rt_frame_export(self._h, buf))
        return bytes(buf)

    def frame_open(self, handle):
        self._ck(self._lib.rt_frame_open(self._h, handle))

    def frame_close(self):
        self._ck(self._lib.rt_frame_close(self._h))

    def resolve_tiles(self, spp_total=0):
        self._ck(self._lib.rt_resolve_tiles(self._h, spp_total))

    # ---- parity hooks ----
    def trace_closest(self, rays, tmin=0.001, tmax=np.inf, precision=L.F64, backend=L.TRACE_BVH):
        rays = np.ascontiguousarray(rays)
        assert rays.dtype.itemsize == 56
        out = np.zeros(len(rays), L.HIT)
        self._ck(self._lib.rt_trace_closest(self._h, L.ptr(rays), len(rays), tmin, tmax, precision, backend, L.ptr(out)))
        return out

    def scatter(self, materials, rays_in, hits, seed=0, precision=L.F64):
        n = len(rays_in)
        materials = np.ascontiguousarray(materials)
        rays_in = np.ascontiguousarray(rays_in)
        hits = np.ascontiguousarray(hits)
        assert materials.dtype.itemsize == 40 and rays_in.dtype.itemsize == 56 and hits.dtype.itemsize == 64
        flag = np.zeros(n, "i4")
        att = np.zeros((n, 3), "f8")
        out = np.zeros(n, L.RAY)
        self._ck(self._lib.rt_scatter(self._h, L.ptr(materials), L.ptr(rays_in), L.ptr(hits), n, seed, precision,
                                      L.ptr(flag), L.ptr(att), L.ptr(out)))
        return flag, att, out

    def gen_rays(self, cam, px_i, px_j, sample, seed=0, precision=L.F64):
        px_i = np.ascontiguousarray(px_i, "i4")
        px_j = np.ascontiguousarray(px_j, "i4")
        sample = np.ascontiguousarray(sample, "u4")
        out = np.zeros(len(px_i), L.RAY)
        self._ck(self._lib.rt_gen_rays(self._h, C.byref(cam), L.ptr(px_i), L.ptr(px_j), L.ptr(sample), len(px_i), seed,
                                       precision, L.ptr(out)))
        return out

    def bench_closest(self, cam, n, seed=1, precision=L.F32, backend=L.TRACE_BVH, repeats=1, want_results=False, dump_rays=0):
        ms = C.c_double()
        prim = np.zeros(n, "i4") if want_results else None
        t = np.zeros(n, "f4") if want_results else None
        rays = np.zeros(dump_rays, L.RAY) if dump_rays else None
        self._ck(self._lib.rt_bench_closest(self._h, C.byref(cam), n, seed, precision, backend, repeats, C.byref(ms),
                                            L.ptr(prim), L.ptr(t), L.ptr(rays), dump_rays))
        return dict(ms=ms.value, prim=prim, t=t, rays=rays)


    def write_color(self, sums, spp):
        sums = np.ascontiguousarray(sums, "f8").reshape(-1, 3)
        out = np.zeros(sums.shape, "u1")
        self._ck(self._lib.rt_write_color(self._h, L.ptr(sums), len(sums), spp, L.ptr(out)))
        return out

    def encode_p3(self, rgb8):
        rgb8 = np.ascontiguousarray(rgb8, "u1")
        H, W, _ = rgb8.shape
        cap = C.c_size_t(H * W * 12 + 64)
        buf = C.create_string_buffer(cap.value)
        self._ck(self._lib.rt_encode_p3(self._h, L.ptr(rgb8), W, H, buf, C.byref(cap)))
        return buf.raw[:cap.value]

    def encode_p3_static(self, dev_rgb8, width, height, dev_ppm, ppm_capacity, dev_ppm_len):
        """rt_encode_p3_static: the allocation-free encoder.  Every buffer is a DEVICE address (int) of the caller's, on this
        context's device; asynchronous on the context's stream (synchronize() before reading the results)."""
        self._ck(self._lib.rt_encode_p3_static(self._h, C.c_void_p(dev_rgb8), width, height, C.c_void_p(dev_ppm),
                                               C.c_size_t(ppm_capacity), C.c_void_p(dev_ppm_len)))


def plan_layers(sample_count, tiles, resident_warps, jobs_per_warp=0):
    """Host logic alone (no device): the sample layers of a wavefront frame in auto mode -> (sizes, n_bulk)."""
    sizes = (C.c_int * 64)()
    n_bulk = C.c_int()
    n = L.lib().rt_plan_layers(int(sample_count), int(tiles), int(resident_warps), int(jobs_per_warp), sizes, C.byref(n_bulk))
    return list(sizes[:n]), n_bulk.value


def build_bvh_host(spheres, max_leaf=4):
    """Host SAH builder alone (no device).  -> dict(nodes (n_inner,16) f4, order, root_ref, depth)."""
    spheres = np.ascontiguousarray(spheres)
    n = len(spheres)
    nodes = np.zeros((max(n, 1), 16), "f4")
    order = np.zeros(max(n, 1), "i4")
    root, depth, n_inner = C.c_int32(), C.c_int32(), C.c_int32()
    rc = L.lib().rt_build_bvh_host(L.ptr(spheres), n, max_leaf, L.ptr(nodes), L.ptr(order), C.byref(root), C.byref(depth), C.byref(n_inner))
    if rc != 0:
        raise RtError("rt_build_bvh_host failed")
    return dict(nodes=nodes[:n_inner.value], order=order[:n], root_ref=root.value, depth=depth.value)


# ----------------------------------------------------------------------------- scenes --
def scene_rtiow(g=11, p_diffuse=0.8, p_metal=0.95, mt_seed=5489):
    """The world of reference src/main.cpp:11-53 (g, thresholds generalised for configs 4/5)."""
    lib = L.lib()
    n = lib.rt_scene_rtiow(g, p_diffuse, p_metal, mt_seed, None, None, 0)
    s = np.zeros(n, L.SPHERE)
    m = np.zeros(n, L.MATERIAL)
    lib.rt_scene_rtiow(g, p_diffuse, p_metal, mt_seed, L.ptr(s), L.ptr(m), n)
    return s, m


def camera_rtiow(image_width=400, spp=30, max_depth=50, scale=1.0):
    """The camera of reference src/main.cpp:55-68."""
    cam = L.Camera()
    L.lib().rt_camera_rtiow(C.byref(cam), image_width, spp, max_depth, scale)
    return cam


# ------------------------------------------------------------------------------ camera --
class Camera:
    """Drop-in for CPUImpl::Camera (reference src/camera_cpu.h:4-28): same public fields
    (src/camera.h:15-26), same render(world) contract (P3 text to a stream, src/camera.h:32-50)."""

    def __init__(self, device=0, precision=L.F64, seed=0):
        self.aspect_ratio = 1.0
        self.image_width = 100
        self.samples_per_pixel = 10
        self.max_depth = 10
        self.vfov = 90.0
        self.lookfrom = (0.0, 0.0, -1.0)
        self.lookat = (0.0, 0.0, 0.0)
        self.vup = (0.0, 1.0, 0.0)
        self.defocus_angle = 0.0
        self.focus_dist = 10.0
        self.device, self.precision, self.seed = device, precision, seed
        self._ctx = None
        self.last_stats = None

    def _pod(self):
        c = L.Camera()
        c.aspect_ratio, c.image_width = self.aspect_ratio, self.image_width
        c.samples_per_pixel, c.max_depth, c.vfov = self.samples_per_pixel, self.max_depth, self.vfov
        c.lookfrom[:] = self.lookfrom
        c.lookat[:] = self.lookat
        c.vup[:] = self.vup
        c.defocus_angle, c.focus_dist = self.defocus_angle, self.focus_dist
        return c

    def image_size(self):              # src/camera.h:28-30
        w, h = C.c_int(), C.c_int()
        c = self._pod()
        if L.lib().rt_image_size(C.byref(c), C.byref(w), C.byref(h)) != 0:
            raise RtError("bad image size")
        return w.value, h.value

    def context(self):
        if self._ctx is None:
            self._ctx = Context(self.device)
        return self._ctx

    def render_image(self, world):
        """Render and return dict(sum, rgb8, ppm) without printing."""
        ctx = self.context()
        s, m = world.to_arrays() if isinstance(world, HittableList) else world
        ctx.upload_scene(s, m)
        ctx.render(self._pod(), precision=self.precision, seed=self.seed)
        out = ctx.readback(self.samples_per_pixel, want_ppm=True)
        self.last_stats = ctx.stats()
        return out

    def render(self, world, out=None):  # src/camera.h:32-50
        img = self.render_image(world)
        stream = out if out is not None else sys.stdout
        data = img["ppm"]
        if hasattr(stream, "buffer"):
            stream.buffer.write(data)
        else:
            try:
                stream.write(data)
            except TypeError:
                stream.write(data.decode("ascii"))
        return img
