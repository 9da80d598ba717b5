"""ctypes binding of librt_b200.so (include/rt_b200.h).  No fallbacks: a missing library or a
missing CUDA device is an error, never a silent CPU path."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(HERE, "librt_b200.so")   # RT_B200_LIB: an A/B or STATS build of the same library

ABI = 2
F32, F64 = 0, 1
MEGAKERNEL, WAVEFRONT, MEGA_LOCKSTEP = 0, 1, 2
TRACE_BVH, TRACE_BRUTE = 0, 1
BVH_SAH_HOST, BVH_LBVH_GPU = 0, 1
LAMBERTIAN, METAL, DIELECTRIC = 0, 1, 2

# numpy mirrors of the POD structs (layouts asserted in rt_api.cu)
SPHERE = np.dtype([("center1", "f8", 3), ("center_vec", "f8", 3), ("radius", "f8"),
                   ("is_moving", "i4"), ("material", "i4")])
IMG_P3, IMG_P6, IMG_PFM = 0, 1, 2
IPC_HANDLE_BYTES = 80
DEMO_SPHERE = np.dtype([("center", "f4", 3), ("radius", "f4"), ("color", "f4", 4)])    # rt_demo_sphere
MATERIAL = np.dtype([("type", "i4"), ("pad", "i4"), ("albedo", "f8", 3), ("param", "f8")])
RAY = np.dtype([("orig", "f8", 3), ("dir", "f8", 3), ("tm", "f8")])
HIT = np.dtype([("p", "f8", 3), ("normal", "f8", 3), ("t", "f8"), ("front_face", "i4"), ("prim", "i4")])


class Camera(C.Structure):
    """rt_camera: the public fields of `camera` (reference src/camera.h:15-26)."""
    _fields_ = [("aspect_ratio", C.c_double), ("image_width", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_depth", C.c_int32), ("pad", C.c_int32), ("vfov", C.c_double),
                ("lookfrom", C.c_double * 3), ("lookat", C.c_double * 3), ("vup", C.c_double * 3),
                ("defocus_angle", C.c_double), ("focus_dist", C.c_double)]


class BvhOpts(C.Structure):
    _fields_ = [("abi", C.c_uint32), ("builder", C.c_int32), ("max_leaf", C.c_int32), ("wide", C.c_int32),
                ("big_radius", C.c_double)]


class RenderOpts(C.Structure):
    _fields_ = [("abi", C.c_uint32), ("precision", C.c_int32), ("variant", C.c_int32),
                ("sample_begin", C.c_int32), ("sample_count", C.c_int32), ("accumulate", C.c_int32),
                ("seed", C.c_uint64), ("slices", C.c_int32), ("sched", C.c_int32),
                ("tile_rank", C.c_int32), ("tile_world", C.c_int32), ("top_cap", C.c_int32), ("shared_jobs", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("samples", C.c_uint64), ("render_ms", C.c_double),
                ("resolve_ms", C.c_double), ("build_ms", C.c_double), ("upload_ms", C.c_double),
                ("n_spheres", C.c_int32), ("n_big", C.c_int32), ("n_nodes", C.c_int32), ("bvh_depth", C.c_int32),
                ("image_width", C.c_int32), ("image_height", C.c_int32),
                ("slices", C.c_int32), ("kernel_launches", C.c_int32),
                ("scene_in_smem", C.c_int32), ("big_in_tree", C.c_int32),
                ("top_nodes_in_smem", C.c_int32), ("pad", C.c_int32)]


# every symbol include/rt_b200.h declares: name -> (restype, argtypes)
_VP, _SZ, _I, _D, _U64 = C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.c_uint64
SYMBOLS = {
    "rt_create": (_I, [C.POINTER(_VP), _I]),
    "rt_destroy": (_I, [_VP]),
    "rt_last_error": (C.c_char_p, [_VP]),
    "rt_abi_version": (C.c_uint32, []),
    "rt_upload_scene": (_I, [_VP, _VP, _SZ, _VP, _SZ, _VP]),
    "rt_update_spheres": (_I, [_VP, _VP, _SZ, _VP, _SZ]),
    "rt_render": (_I, [_VP, _VP, _VP]),
    "rt_readback": (_I, [_VP, _I, _VP, _VP, _VP, _VP]),
    "rt_readback_begin": (_I, [_VP, _I, _VP, _VP, _SZ]),
    "rt_readback_end": (_I, [_VP, _VP]),
    "rt_readback_image": (_I, [_VP, _I, _I, _VP, _VP]),
    "rt_preview": (_I, [_VP, _I, _VP, _VP]),
    "rt_image_size": (_I, [_VP, _VP, _VP]),
    "rt_synchronize": (_I, [_VP]),
    "rt_debug_counters": (_I, [_VP, _VP]),
    "rt_debug_warp_times": (_I, [_VP, _VP, _SZ]),
    "rt_get_stats": (_I, [_VP, _VP]),
    "rt_accum_buffer": (_I, [_VP, _VP, _VP, _VP]),
    "rt_stream": (_I, [_VP, _VP]),
    "rt_fold_slices": (_I, [_VP]),
    "rt_plan_layers": (_I, [_I, C.c_longlong, C.c_longlong, _I, _VP, _VP]),
    "rt_ipc_export": (_I, [_VP, _VP]),
    "rt_ipc_open_peers": (_I, [_VP, _VP, _I]),
    "rt_ipc_close_peers": (_I, [_VP]),
    "rt_jobs_export": (_I, [_VP, _VP]),
    "rt_jobs_open": (_I, [_VP, _VP]),
    "rt_jobs_share": (_I, [_VP, _VP]),
    "rt_jobs_close": (_I, [_VP]),
    "rt_frame_export": (_I, [_VP, _VP]),
    "rt_frame_open": (_I, [_VP, _VP]),
    "rt_frame_close": (_I, [_VP]),
    "rt_resolve_tiles": (_I, [_VP, _I]),
    "rt_trace_closest": (_I, [_VP, _VP, _SZ, _D, _D, _I, _I, _VP]),
    "rt_scatter": (_I, [_VP, _VP, _VP, _VP, _SZ, _U64, _I, _VP, _VP, _VP]),
    "rt_gen_rays": (_I, [_VP, _VP, _VP, _VP, _VP, _SZ, _U64, _I, _VP]),
    "rt_bench_closest": (_I, [_VP, _VP, _SZ, _U64, _I, _I, _I, _VP, _VP, _VP, _VP, _SZ]),
    "rt_write_color": (_I, [_VP, _VP, _SZ, _I, _VP]),
    "rt_encode_p3": (_I, [_VP, _VP, _I, _I, _VP, _VP]),
    "rt_encode_p3_static": (_I, [_VP, _VP, _I, _I, _VP, _SZ, _VP]),
    "rt_demo_compute": (_I, [_VP, _VP, _SZ, _I, _I, _VP]),
    "rt_build_bvh_host": (_I, [_VP, _SZ, _I, _VP, _VP, _VP, _VP, _VP]),
    "rt_scene_rtiow": (_SZ, [_I, _D, _D, C.c_uint32, _VP, _VP, _SZ]),
    "rt_camera_rtiow": (_I, [_VP, _I, _I, _I, _D]),
}

_lib = None


def lib():
    """Load librt_b200.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(make -C raytracingproject_b200/csrc).  There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)          # AttributeError if the library does not export it
            fn.restype = res
            fn.argtypes = args
        if L.rt_abi_version() != ABI:
            raise RuntimeError("librt_b200.so ABI version mismatch")
        _lib = L
    return _lib


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)
