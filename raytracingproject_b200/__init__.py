"""raytracingproject_b200 -- B200-native (sm_100a CUDA) path tracer behind the camera /
hittable_list / sphere / material API of ShuvalovIvan/RayTracingProject.  Only the reference's
CPU path-tracing hot path is in scope."""
from ._lib import (ABI, F32, F64, MEGAKERNEL, WAVEFRONT, MEGA_LOCKSTEP, TRACE_BVH, TRACE_BRUTE, BVH_SAH_HOST, BVH_LBVH_GPU,
                   LAMBERTIAN, METAL, DIELECTRIC, SPHERE, MATERIAL, DEMO_SPHERE, IMG_P3, IMG_P6, IMG_PFM, RAY, HIT, LIB_PATH, SYMBOLS, lib)
from .api import (Camera, Context, Dielectric, HittableList, Lambertian, Metal, RtError, Sphere,
                  build_bvh_host, camera_rtiow, plan_layers, scene_rtiow)

__all__ = [n for n in dir() if not n.startswith("_")]
