"""Small invocation of every kernel family for compute-sanitizer (memcheck): python tools/sanitize_case.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
s, m = rt.scene_rtiow()
ctx = rt.Context(0)
for builder in (rt.BVH_SAH_HOST, rt.BVH_LBVH_GPU):
    ctx.upload_scene(s, m, builder=builder)
    cam = rt.camera_rtiow(97, 3, 12)            # ragged tiles
    for prec in (rt.F32, rt.F64):
        for v in (rt.WAVEFRONT, rt.MEGAKERNEL, rt.MEGA_LOCKSTEP):
            ctx.render(cam, precision=prec, seed=3, variant=v)
            out = ctx.readback(3, want_ppm=True)
    r = ctx.bench_closest(cam, 1 << 14, seed=1, precision=rt.F32, want_results=True, dump_rays=256)
    h = ctx.trace_closest(r["rays"].view(rt.RAY), precision=rt.F32)
# a scene too large for shared memory (unstaged path)
s5, m5 = rt.scene_rtiow(40)
ctx.upload_scene(s5, m5, builder=rt.BVH_LBVH_GPU)
for v in (rt.WAVEFRONT, rt.MEGAKERNEL):
    ctx.render(rt.camera_rtiow(64, 2, 8, scale=40 / 11), precision=rt.F32, seed=0, variant=v)
    ctx.readback(2)
print("sanitize case done", ctx.stats()["rays"])
