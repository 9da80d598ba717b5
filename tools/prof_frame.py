"""Framebuffer passes alone, for ncu: python tools/prof_frame.py [width] [slices]   (RT_NO_FUSED_ENCODE=1: the three-kernel encoder)"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
width = int(sys.argv[1]) if len(sys.argv) > 1 else 3840
slices = int(sys.argv[2]) if len(sys.argv) > 2 else 1
ctx = rt.Context(0)
s, m = rt.scene_rtiow()
ctx.upload_scene(s, m)
spp = max(2, slices)
ctx.render(rt.camera_rtiow(width, spp, 4), precision=rt.F32, seed=0, slices=slices)
ctx.synchronize()
n = C.c_size_t(0)
for _ in range(2):
    ctx._ck(ctx._lib.rt_readback(ctx._h, spp, None, None, None, None))              # resolve alone
    ctx._ck(ctx._lib.rt_readback(ctx._h, spp, None, None, None, C.byref(n)))        # resolve + P3 text
    print("resolve_ms", ctx.stats()["resolve_ms"], "text bytes", n.value)
