"""Where the end-to-end step goes: python tools/e2e_breakdown.py [g] [width] [spp]   (g = 11: C3; 50 / 159 / 500: C5)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import raytracingproject_b200 as rt
g = int(sys.argv[1]) if len(sys.argv) > 1 else 11
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 500
H = W * 9 // 16
builder = rt.BVH_SAH_HOST if g == 11 else rt.BVH_LBVH_GPU
s, m = rt.scene_rtiow(g)
ctx = rt.Context(0)
cam = rt.camera_rtiow(W, spp, 50, scale=g / 11.0)
rgb = torch.empty(H * W * 3, dtype=torch.uint8).pin_memory().numpy()
ppm = torch.empty(H * W * 12 + 64, dtype=torch.uint8).pin_memory().numpy()
for it in range(4):
    t0 = time.perf_counter(); ctx.upload_scene(s, m, builder=builder); t1 = time.perf_counter()
    ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT); t2 = time.perf_counter()
    ctx.synchronize(); t3 = time.perf_counter()
    n = ctx.readback_into(spp, rgb, ppm); t4 = time.perf_counter()
    st = ctx.stats()
    print(f"upload {1e3*(t1-t0):.2f} ms (build {st['build_ms']:.2f}, upload_ms {st['upload_ms']:.2f})  launch {1e3*(t2-t1):.2f}  render wait {1e3*(t3-t2):.2f} (kernel {st['render_ms']:.2f})  readback {1e3*(t4-t3):.2f} (resolve+encode {st['resolve_ms']:.3f})  total {1e3*(t4-t0):.2f}  ppm bytes {n}")
