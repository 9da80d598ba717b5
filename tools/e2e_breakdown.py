import sys, os, time
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import raytracingproject_b200 as rt
s, m = rt.scene_rtiow()
ctx = rt.Context(0)
cam = rt.camera_rtiow(1920, 500, 50)
W, H = 1920, 1080
rgb = torch.empty(H * W * 3, dtype=torch.uint8).pin_memory().numpy()
ppm = torch.empty(H * W * 12 + 64, dtype=torch.uint8).pin_memory().numpy()
for it in range(4):
    t0 = time.perf_counter(); ctx.upload_scene(s, m); t1 = time.perf_counter()
    ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT); t2 = time.perf_counter()
    ctx.synchronize(); t3 = time.perf_counter()
    n = ctx.readback_into(500, rgb, ppm); t4 = time.perf_counter()
    st = ctx.stats()
    print(f"upload {1e3*(t1-t0):.2f} ms (build {st['build_ms']:.2f}, upload_ms {st['upload_ms']:.2f})  launch {1e3*(t2-t1):.2f}  render wait {1e3*(t3-t2):.2f} (kernel {st['render_ms']:.2f})  readback {1e3*(t4-t3):.2f} (resolve+encode {st['resolve_ms']:.3f})  total {1e3*(t4-t0):.2f}  ppm bytes {n}")
