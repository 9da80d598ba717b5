"""dynamic sharding on one GPU: the cost of the mode itself (shared counter in this GPU's memory): python tools/jobs_1gpu.py [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 125
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
cam = rt.camera_rtiow(1920, spp, 50)
ctx.render(cam, precision=rt.F32, seed=0); ctx.synchronize()
ctx.jobs_export()
for name, kw in (("own counter", {}), ("shared counter", {"shared_jobs": True}), ("own counter", {}), ("shared counter", {"shared_jobs": True})):
    ts = []
    for _ in range(4):
        ctx.render(cam, precision=rt.F32, seed=0, **kw); ctx.synchronize(); st = ctx.stats(); ts.append(st["render_ms"])
    print(f"{name:15s} spp {spp} best {min(ts):8.3f} ms  all {[round(t,2) for t in ts]} slices {st['slices']} rays {st['rays']}", flush=True)
