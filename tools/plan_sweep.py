"""bulk jobs per warp of the layer plan (rt_render_opts.sched bits 24..31) on the C3 frame: python tools/plan_sweep.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
for spp in (500, 62, 250):
    cam = rt.camera_rtiow(1920, spp, 50)
    for rnd in range(2):
        for jobs in (20, 24, 28, 32, 40):
            ts = []
            for _ in range(3 if spp == 500 else 5):
                ctx.render(cam, precision=rt.F32, seed=0, sched=jobs << 24); ctx.synchronize(); st = ctx.stats(); ts.append(st["render_ms"])
            print(f"spp {spp} jobs/warp {jobs:3d}: best {min(ts):8.3f} ms  slices {st['slices']}", flush=True)
