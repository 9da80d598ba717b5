import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
for spp, stride in ((125, 1), (250, 2), (125, 1), (250, 2), (500, 4), (250,1)):
    os.environ["RT_JOB_STRIDE"] = str(stride)
    cam = rt.camera_rtiow(1920, spp, 50)
    ts = []
    for _ in range(4):
        ctx.render(cam, precision=rt.F32, seed=0); ctx.synchronize(); st = ctx.stats(); ts.append(st["render_ms"])
    print(f"spp {spp} stride {stride}: best {min(ts):8.3f} ms rays {st['rays']/1e6:.1f} M  -> {st['rays']/1e6/min(ts):.2f} Mrays/ms  slices {st['slices']}", flush=True)
