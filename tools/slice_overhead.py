"""Cost of cutting a frame into jobs: render-kernel time of the C3 scene at 1920x1080 over spp x slices.
A job = one 16x8 tile x one sample slice; every job fills and drains its warp's path pool once, and the
frame ends when the last job ends.  python tools/slice_overhead.py [spp,spp,...] [slices,slices,...]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
spps = [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "8,16,31,62,125,250").split(",")]
sls = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "0,1,2,4,8,16").split(",")]
jpw = [int(x) for x in (sys.argv[3] if len(sys.argv) > 3 else "0").split(",")]      # auto layers: bulk jobs per warp (0 = default)
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
for spp in spps:
    cam = rt.camera_rtiow(1920, spp, 50)
    for sl, jp in [(a, b) for a in sls for b in (jpw if a == 0 else [0])]:
        if sl > spp:
            continue
        best = 1e30
        for _ in range(3):
            ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT, slices=sl, sched=jp << 24)
            ctx.synchronize(); st = ctx.stats(); best = min(best, st["render_ms"])
        print(f"spp {spp:4d} slices {st['slices']:2d} ({'auto' if sl == 0 else 'set'}{f' jobs/warp {jp}' if jp else ''})  ms {best:8.3f}  ms/spp {best / spp:7.4f}  Msamples/s {st['samples'] / best / 1e3:8.1f}", flush=True)
