"""One C3-shaped render for profiling: python tools/prof_render.py [spp] [f32|f64] [repeats] [variant] [sched]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 32
prec = rt.F64 if len(sys.argv) > 2 and sys.argv[2] == "f64" else rt.F32
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
variant = int(sys.argv[4]) if len(sys.argv) > 4 else rt.MEGAKERNEL
sched = int(sys.argv[5], 0) if len(sys.argv) > 5 else 0
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
cam = rt.camera_rtiow(1920, spp, 50)
for _ in range(reps):
    ctx.render(cam, precision=prec, seed=0, variant=variant, sched=sched)
    ctx.readback(spp, want_sum=False, want_rgb8=False)
    st = ctx.stats()
    print(f"spp {spp} variant {variant} sched {sched:#x} render_ms {st['render_ms']:.3f} Msamples/s {st['samples']/st['render_ms']/1e3:.1f} Mrays/s {st['rays']/st['render_ms']/1e3:.1f} resolve_ms {st['resolve_ms']:.3f}")
