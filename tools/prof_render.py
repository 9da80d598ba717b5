"""One render for profiling: python tools/prof_render.py [spp] [f32|f64] [repeats] [variant] [sched] [g] [width] [max_leaf]
(g = half-width of the scene generator: 11 = the final scene / C3, 50 / 159 / 500 = C5's 10 k / 100 k / 1 M spheres, GPU LBVH)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 32
prec = rt.F64 if len(sys.argv) > 2 and sys.argv[2] == "f64" else rt.F32
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
variant = int(sys.argv[4]) if len(sys.argv) > 4 else rt.MEGAKERNEL
sched = int(sys.argv[5], 0) if len(sys.argv) > 5 else 0
g = int(sys.argv[6]) if len(sys.argv) > 6 else 11
width = int(sys.argv[7]) if len(sys.argv) > 7 else 1920
s, m = rt.scene_rtiow(g)
max_leaf = int(sys.argv[8]) if len(sys.argv) > 8 else 0
ctx = rt.Context(0); ctx.upload_scene(s, m, builder=rt.BVH_SAH_HOST if g == 11 else rt.BVH_LBVH_GPU, max_leaf=max_leaf)
cam = rt.camera_rtiow(width, spp, 50, scale=g / 11.0)
for _ in range(reps):
    ctx.render(cam, precision=prec, seed=0, variant=variant, sched=sched)
    ctx.readback(spp, want_sum=False, want_rgb8=False)
    st = ctx.stats()
    print(f"spp {spp} variant {variant} sched {sched:#x} render_ms {st['render_ms']:.3f} Msamples/s {st['samples']/st['render_ms']/1e3:.1f} Mrays/s {st['rays']/st['render_ms']/1e3:.1f} resolve_ms {st['resolve_ms']:.3f} bvh_depth {st['bvh_depth']} build_ms {st['build_ms']:.2f}")
