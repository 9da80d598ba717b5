"""A/B of the 4-wide node records on ONE box: C5 scenes (GPU LBVH, tree in global memory), 3840x2160.
python tools/wide_ab.py [spp] [g,g,...]   (g = 50 / 159 / 500: 10 k / 100 k / 1 M spheres)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
gs = [int(x) for x in (sys.argv[2] if len(sys.argv) > 2 else "50,159,500").split(",")]
ctx = rt.Context(0)
for g in gs:
    s, m = rt.scene_rtiow(g)
    cam = rt.camera_rtiow(3840, spp, 50, scale=g / 11.0)
    imgs = {}
    for wide, name in ((-1, "binary 64 B"), (1, "4-wide 128 B")):
        ctx.upload_scene(s, m, builder=rt.BVH_LBVH_GPU, wide=wide)
        best = 1e30
        for _ in range(3):
            ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT)
            ctx.synchronize(); st = ctx.stats(); best = min(best, st["render_ms"])
        imgs[wide] = ctx.readback(spp, want_sum=False)["rgb8"]
        print(f"g {g:4d} spheres {st['n_spheres']:8d} depth {st['bvh_depth']:2d} {name:13s} ms {best:8.3f} Msamples/s {st['samples'] / best / 1e3:8.1f} "
              f"Mrays/s {st['rays'] / best / 1e3:8.1f} rays {st['rays']}", flush=True)
    d = np.abs(imgs[-1].astype(int) - imgs[1].astype(int))
    print(f"       images identical: {bool((d == 0).all())} (max diff {d.max()})", flush=True)
