"""Exploratory GPU run: parity numbers + timings printed as text (not a test, not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
from oracle import oracle as O

def psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)

s, m = rt.scene_rtiow()
ctx = rt.Context(0)
ctx.upload_scene(s, m)
print("stats after upload", ctx.stats())

# C1: 400x225, 10 spp vs oracle (philox, seed 0)
cam = rt.camera_rtiow(400, 10, 50)
ocam = O.camera_rtiow(400, 10, 50)
t = time.time(); ref = O.render(s.view(O.SPHERE), m.view(O.MATERIAL), ocam, rng_mode=O.RNG_PHILOX, seed=0, use_bvh=True); print("oracle s", time.time() - t, "rays", ref["rays"])
for prec, name in ((rt.F64, "f64"), (rt.F32, "f32")):
    ctx.render(cam, precision=prec, seed=0)
    out = ctx.readback(10, want_ppm=True)
    st = ctx.stats()
    d = np.abs(out["rgb8"].astype(int) - ref["rgb8"].astype(int)).max(-1)
    rel = np.abs(out["sum"] - ref["sum"]).max() 
    print(name, "rays", st["rays"], "ms", st["render_ms"], "slices", st["slices"], "smem", st["scene_in_smem"],
          "| px differing", int((d > 0).sum()), "px >1:", int((d > 1).sum()), "psnr", psnr(out["rgb8"], ref["rgb8"]),
          "max |sum diff|", rel, "ppm equal", out["ppm"] == O.ppm_p3(ref["rgb8"]))

# closest hit parity on oracle-generated camera rays + bounce-like rays
rng = np.random.default_rng(0)
n = 20000
pi = rng.integers(0, 400, n); pj = rng.integers(0, 225, n); sm = rng.integers(0, 10, n)
rays, _ = O.get_rays(ocam, pi, pj, sm, seed=0)
g = ctx.gen_rays(cam, pi, pj, sm, seed=0, precision=rt.F64)
print("gen_rays f64 bit-equal:", g.tobytes() == rays.tobytes())
g32 = ctx.gen_rays(cam, pi, pj, sm, seed=0, precision=rt.F32)
print("gen_rays f32 max rel err", np.abs(g32["dir"] - O.vec3s(rays["dir"])).max() / np.abs(O.vec3s(rays["dir"])).max())
oh = O.closest_hit(s.view(O.SPHERE), rays)
for prec, name in ((rt.F64, "f64"), (rt.F32, "f32")):
    for be, bn in ((rt.TRACE_BVH, "bvh"), (rt.TRACE_BRUTE, "brute")):
        h = ctx.trace_closest(rays.view(rt.RAY), precision=prec, backend=be)
        hitm = oh["prim"] >= 0
        same = h["prim"] == oh["prim"]
        terr = np.abs(h["t"][hitm & same] - oh["t"][hitm & same]) / oh["t"][hitm & same]
        print(name, bn, "prim mismatches", int((~same).sum()), "of", n, "max rel t err", terr.max() if len(terr) else 0,
              "t bit-equal" if np.array_equal(h["t"][same], oh["t"][same]) else "")

# C3-like timing: 1920x1080 at a few spp
for prec, name in ((rt.F32, "f32"), (rt.F64, "f64")):
    for spp in (8, 32):
        cam3 = rt.camera_rtiow(1920, spp, 50)
        ctx.render(cam3, precision=prec, seed=0)
        st = ctx.stats()
        print("C3", name, "spp", spp, "ms", round(st["render_ms"], 2), "Msamples/s", round(st["samples"] / st["render_ms"] / 1e3, 1),
              "Mrays/s", round(st["rays"] / st["render_ms"] / 1e3, 1), "rays/sample", round(st["rays"] / st["samples"], 3), "slices", st["slices"])
# C2-like: closest-hit microbench
for prec, name in ((rt.F32, "f32"), (rt.F64, "f64")):
    r = ctx.bench_closest(cam, 1 << 24, seed=1, precision=prec, repeats=3)
    print("C2", name, "2^24 rays ms", round(r["ms"], 3), "Mrays/s", round((1 << 24) / r["ms"] / 1e3, 1))
