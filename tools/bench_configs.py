"""BASELINE.json configs other than the headline one, measured on ONE B200 (not the bench contract;
writes one JSON line per config): python tools/bench_configs.py [c1] [c2] [c4] [c5] [variants]
  C1  400x225, 10 spp                 (parity-test size)
  C2  closest-hit microbench, 2^26 synthetic rays vs the final-scene BVH (+ exact-id check of 2^20 of them vs brute force)
  C4  80 % glass/metal scene, 1920x1080, 1000 spp
  C5  10k / 100k / 1M spheres, 3840x2160, 64 spp (GPU LBVH; nodes in global memory)"""
import json, math, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt

which = set(a.lower() for a in sys.argv[1:]) or {"c1", "c2", "c4", "c5", "variants"}
ctx = rt.Context(0)
VAR = {"wavefront": rt.WAVEFRONT, "megakernel": rt.MEGAKERNEL, "lockstep": rt.MEGA_LOCKSTEP}


def timed_render(cam, prec, variant, reps=3):
    best, st = 1e30, None
    for _ in range(reps):
        ctx.render(cam, precision=prec, seed=0, variant=variant)
        ctx.synchronize(); st = ctx.stats(); best = min(best, st["render_ms"])
    return best, st


def line(name, ms, st, **kw):
    d = {"config": name, "render_ms": round(ms, 3), "msamples_per_s": round(st["samples"] / ms / 1e3, 1),
         "mrays_per_s": round(st["rays"] / ms / 1e3, 1), "rays_per_sample": round(st["rays"] / max(1, st["samples"]), 3),
         "n_spheres": st["n_spheres"], "bvh_depth": st["bvh_depth"], "scene_in_smem": st["scene_in_smem"], "slices": st["slices"]}
    d.update(kw)
    print(json.dumps(d), flush=True)


s, m = rt.scene_rtiow()
if "c1" in which:
    ctx.upload_scene(s, m)
    for name, v in VAR.items():
        for prec, pn in ((rt.F32, "f32"), (rt.F64, "f64")):
            ms, st = timed_render(rt.camera_rtiow(400, 10, 50), prec, v)
            line("C1 400x225x10spp", ms, st, variant=name, dtype=pn)

if "variants" in which:
    ctx.upload_scene(s, m)
    for name, v in VAR.items():
        for prec, pn, spp in ((rt.F32, "f32", 64), (rt.F64, "f64", 8)):
            ms, st = timed_render(rt.camera_rtiow(1920, spp, 50), prec, v)
            line(f"C3-shaped 1920x1080x{spp}spp", ms, st, variant=name, dtype=pn)

if "c2" in which:
    ctx.upload_scene(s, m)
    cam = rt.camera_rtiow(400, 10, 50)
    n = 1 << 26
    for prec, pn in ((rt.F32, "f32"), (rt.F64, "f64")):
        r = ctx.bench_closest(cam, n, seed=1, precision=prec, repeats=3)
        print(json.dumps({"config": "C2 closest-hit 2^26 rays, BVH", "dtype": pn, "ms": round(r["ms"], 3),
                          "mrays_per_s": round(n / r["ms"] / 1e3, 1)}), flush=True)
    nb = 1 << 20
    a = ctx.bench_closest(cam, nb, seed=1, precision=rt.F64, want_results=True)
    b = ctx.bench_closest(cam, nb, seed=1, precision=rt.F64, backend=rt.TRACE_BRUTE, want_results=True, repeats=1)
    f = ctx.bench_closest(cam, nb, seed=1, precision=rt.F32, want_results=True)
    print(json.dumps({"config": "C2 exact-id check, 2^20 rays", "f64_bvh_vs_brute_prim_mismatches": int((a["prim"] != b["prim"]).sum()),
                      "f64_bvh_vs_brute_t_bit_equal": bool(np.array_equal(a["t"], b["t"])),
                      "f32_vs_f64_prim_mismatches": int((f["prim"] != a["prim"]).sum()),
                      "brute_ms": round(b["ms"], 3), "brute_mrays_per_s": round(nb / b["ms"] / 1e3, 1)}), flush=True)

if "c4" in which:
    s4, m4 = rt.scene_rtiow(11, 0.2, 0.6)
    ctx.upload_scene(s4, m4)
    ty = m4["type"][s4["material"]]
    for name in ("wavefront", "megakernel"):
        ms, st = timed_render(rt.camera_rtiow(1920, 100, 50), rt.F32, VAR[name], reps=2)
        line("C4 80% glass/metal 1920x1080, timed at 100 of 1000 spp", ms, st, variant=name, dtype="f32",
             mix={"lambertian": int((ty == 0).sum()), "metal": int((ty == 1).sum()), "dielectric": int((ty == 2).sum())},
             full_frame_ms_scaled=round(ms * 10, 1))

if "c5" in which:
    for n_target in (10_000, 100_000, 1_000_000):
        g = math.ceil(math.sqrt(n_target) / 2)
        s5, m5 = rt.scene_rtiow(g)
        t0 = time.perf_counter()
        ctx.upload_scene(s5, m5, builder=rt.BVH_LBVH_GPU)
        up = (time.perf_counter() - t0) * 1e3
        st0 = ctx.stats()
        cam = rt.camera_rtiow(3840, 8, 50, scale=g / 11.0)
        for name in ("wavefront", "megakernel"):
            ms, st = timed_render(cam, rt.F32, VAR[name], reps=2)
            line(f"C5 N={len(s5)} (g={g}) 3840x2160, timed at 8 of 64 spp", ms, st, variant=name, dtype="f32", builder="gpu lbvh",
                 build_ms=round(st0["build_ms"], 2), upload_ms=round(up, 1), full_frame_ms_scaled=round(ms * 8, 1))
        if n_target <= 100_000:
            t0 = time.perf_counter(); ctx.upload_scene(s5, m5); up = (time.perf_counter() - t0) * 1e3
            ms, st = timed_render(cam, rt.F32, rt.WAVEFRONT, reps=2)
            line(f"C5 N={len(s5)} (g={g}) 3840x2160, timed at 8 of 64 spp", ms, st, variant="wavefront", dtype="f32", builder="host sah",
                 build_ms=round(ctx.stats()["build_ms"], 2), upload_ms=round(up, 1), full_frame_ms_scaled=round(ms * 8, 1))
