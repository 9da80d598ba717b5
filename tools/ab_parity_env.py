"""Parity of an environment knob: python tools/ab_parity_env.py KNOB v1,v2,...  -- the f64 wavefront image and ray count of a small
frame under every value must equal the CPU oracle's, the f32 image must stay within 1/255 of it on > 99.9 % of pixels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
from oracle import oracle as O
knob = sys.argv[1]; vals = sys.argv[2].split(",")
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
W, spp = 200, 6
cam = rt.camera_rtiow(W, spp, 50)
ref = O.render(s.view(O.SPHERE), m.view(O.MATERIAL), O.camera_rtiow(W, spp, 50), rng_mode=O.RNG_PHILOX, seed=0, use_bvh=True, nthreads=os.cpu_count())
ok = True
for v in vals:
    if v == "-": os.environ.pop(knob, None)
    else: os.environ[knob] = v
    ctx.render(cam, precision=rt.F64, seed=0, variant=rt.WAVEFRONT)
    o = ctx.readback(spp); st = ctx.stats()
    same = np.array_equal(o["rgb8"], ref["rgb8"]) and st["rays"] == ref["rays"]
    ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT)
    o32 = ctx.readback(spp); r32 = ctx.stats()["rays"]
    frac = float((np.abs(o32["rgb8"].astype(int) - ref["rgb8"].astype(int)).max(-1) <= 1).mean())
    print(f"{knob}={v}: f64 image+rays == oracle: {same} (rays {st['rays']} vs {ref['rays']}); f32 within 1/255: {frac:.5f}, f32 rays {r32}", flush=True)
    ok = ok and same and frac > 0.999
sys.exit(0 if ok else 1)
