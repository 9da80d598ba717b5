import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
from oracle import oracle as O
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
cam = rt.camera_rtiow(200, 8, 50)
ref = O.render(s.view(O.SPHERE), m.view(O.MATERIAL), O.camera_rtiow(200, 8, 50), rng_mode=O.RNG_PHILOX, seed=4, use_bvh=True)
for sl in (1, 2, 4, 8, 0):
    ctx.render(cam, precision=rt.F64, seed=4, slices=sl)
    o = ctx.readback(8)
    d = (o["rgb8"] != ref["rgb8"]).any(-1)
    print("slices", sl, ctx.stats()["slices"], "rays", ctx.stats()["rays"], ref["rays"], "diff px", d.sum(), np.argwhere(d)[:4].tolist(),
          "max rel sum", np.abs(o["sum"] - ref["sum"]).max())
    if d.any():
        j, i = np.argwhere(d)[0]
        print(o["sum"][j, i], ref["sum"][j, i], o["rgb8"][j, i], ref["rgb8"][j, i])
