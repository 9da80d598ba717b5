#!/usr/bin/env python
"""Where the host-side milliseconds of an end-to-end frame go: upload, render, read-back (whole / split), per config.
usage: python tools/e2e_breakdown.py [c3|c5-10k|c5-100k|c5-1m] ..."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import raytracingproject_b200 as rt
import bench

def pinned(a):
    t = torch.empty(a.nbytes, dtype=torch.uint8, pin_memory=True)
    v = t.numpy().view(a.dtype).reshape(a.shape); v[...] = a
    return v, t

def main():
    for name in (sys.argv[1:] or ["c3"]):
        workload, W, H, spp, g, pd, pm, builder = bench.CONFIGS[name]
        ctx = rt.Context(0)
        s, m = rt.scene_rtiow(g, pd, pm)
        (s, ks), (m, km) = pinned(s), pinned(m)
        bid = rt.BVH_LBVH_GPU if builder == "lbvh" else rt.BVH_SAH_HOST
        cam = rt.camera_rtiow(W, spp, 50, scale=g / 11.0)
        rgb = torch.empty(W * H * 3, dtype=torch.uint8, pin_memory=True).numpy()
        ppm = torch.empty(W * H * 12 + 64, dtype=torch.uint8, pin_memory=True).numpy()
        def t(f):
            t0 = time.perf_counter(); r = f(); return (time.perf_counter() - t0) * 1e3, r
        rows = []
        for it in range(4):
            up, _ = t(lambda: ctx.upload_scene(s, m, builder=bid))
            rl, _ = t(lambda: ctx.render(cam, precision=rt.F32, seed=0))
            rw, _ = t(ctx.synchronize)
            rb, _ = t(lambda: ctx.readback_into(spp, rgb, ppm))
            rows.append(("whole", up, rl, rw, rb, 0.0))
        for it in range(4):
            up, _ = t(lambda: ctx.upload_scene(s, m, builder=bid))
            rl, _ = t(lambda: ctx.render(cam, precision=rt.F32, seed=0))
            rw, _ = t(ctx.synchronize)
            bg, _ = t(lambda: ctx.readback_begin(spp, rgb, ppm))
            up2, _ = t(lambda: ctx.upload_scene(s, m, builder=bid))          # the next frame's upload, copies in flight
            en, _ = t(ctx.readback_end)
            rows.append(("split", up, rl, rw, bg, en, up2))
        print(name, "kind upload render_launch render_wait readback|begin end [next_upload]  (host ms)")
        for r in rows:
            print("  ", r[0], " ".join(f"{x:8.3f}" for x in r[1:]))
        ctx.close()

if __name__ == "__main__":
    main()
