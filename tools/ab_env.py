"""Same-process A/B of environment knobs the library reads at every rt_render (RT_TH_X ...):
    python tools/ab_env.py KNOB v1,v2,... [config[:spp]] [reps] [sched]
renders the config's frame with every value in turn (two interleaved rounds), device time from rt_stats.render_ms."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
knob = sys.argv[1]; vals = sys.argv[2].split(",")
cfg = (sys.argv[3] if len(sys.argv) > 3 else "c3:125").split(":")
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
sched = int(sys.argv[5], 0) if len(sys.argv) > 5 else 0
SC = {"c3": (11, .8, .95, 1920, "sah"), "c4": (11, .2, .6, 1920, "sah"), "c5-10k": (50, .8, .95, 3840, "lbvh"),
      "c5-100k": (159, .8, .95, 3840, "lbvh"), "c5-1m": (500, .8, .95, 3840, "lbvh")}
g, pd, pm, W, b = SC[cfg[0]]; spp = int(cfg[1]) if len(cfg) > 1 else 125
s, m = rt.scene_rtiow(g, pd, pm)
ctx = rt.Context(0)
ctx.upload_scene(s, m, builder=rt.BVH_LBVH_GPU if b == "lbvh" else rt.BVH_SAH_HOST)
cam = rt.camera_rtiow(W, spp, 50, scale=g / 11.0)
res = {v: [] for v in vals}; rays = {}
for rnd in range(2):
    for v in vals:
        if v == "-": os.environ.pop(knob, None)
        else: os.environ[knob] = v
        for r in range(reps + (1 if rnd == 0 else 0)):
            ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT, sched=sched)
            st = ctx.stats()
            if r or rnd: res[v].append(st["render_ms"])
        rays[v] = st["rays"]
base = min(res[vals[0]])
for v in vals:
    t = min(res[v])
    print(f"{cfg[0]} {spp} spp  {knob}={v:>4s}  best {t:8.3f} ms  median {statistics.median(res[v]):8.3f} ms  {st['samples']/t/1e3:8.1f} Msamples/s  vs first {100*(t/base-1):+5.1f}%  rays {rays[v]}", flush=True)
