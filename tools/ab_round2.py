#!/usr/bin/env python
"""Same-box A/B of library builds (round 2): every library named on the command line renders the same frames in the
same process order, device time from the library's own CUDA events (rt_stats.render_ms), best and median of `reps`.

  python tools/ab_round2.py [--configs c3,c4,c5-10k,c5-100k,c5-1m] [--reps 5] name=path/to/librt_b200_x.so ...

A library is loaded in a fresh subprocess (RT_B200_LIB), so builds with different ABI-compatible internals can be
compared; `name=@r1` runs the round-1 tree copied under ab/r1 (its own Python bindings: the ABI differs).
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import json, os, sys, statistics
sys.path.insert(0, os.environ["AB_ROOT"])
import numpy as np
import raytracingproject_b200 as rt
cfgs = json.loads(os.environ["AB_CFGS"]); reps = int(os.environ["AB_REPS"])
SC = {"c3": (11, .8, .95, 1920, "sah"), "c4": (11, .2, .6, 1920, "sah"), "c5-10k": (50, .8, .95, 3840, "lbvh"),
      "c5-100k": (159, .8, .95, 3840, "lbvh"), "c5-1m": (500, .8, .95, 3840, "lbvh")}
ctx = rt.Context(0)
for name, spp in cfgs:
    g, pd, pm, W, b = SC[name]
    s, m = rt.scene_rtiow(g, pd, pm)
    kw = json.loads(os.environ.get("AB_RENDER_KW", "{}"))
    ukw = {k[2:]: v for k, v in kw.items() if k.startswith("u_")}            # u_max_leaf etc.: upload_scene keywords
    kw = {k: v for k, v in kw.items() if not k.startswith("u_")}
    if "builder" in ukw: b = ukw.pop("builder")
    if "width" in kw: W = kw.pop("width")
    ctx.upload_scene(s, m, builder=rt.BVH_LBVH_GPU if b == "lbvh" else rt.BVH_SAH_HOST, **ukw)
    cam = rt.camera_rtiow(W, spp, 50, scale=g / 11.0)
    ts = []
    for r in range(reps + 1):
        ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT, **kw)
        st = ctx.stats()
        if r: ts.append(st["render_ms"])
    print(json.dumps({"lib": os.environ["AB_NAME"], "config": name, "spp": spp, "best_ms": min(ts), "median_ms": statistics.median(ts),
                      "msamples_per_s": st["samples"] / min(ts) / 1e3, "mrays_per_s": st["rays"] / min(ts) / 1e3, "rays": st["rays"],
                      "scene_in_smem": st.get("scene_in_smem"), "top_nodes": st.get("top_nodes_in_smem")}), flush=True)
'''


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c3:125,c4:125,c5-10k:16,c5-100k:16,c5-1m:16")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--rounds", type=int, default=2, help="the whole list of libraries is run this many times, interleaved")
    ap.add_argument("libs", nargs="+")
    a = ap.parse_args()
    cfgs = [(c.split(":")[0], int(c.split(":")[1])) for c in a.configs.split(",")]
    for rnd in range(a.rounds):
        for spec in a.libs:
            name, path = spec.split("=", 1)
            env = dict(os.environ, AB_CFGS=json.dumps(cfgs), AB_REPS=str(a.reps), AB_NAME=name)
            kw = {}
            if "+" in path:                                   # lib+{"top_cap": -1}: extra render keywords
                path, extra = path.split("+", 1)
                kw = json.loads(extra)
            for k in [k for k in kw if k.startswith("e_")]:           # e_NAME: environment variable NAME for the child
                env[k[2:]] = str(kw.pop(k))
            env["AB_RENDER_KW"] = json.dumps(kw)
            if path.startswith("@"):
                env["AB_ROOT"] = os.path.join(ROOT, "ab", path[1:])
                env.pop("RT_B200_LIB", None)
            else:
                env["AB_ROOT"] = ROOT
                env["RT_B200_LIB"] = os.path.abspath(path)
            for k in [k for k in env if k.startswith("RT_ENV_")]:
                pass
            if name.endswith("!notop"):
                env["RT_NO_TOP_FIRST"] = "1"
            r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
            sys.stdout.write(r.stdout)
            if r.returncode != 0:
                sys.stdout.write(json.dumps({"lib": name, "error": r.stderr[-800:]}) + "\n")
            sys.stdout.flush()


if __name__ == "__main__":
    main()
