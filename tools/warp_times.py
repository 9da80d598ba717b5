"""The frame's tail, measured: python tools/warp_times.py [spp ...]   (library built with `make TIMES=1`, via RT_B200_LIB)
Per warp of render_pool_kernel: when it started, when the frame ran out of jobs for it, when it exited, and how full its pool
was (rays per round) before and after that moment."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
spps = [int(a) for a in sys.argv[1:]] or [62, 500]
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
for spp in spps:
    cam = rt.camera_rtiow(1920, spp, 50)
    for _ in range(2):
        ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT)
        st = ctx.stats()
    w = ctx.debug_warp_times().astype(np.int64)
    w = w[w[:, 7] == 1]
    if len(w) == 0:
        sys.exit("library was built without TIMES=1")
    t0 = w[:, 0].min()
    first, exh, ext = (w[:, 0] - t0) / 1e6, (w[:, 1] - t0) / 1e6, (w[:, 2] - t0) / 1e6
    end = ext.max()
    q = lambda a: " ".join(f"{np.percentile(a, p):8.3f}" for p in (0, 10, 50, 90, 100))
    print(f"spp {spp}: kernel {st['render_ms']:.3f} ms, {len(w)} warps, slices {st['slices']}, last exit at {end:.3f} ms")
    print(f"  start            (ms, pct 0/10/50/90/100): {q(first)}")
    print(f"  out of jobs      (ms)                    : {q(exh)}")
    print(f"  exit             (ms)                    : {q(ext)}")
    print(f"  exit - out of jobs (a warp's own drain)  : {q(ext - exh)}")
    print(f"  idle before the last warp exits, mean over warps: {np.mean(end - ext):.3f} ms = {100 * np.mean(end - ext) / end:.2f} % of the frame")
    print(f"  rays per round while jobs remain: {w[:, 4].sum() / max(1, w[:, 3].sum()):.1f} of 128   after: {w[:, 6].sum() / max(1, w[:, 5].sum()):.1f}"
          f"   rounds after / all: {100 * w[:, 5].sum() / (w[:, 3].sum() + w[:, 5].sum()):.2f} %")
