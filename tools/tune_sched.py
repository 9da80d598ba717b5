"""Sweep of the warp scheduler's knobs on a C3-shaped frame: python tools/tune_sched.py [spp] [f32|f64]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracingproject_b200 as rt
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 32
prec = rt.F64 if len(sys.argv) > 2 and sys.argv[2] == "f64" else rt.F32
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
cam = rt.camera_rtiow(1920, spp, 50)
def run(variant, sched=0, slices=0, reps=3):
    best = 1e30
    for _ in range(reps):
        ctx.render(cam, precision=prec, seed=0, variant=variant, sched=sched, slices=slices)
        ctx.synchronize(); st = ctx.stats(); best = min(best, st["render_ms"])
    return best, st
VAR = int(os.environ.get("VARIANT", str(rt.MEGAKERNEL)))
ms, st = run(rt.MEGA_LOCKSTEP)
ref = ctx.readback(spp, want_sum=False)["rgb8"]
print(f"lockstep            ms {ms:8.3f} Msamples/s {st['samples']/ms/1e3:8.1f} Mrays/s {st['rays']/ms/1e3:8.1f} slices {st['slices']}")
shapes = [int(x) for x in os.environ.get("SHAPES", "1,2,3,4").split(",")]
ths = [int(x) for x in os.environ.get("THS", "1,8,12,16,20,24,28,32").split(",")]
thl = [int(x) for x in os.environ.get("THL", "33").split(",")]
sls = [int(x) for x in os.environ.get("SLICES", "0").split(",")]
thn = [int(x) for x in os.environ.get("THN", "0").split(",")]
for shape in shapes:
    for t in ths:
        for l in thl:
            for sl, tn in [(a, b) for a in sls for b in thn]:
                ms, st = run(VAR, sched=t | (l << 8) | (shape << 16) | (tn << 20), slices=sl)
                img = ctx.readback(spp, want_sum=False)["rgb8"]
                same = float((np.abs(img.astype(int) - ref.astype(int)).max(-1) <= 1).mean())
                print(f"shape {shape} th_s {t:2d} th_l {l:2d} th_n {tn} slices {st['slices']:2d} ms {ms:8.3f} Msamples/s {st['samples']/ms/1e3:8.1f} Mrays/s {st['rays']/ms/1e3:8.1f} same_image {same}", flush=True)
