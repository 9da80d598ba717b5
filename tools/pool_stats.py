"""Scheduler statistics of the wavefront kernel's TRACE stage.  Needs a library built with the counters:
    make -C raytracingproject_b200/csrc clean && make -C raytracingproject_b200/csrc -j8 STATS=1
    python tools/pool_stats.py [spp] [sched]
(rebuild without STATS=1 afterwards: the counters cost registers and time)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracingproject_b200 as rt
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
sched = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0
s, m = rt.scene_rtiow()
ctx = rt.Context(0); ctx.upload_scene(s, m)
cam = rt.camera_rtiow(1920, spp, 50)
ctx.render(cam, precision=rt.F32, seed=0, variant=rt.WAVEFRONT, sched=sched)
c = [int(v) for v in ctx.debug_counters()]
rays, st = c[0], c[1:]
if st[0] == 0:
    sys.exit("library was built without STATS=1")
d = st[0]
print(f"rays {rays}  scheduler decisions {d} ({d / rays * 32:.2f} per 32 rays)")
print(f"lanes at a decision: node {st[1]/d:.1f}  leaf {st[2]/d:.1f}  idle {st[3]/d:.1f}   queue not empty at {100*st[4]/d:.0f}% of them")
print(f"pull  sections {st[5]:>10} ({100*st[5]/d:4.1f}%)  rays taken {st[6]/max(st[5],1):.1f}  lanes busy meanwhile {st[7]/max(st[5],1):.1f}")
print(f"leaf  sections {st[8]:>10} ({100*st[8]/d:4.1f}%)  leaf lanes {st[9]/max(st[8],1):.1f}  node lanes waiting {st[10]/max(st[8],1):.1f}  idle {st[11]/max(st[8],1):.1f}")
print(f"node  sections {st[12]:>10} ({100*st[12]/d:4.1f}%)  node lanes at entry {st[13]/max(st[12],1):.1f}  leaf lanes waiting {st[14]/max(st[12],1):.1f}  idle {st[15]/max(st[12],1):.1f}")
print(f"      visit pairs {st[16]} ({st[16]/max(st[12],1):.2f} per section)  lanes at the head of a pair {st[17]/max(st[16],1):.1f}")
print(f"node visits per ray (upper bound, 2 per lane per pair): {2*st[17]/rays:.1f}   leaf visits per ray {st[9]/rays:.2f}")
