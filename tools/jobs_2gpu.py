"""dynamic sharding diagnostics, torchrun with 2+ ranks: per-rank rays and kernel time of a shared-jobs frame"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import raytracingproject_b200 as rt
from raytracingproject_b200 import sharding
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 250
s, m = rt.scene_rtiow()
ctx = rt.Context(local); ctx.upload_scene(s, m)
cam = rt.camera_rtiow(1920, spp, 50)
sharding.setup_shared_jobs(ctx, rank, world)
for mode in ("shared", "samples", "shared", "samples"):
    for rep in range(3):
        dist.barrier(); torch.cuda.synchronize()
        if mode == "shared": sharding.render_shared(ctx, cam, rt.F32, world)
        else: sharding.render_sharded(ctx, cam, rank, world, rt.F32)
        ctx.synchronize(); st = ctx.stats()
        t = torch.zeros(world * 2, dtype=torch.float64, device="cuda"); t[2 * rank] = st["render_ms"]; t[2 * rank + 1] = st["rays"]
        dist.all_reduce(t)
        if rank == 0 and rep == 2:
            tt = t.tolist()
            print(mode, "group", os.environ.get("RT_JOB_GROUP", "2"), "ms", [round(tt[2 * r], 2) for r in range(world)], "Mrays", [round(tt[2 * r + 1] / 1e6, 1) for r in range(world)],
                  "Mrays/ms per rank", [round(tt[2 * r + 1] / 1e6 / tt[2 * r], 2) for r in range(world)], flush=True)
dist.barrier(); dist.destroy_process_group()
