"""GPU, 2 ranks (one process per GPU, spawn, NCCL over NVLink): the N > 1 path of SURVEY.md 8e on real devices.
Skipped on a box with fewer than 2 GPUs (the CPU twin is tests/test_sharding_gloo.py).

Every combine the library offers is rendered by two ranks and compared, on rank 0, with the SAME frame rendered by
rank 0 alone and with the CPU oracle:
  samples+nccl   sample-range shards, fold, torch.distributed.reduce(SUM) on a tensor aliasing the library's buffer
  samples+ipc    sample-range shards, fold, rank 0's resolve kernel reads the peer's buffer over NVLink (CUDA IPC)
  tiles          tile-interleaved shards: every rank renders all samples of its tiles, resolves them and writes the
                 bytes straight into rank 0's image over NVLink -- no arithmetic in the combine
  jobs+nccl      dynamic shards (fast mode): both GPUs take the (tile, sample layer) jobs of the frame from one counter in
                 rank 0's memory (system-scope atomics over NVLink), fold, NCCL reduce
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import raytracingproject_b200 as rt
    from raytracingproject_b200 import sharding
    res = {}
    try:
        ctx = rt.Context(rank)
        s, m = rt.scene_rtiow()
        ctx.upload_scene(s, m)
        stream = torch.cuda.ExternalStream(ctx.stream(), device=torch.device("cuda", rank))
        for prec, pname, width, spp in ((rt.F32, "f32", 400, 9), (rt.F64, "f64", 160, 5)):
            cam = rt.camera_rtiow(width, spp, 50)
            # --- sample shards, NCCL reduce of the folded layer
            sharding.render_sharded(ctx, cam, rank, world, prec, seed=5)
            rays = torch.tensor([ctx.stats()["rays"]], dtype=torch.int64, device="cuda")
            ctx.fold_slices()
            acc = sharding.alias_accum_tensor(ctx, layers=1)
            with torch.cuda.stream(stream):
                sharding.reduce_partials(acc, dst=0)
            stream.synchronize()
            dist.all_reduce(rays)
            if rank == 0:
                res[pname + "_nccl"] = ctx.readback(spp, want_sum=False)["rgb8"]
                res[pname + "_rays"] = int(rays.item())
            dist.barrier()
            # --- sample shards, fused IPC peer-reduce resolve
            sharding.render_sharded(ctx, cam, rank, world, prec, seed=5)
            ctx.synchronize()
            handles = sharding.exchange_ipc_handles(ctx.ipc_export())
            if rank == 0:
                ctx.ipc_open_peers([h for r, h in enumerate(handles) if r != 0])
                try:
                    ctx.readback(spp)                      # peers mapped, not folded: must be refused
                    res[pname + "_unfolded_refused"] = False
                except rt.RtError:
                    res[pname + "_unfolded_refused"] = True
            ctx.fold_slices()
            ctx.synchronize()
            dist.barrier()                                 # the peer's folded layer is complete before rank 0 reads it
            if rank == 0:
                res[pname + "_ipc"] = ctx.readback(spp, want_sum=False)["rgb8"]
                ctx.ipc_close_peers()
            dist.barrier()
            # --- tile shards, bytes written into rank 0's image over NVLink
            res_tiles = sharding.render_tiles_combined(ctx, cam, rank, world, prec, seed=5)
            if rank == 0:
                res[pname + "_tiles"] = res_tiles["rgb8"]
                res[pname + "_tiles_ppm"] = res_tiles["ppm"]
                res[pname + "_tiles_rays"] = res_tiles["rays"]
            dist.barrier()
            # --- dynamic shards: one job counter for both GPUs (fast mode only), three frames in a row
            if prec == rt.F32:
                sharding.setup_shared_jobs(ctx, rank, world)
                for frame in range(3):
                    sharding.render_shared(ctx, cam, prec, world, seed=5)
                    rays = torch.tensor([ctx.stats()["rays"]], dtype=torch.int64, device="cuda")
                    ctx.fold_slices()
                    acc = sharding.alias_accum_tensor(ctx, layers=1)
                    with torch.cuda.stream(stream):
                        sharding.reduce_partials(acc, dst=0)
                    stream.synchronize()
                    both = torch.tensor([1 if ctx.stats()["rays"] > 0 else 0], dtype=torch.int64, device="cuda")
                    dist.all_reduce(rays); dist.all_reduce(both)
                    if rank == 0:
                        res[f"f32_jobs{frame}"] = ctx.readback(spp, want_sum=False)["rgb8"]
                        res[f"f32_jobs{frame}_rays"] = int(rays.item())
                        res[f"f32_jobs{frame}_ranks_with_work"] = int(both.item())
                    dist.barrier()
                ctx.jobs_close()
                try:
                    ctx.render(cam, precision=prec, seed=5, shared_jobs=True)      # counters closed: must be refused
                    res["jobs_closed_refused"] = False
                except rt.RtError:
                    res["jobs_closed_refused"] = True
                dist.barrier()
            # --- the same frame on rank 0 alone
            if rank == 0:
                ctx.render(cam, precision=prec, seed=5)
                res[pname + "_solo"] = ctx.readback(spp, want_sum=False)["rgb8"]
                res[pname + "_solo_rays"] = ctx.stats()["rays"]
            dist.barrier()
        # a peer that rendered another frame size is refused
        ctx.render(rt.camera_rtiow(64 if rank == 0 else 96, 1, 5), precision=rt.F32)
        ctx.synchronize()
        handles = sharding.exchange_ipc_handles(ctx.ipc_export())
        if rank == 0:
            try:
                ctx.ipc_open_peers([handles[1]])
                res["mismatch_refused"] = False
            except rt.RtError:
                res["mismatch_refused"] = True
        dist.barrier()
        ctx.close()
    except Exception as e:                                  # report instead of hanging the other rank
        import traceback
        res["error"] = f"rank {rank}: {e}\n{traceback.format_exc()}"
    if rank == 0 or "error" in res:
        q.put(res)
    try:
        dist.destroy_process_group()
    except Exception:
        pass


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_gpus_every_combine_gives_the_single_gpu_frame():
    import torch.multiprocessing as mp
    from oracle import oracle as O
    mpctx = mp.get_context("spawn")
    q = mpctx.Queue()
    port = _free_port()
    procs = [mpctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
    assert "error" not in res, res.get("error")
    assert all(p.exitcode == 0 for p in procs)
    assert res["mismatch_refused"] and res["f32_unfolded_refused"] and res["f64_unfolded_refused"]
    s, m = O.scene_rtiow()
    for pname, width, spp in (("f32", 400, 9), ("f64", 160, 5)):
        ref = O.render(s, m, O.camera_rtiow(width, spp, 50), rng_mode=O.RNG_PHILOX, seed=5, use_bvh=True)
        solo = res[pname + "_solo"]
        if pname == "f64":
            assert np.array_equal(solo, ref["rgb8"]) and res["f64_solo_rays"] == ref["rays"] == res["f64_rays"] == res["f64_tiles_rays"]
            # tile shards add nothing up: every pixel is one rank's own sequential sum -> the oracle's image, bit for bit
            assert np.array_equal(res["f64_tiles"], ref["rgb8"])
            assert res["f64_tiles_ppm"] == O.ppm_p3(ref["rgb8"])
        else:
            assert abs(res["f32_rays"] - ref["rays"]) / ref["rays"] < 5e-3
            assert (np.abs(solo.astype(int) - ref["rgb8"].astype(int)).max(-1) <= 1).mean() >= 0.999
        if pname == "f32":
            assert res["jobs_closed_refused"]
            for frame in range(3):
                d = np.abs(res[f"f32_jobs{frame}"].astype(int) - solo.astype(int))
                assert d.max() <= 1 and (d > 0).any(-1).mean() < 2e-3, ("jobs", frame, int(d.max()))
                assert res[f"f32_jobs{frame}_rays"] == res["f32_solo_rays"]       # every job rendered exactly once, by one GPU or the other
                assert res[f"f32_jobs{frame}_ranks_with_work"] >= 1                # (a frame this small may go to whichever GPU starts first)
        for mode in ("nccl", "ipc", "tiles"):
            d = np.abs(res[f"{pname}_{mode}"].astype(int) - solo.astype(int))
            # two partial sums re-associate the per-pixel sum (float: the last bits; double: 1e-16): a byte may move by one
            assert d.max() <= 1, (pname, mode, int(d.max()))
            assert (d > 0).any(-1).mean() < (2e-3 if pname == "f32" else 1e-4), (pname, mode, float((d > 0).any(-1).mean()))
