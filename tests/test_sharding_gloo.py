"""CPU, world_size 2 over gloo: the N>1 host path -- sample-range shard planning, the SUM reduce of
partial accumulation buffers to rank 0, and the resolve of the combined buffer.  The oracle stands in
for the device render here (this is a test of the host logic; the GPU version is test_gpu_multi)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    from raytracingproject_b200.sharding import reduce_partials, sample_shard
    s, m = O.scene_rtiow()
    spp = 5
    cam = O.camera_rtiow(96, spp, 50)
    begin, count = sample_shard(spp, rank, world)
    part = O.render(s, m, cam, rng_mode=O.RNG_PHILOX, seed=3, use_bvh=True, nthreads=2, sample0=begin, nsamples=count)
    H, W = part["height"], part["width"]
    buf = np.zeros((H, W, 4), "f8")                      # same layout as the device buffer: (sum.rgb, samples)
    buf[..., :3] = part["sum"]
    buf[..., 3] = count
    t = torch.from_numpy(buf.reshape(-1))
    reduce_partials(t, dst=0)
    rays = torch.tensor([part["rays"]], dtype=torch.int64)
    dist.reduce(rays, dst=0)
    if rank == 0:
        full = O.render(s, m, cam, rng_mode=O.RNG_PHILOX, seed=3, use_bvh=True, nthreads=2)
        tot = t.numpy().reshape(H, W, 4)
        ok_count = bool(np.all(tot[..., 3] == spp))
        # partial sums are added in a different order than the sequential per-pixel sum: 1e-13 relative
        ok_sum = bool(np.allclose(tot[..., :3], full["sum"], rtol=1e-13, atol=1e-13))
        rgb = O.write_color(tot[..., :3].reshape(-1, 3), spp).reshape(H, W, 3)
        q.put((ok_count, ok_sum, bool(np.array_equal(rgb, full["rgb8"])), int(rays.item()) == full["rays"]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sample_sharding_reduces_to_the_full_image():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res == (True, True, True, True)
