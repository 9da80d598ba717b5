"""Multi-GPU host logic: one process per GPU, sample-range sharding.

Every (pixel, sample) is independent once the RNG is keyed, so rank g renders samples
[begin_g, begin_g + count_g) of EVERY pixel into its own accumulation buffer and there is no
exchange while tracing.  The partial buffers are combined once at the end, either

* ``combine="nccl"``: ``torch.distributed.reduce(SUM)`` to rank 0 over NVLink/NVSwitch on a tensor
  that aliases the library's device buffer (torch is plumbing only); every rank first folds its
  sample layers into layer 0 (``ctx.fold_slices()``) so that one layer per rank is reduced, or
* ``combine="ipc"``: rank 0 maps every peer's buffer through CUDA IPC and its resolve kernel reads
  them over NVLink while it resolves -- one fused reduce+resolve kernel, fixed summation order.

The scene and BVH are replicated (15.5 kB + 31 kB for the final scene).
"""
import numpy as np


def sample_shard(spp, rank, world):
    """Samples [begin, begin+count) of each pixel owned by `rank`: contiguous, sizes differ by <= 1."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    begin = (spp * rank) // world
    end = (spp * (rank + 1)) // world
    return begin, end - begin


class _CudaAlias:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can alias it."""

    def __init__(self, ptr, n_elems, elem_bytes):
        self.__cuda_array_interface__ = {
            "shape": (int(n_elems),), "typestr": "<f4" if elem_bytes == 4 else "<f8",
            "data": (int(ptr), False), "version": 3, "strides": None}


def alias_accum_tensor(ctx, layers=None):
    """torch tensor (cuda) aliasing the context's accumulation buffer [slices*H*W*4], or its first
    `layers` sample layers (layers=1 after ctx.fold_slices(): the rank's whole partial sum)."""
    import torch
    ptr, n, eb = ctx.accum_buffer()
    if layers is not None:
        n = min(n, int(layers) * ctx.width * ctx.height * 4)
    return torch.as_tensor(_CudaAlias(ptr, n, eb), device="cuda")


def reduce_partials(t, dst=0, group=None):
    """SUM-reduce the per-rank partial accumulation tensors to `dst` (NCCL on GPU, gloo in tests)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return t


def exchange_ipc_handles(handle, group=None):
    """all_gather of the 64-byte CUDA IPC handles; returns the list indexed by rank."""
    import torch.distributed as dist
    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, handle, group=group)
    return out


def render_sharded(ctx, cam, rank, world, precision, seed=0, slices=0, variant=1):
    """Render this rank's sample shard of `cam` (asynchronous); returns (begin, count).
    variant: 1 = SM-resident wavefront (default), 0 = warp-scheduled megakernel, 2 = lock-step megakernel."""
    begin, count = sample_shard(cam.samples_per_pixel, rank, world)
    if count > 0:
        ctx.render(cam, precision=precision, seed=seed, sample_begin=begin, sample_count=count, slices=slices, variant=variant)
    return begin, count


# ------------------------------------------------------------------------- dynamic sharding --
# The third partition: the ranks take the (tile, sample layer) jobs of ONE frame from a counter in rank 0's memory, advanced
# by every warp of every GPU with a system-scope atomic over NVLink (rt_jobs_*).  A slower GPU takes fewer jobs, all ranks
# finish together; the partial buffers are combined like sample shards (fold + reduce).

def setup_shared_jobs(ctx, rank, world, group=None):
    """Once: rank 0 creates the frame counters and exports them, every other rank maps them."""
    import torch.distributed as dist
    blob = [ctx.jobs_export() if rank == 0 else None]
    dist.broadcast_object_list(blob, src=0, group=group)
    if rank != 0:
        ctx.jobs_open(blob[0])
    dist.barrier(group=group)


def render_shared(ctx, cam, precision, world=1, seed=0):
    """This rank's share of the frame's jobs -- whatever it gets -- of ALL samples (asynchronous).  world = how many GPUs share
    the list (the frame is cut into jobs for all their warps)."""
    ctx.render(cam, precision=precision, seed=seed, shared_jobs=max(1, int(world)))


# ------------------------------------------------------------------ tile-interleaved sharding --
# The alternative partition of SURVEY.md 8e: rank g renders ALL samples of the tile pairs whose index is g modulo the
# world size (pairs of 16 x 8 pixel tiles, interleaved finely enough that sky and sphere rows balance out), so the combine
# has no arithmetic at all: each rank's resolve kernel writes the 8-bit pixels of its tiles straight into rank 0's image
# over NVLink (CUDA IPC mapping) and bumps a counter there; rank 0's stream waits for the counter, then encodes.
# Per frame that is 6 MB / world of stores per rank instead of a fold + a 33 MB float reduce.

def setup_tile_frame(ctx, rank, world, group=None):
    """Once per image size, after this context has rendered a frame of that size: rank 0 exports its image,
    every other rank maps it."""
    import torch.distributed as dist
    blob = [ctx.frame_export() if rank == 0 else None]
    dist.broadcast_object_list(blob, src=0, group=group)
    if rank != 0:
        ctx.frame_open(blob[0])
    dist.barrier(group=group)


def render_tiles(ctx, cam, rank, world, precision, seed=0):
    """This rank's tiles of `cam`'s frame, all samples (asynchronous).  Ranks other than 0 also resolve their tiles into
    rank 0's image; rank 0 calls ctx.readback(...) afterwards, which waits for them on its stream."""
    ctx.render(cam, precision=precision, seed=seed, tile_rank=rank, tile_world=world)
    if rank != 0 and world > 1:
        ctx.resolve_tiles(cam.samples_per_pixel)


def render_tiles_combined(ctx, cam, rank, world, precision, seed=0, group=None):
    """Convenience for tests / one-off frames: set-up, one tile-sharded frame, the combined image on rank 0
    (dict(rgb8, ppm, rays); None elsewhere)."""
    import torch
    import torch.distributed as dist
    ctx.render(cam, precision=precision, seed=seed, tile_rank=rank, tile_world=world, sample_count=1)   # defines the image size
    ctx.synchronize()
    setup_tile_frame(ctx, rank, world, group)
    render_tiles(ctx, cam, rank, world, precision, seed)
    rays = torch.tensor([ctx.stats()["rays"]], dtype=torch.int64, device="cuda")
    dist.all_reduce(rays, group=group)
    out = None
    if rank == 0:
        out = ctx.readback(cam.samples_per_pixel, want_sum=False, want_ppm=True)
        out["rays"] = int(rays.item())
    dist.barrier(group=group)
    return out
