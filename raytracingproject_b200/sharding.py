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


def row_shard(height, rank, world):
    """Alternative tile sharding: rows j with j % world == rank (no arithmetic in the combine)."""
    return np.arange(rank, height, world)


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
