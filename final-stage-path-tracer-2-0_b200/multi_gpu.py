"""Multi-GPU rendering: the scene is replicated, sample indices are sharded across ranks,
and the per-rank accumulation films (rgb sum + sample count per pixel) are combined with one
reduce over NCCL / NVLink (gloo on CPU in the tests).  torch.distributed is plumbing only."""
import torch
import torch.distributed as dist


def shard_samples(total, rank, world):
    """Rank r renders the samples congruent to (total - 1 - r) modulo world of every pixel, so
    rank 0 — where the film is resolved and read — owns the last sample, whose depth-0 hit fills
    the aux buffers (reference frame.cpp:78-85: last sample wins).  Returns (first_sample,
    n_samples, stride); per-sample random streams depend only on (seed, pixel, sample index),
    so the union over ranks is exactly the single-GPU sample set."""
    if world <= 0 or not 0 <= rank < world or total < 0:
        raise ValueError("bad shard request")
    if total == 0:
        return 0, 0, world
    first = (total - 1 - rank) % world
    n = (total - first + world - 1) // world if total > first else 0
    return first, n, world


class _DevicePtr:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can wrap it
    without a copy."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4",
                                         "data": (ptr, False), "version": 3, "strides": None}


def film_tensor(ctx):
    """torch view (no copy) of the context's device film: 4 floats per pixel, rgb sum + count."""
    ptr, n = ctx.film_device()
    return torch.as_tensor(_DevicePtr(ptr, n), device="cuda:%d" % ctx.device)


def reduce_film(film, dst=0, group=None):
    """Sums the films of all ranks into rank `dst` (in place).  Counts add like colours, so
    the resolve after the reduce divides by the total sample count."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return film


def render_sharded(ctx, cam, total_samples, seed, rank=0, world=1, flags=0, group=None):
    """One image of `total_samples` per pixel over `world` ranks: every rank traces its shard
    into its own film (unresolved), the films are reduced to rank 0, rank 0 resolves."""
    from . import abi
    first, n, stride = shard_samples(total_samples, rank, world)
    ctx.reset()
    ctx.trace(cam, n, first_sample=first, stride=stride, seed=seed,
              flags=flags | abi.FS_TRACE_NO_RESOLVE | abi.FS_TRACE_ASYNC)
    if world > 1:
        reduce_film(film_tensor(ctx), 0, group)
    if rank == 0:
        ctx.resolve()
