"""Multi-process multi-GPU rendering (one process per GPU, torch.distributed): the scene is
replicated, sample indices — or, for fewer samples than ranks, interleaved pixel tiles — are
sharded across ranks, and the per-rank accumulation films (rgb sum + sample count per pixel) are
combined with one reduce over NCCL / NVLink (gloo on CPU in the tests).  torch.distributed is
plumbing only.  (Single-process callers use fs_create_multi / base::SetTraceDevices instead:
host threads + a peer-memory combine kernel, no NCCL.)

Stream order.  A context launches on its own non-blocking stream; NCCL collectives run on
NCCL's stream and synchronise with torch's *current* stream only.  Every collective here is
therefore issued with the context's stream made current (torch.cuda.ExternalStream), so the
reduce waits for the tracer's kernels and the resolve / read-back / next reset that follow on
the context's stream wait for the reduce."""
import contextlib

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


def tile_owner(x, y, world):
    """Rank that renders pixel (x, y) under tile sharding: 16x8-pixel tiles, diagonal
    interleave (tx + ty) mod world (csrc/fs2_wavefront.cuh wf_slot_pixel)."""
    return ((x >> 4) + (y >> 3)) % world


class _DevicePtr:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can wrap it
    without a copy."""

    def __init__(self, ptr, n, typestr="<f4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr,
                                         "data": (ptr, False), "version": 3, "strides": None}


def film_tensor(ctx):
    """torch view (no copy) of the context's device film: 4 floats per pixel, rgb sum + count."""
    ptr, n = ctx.film_device()
    return torch.as_tensor(_DevicePtr(ptr, n), device="cuda:%d" % ctx.device)


def aux_tensors(ctx):
    """torch views of the aux planes: float4 (normal, depth) and material id per pixel, both as
    int32 words.  Under tile sharding every pixel's record lives on exactly one rank and is zero
    elsewhere, so an integer sum gathers it with its bit pattern intact (a float sum would turn a
    -0.0 normal component into +0.0)."""
    aux, mat, n = ctx.film_device_aux()
    dev = "cuda:%d" % ctx.device
    return (torch.as_tensor(_DevicePtr(aux, 4 * n, "<i4"), device=dev),
            torch.as_tensor(_DevicePtr(mat, n, "<i4"), device=dev))


def context_stream(ctx):
    """Context manager that makes the context's CUDA stream torch's current stream."""
    handle = ctx.stream()
    if not handle or not torch.cuda.is_available():
        return contextlib.nullcontext()
    return torch.cuda.stream(torch.cuda.ExternalStream(handle, device=torch.device("cuda", ctx.device)))


def reduce_film(film, dst=0, group=None, ctx=None):
    """Sums the films of all ranks into rank `dst` (in place).  Counts add like colours, so
    the resolve after the reduce divides by the total sample count.  With `ctx`, the reduce is
    ordered on the context's stream (see the module docstring)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        with (context_stream(ctx) if ctx is not None else contextlib.nullcontext()):
            dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return film


def render_sharded(ctx, cam, total_samples, seed, rank=0, world=1, flags=0, group=None, tiles=None):
    """One image of `total_samples` per pixel over `world` ranks: every rank traces its shard
    into its own film (unresolved), the films are reduced to rank 0, rank 0 resolves.
    tiles=None: sample sharding unless there are fewer samples than ranks."""
    from . import abi
    if tiles is None:
        tiles = 0 < total_samples < world
    ctx.reset()
    if tiles and world > 1:
        ctx.set_tile_shard(rank, world)
        try:
            ctx.trace(cam, total_samples, seed=seed, flags=flags | abi.FS_TRACE_NO_RESOLVE | abi.FS_TRACE_ASYNC)
        finally:
            ctx.set_tile_shard(0, 1)
    else:
        first, n, stride = shard_samples(total_samples, rank, world)
        ctx.trace(cam, n, first_sample=first, stride=stride, seed=seed,
                  flags=flags | abi.FS_TRACE_NO_RESOLVE | abi.FS_TRACE_ASYNC)
    if world > 1:
        reduce_film(film_tensor(ctx), 0, group, ctx)
        if tiles:  # every pixel's aux record lives on exactly one rank: a sum gathers it
            aux, mat = aux_tensors(ctx)
            reduce_film(aux, 0, group, ctx)
            reduce_film(mat, 0, group, ctx)
    if rank == 0:
        ctx.resolve()
