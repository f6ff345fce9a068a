"""Sample sharding over GPUs and the per-frame accumulation reduce.

The reference is single-device; paths never interact, and the only shared state is the float4
accumulation buffer `pixels` (reference src/clcontext.cpp:347).  Each rank therefore renders its
own N paths against a full scene replica with
    seed       = gid + rank * N           (reference: seed = gid, src/wf_reset.cl:61)
    cursor_0   = rank * (w*h) / world     (reference: 0, src/clcontext.cpp:923-927)
    spp budget = spp / world (+1 on the low ranks for the remainder)
and the buffers are summed with ONE all-reduce per frame; `.w` carries the sample count, so the
normalisation in the postprocess stage (src/mk_postprocess.cl:20-24) is unchanged.

torch.distributed is plumbing only: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
import numpy as np


def shard_plan(rank, world_size, num_tasks, num_pixels, spp=None):
    """Decorrelation parameters of one rank: dict(seed_offset, cursor_start, spp).  A per-pixel budget (maxSpp > 0,
    CHECK_SPP) is per context and NOT combined by reduce_frames / allreduce_pixels: give every rank ceil(maxSpp / world)
    (Tracer(world_size=...) and fluctus_headless do), or N ranks deliver N * maxSpp samples per pixel."""
    rank, world_size = int(rank), int(world_size)
    if not (0 <= rank < world_size):
        raise ValueError("rank %d outside world of %d" % (rank, world_size))
    plan = dict(seed_offset=(rank * int(num_tasks)) & 0xffffffff,
                cursor_start=(rank * int(num_pixels)) // world_size, spp=None)
    if spp is not None:
        base, rem = divmod(int(spp), world_size)
        plan["spp"] = base + (1 if rank < rem else 0)
    return plan


def split_rays(n, rank, world_size):
    """Contiguous chunk [first, first+count) of an n-ray batch owned by `rank` (ray-cast microbench:
    scene replicated, no collective; Mrays/s add)."""
    base, rem = divmod(int(n), int(world_size))
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


class DevicePixels:
    """Zero-copy torch view of a context's accumulation buffer through __cuda_array_interface__."""

    def __init__(self, ptr, num_pixels):
        self.__cuda_array_interface__ = {"shape": (int(num_pixels), 4), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def pixels_tensor(ctx, device):
    """torch tensor aliasing FLX_BUF_PIXELS of a CudaContext (no copy)."""
    import torch
    from .context import BUF_PIXELS
    ptr = ctx.devicePtr(BUF_PIXELS)
    if not ptr:
        raise RuntimeError("context has no frame buffers yet (updateParams first)")
    return torch.as_tensor(DevicePixels(ptr, ctx.width * ctx.height), device=device)


def allreduce_pixels(pixels, group=None):
    """Sum the accumulation buffers of all ranks in place.  `pixels` is a torch tensor (device tensor
    on the NCCL path, CPU tensor under gloo) or a numpy array (reduced through a CPU tensor).
    Must be ordered after the render on the same stream: the caller makes the context's stream the
    current torch stream (see bench.py), which is also the stream NCCL orders against."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return pixels
    if isinstance(pixels, np.ndarray):
        t = torch.from_numpy(pixels)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return pixels
    dist.all_reduce(pixels, op=dist.ReduceOp.SUM, group=group)
    return pixels


def reduce_pixels(pixels, dst=0, group=None):
    """Sum the accumulation buffers of all ranks INTO rank `dst` (one NCCL reduce): for frames that only the root
    tonemaps and reads back.  The other ranks' buffers keep their own samples and they may start the next frame at once."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return pixels
    dist.reduce(pixels, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return pixels


def reduce_frames(contexts):
    """Several CudaContexts in THIS process, one per GPU (the layout the C++ host uses for `--gpus N`): sum their
    accumulation buffers with the peer-memory kernel of csrc/flx_multi.cu (flx_reduce_frames) — the total ends up
    in contexts[0], the others are zeroed.  No NCCL involved; for one process per GPU use allreduce_pixels."""
    import ctypes as C
    from ._lib import FluctusError
    root = contexts[0]
    handles = (C.c_void_p * len(contexts))(*[c.h for c in contexts])
    rc = root.L.flx_reduce_frames(handles, len(contexts))
    if rc != 0:
        msgs = [c.L.flx_last_error(c.h).decode() for c in contexts] + [root.L.flx_last_error(None).decode()]
        raise FluctusError("; ".join(m for m in msgs if m) or "flx_reduce_frames failed (%d)" % rc)
    return root


def normalise(pixels):
    """rgb / sample count — what CLContext::saveImage does for .hdr output (clcontext.cpp:455-470)."""
    px = np.asarray(pixels, np.float32).reshape(-1, 4)
    return px[:, :3] / np.maximum(px[:, 3:4], 1.0)
