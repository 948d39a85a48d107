"""Multi-GPU layout of the fleet: one process per GPU, aircraft sharded by GLOBAL id, no data-path collective.

No aircraft reads another aircraft's state anywhere on the path, so the only exchange is the
episode-statistics all-reduce (16 doubles, once per report interval, NCCL over NVLink on GPUs; gloo in the CPU
tests).  The Philox counter of a reset is (seed, global id, episode), so every trajectory is independent of how many
ranks the fleet is cut into -- tests/test_gpu_sharding.py checks that bit for bit.
"""
from __future__ import annotations

import torch


def shard_range(n_total: int, rank: int, world: int):
    """Contiguous slice [first, first + count) of the global ids owned by `rank`; the first n_total % world ranks get one more."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(int(n_total), int(world))
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def init_comm(ctx, group=None) -> bool:
    """Give `ctx` (an ``_cabi.Context``) the library's own NCCL communicator over the ranks of the (initialised)
    ``torch.distributed`` group: rank 0 draws the unique id through the C ABI, torch.distributed is only the plumbing that
    carries its 128 bytes, every rank calls ``acb200_comm_init``.  Returns False (and leaves the context without a
    communicator) when there is one rank, no process group, or a CPU-only (gloo) world."""
    import ctypes as C
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) < 2 or dist.get_backend(group) != "nccl":
        return False
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    buf = (C.c_ubyte * 128)()
    if rank == 0:
        ctx.check(ctx.lib.acb200_comm_unique_id(ctx.handle, buf))
    t = torch.tensor(list(buf), dtype=torch.uint8, device=torch.device("cuda", ctx.device))
    dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    raw = bytes(t.cpu().tolist())
    ctx.check(ctx.lib.acb200_comm_init(ctx.handle, world, rank, raw))
    return True


def comm_info(ctx) -> dict:
    import ctypes as C
    w, r, v = C.c_int(), C.c_int(), C.c_int()
    ctx.check(ctx.lib.acb200_comm_info(ctx.handle, C.byref(w), C.byref(r), C.byref(v)))
    return {"world": w.value, "rank": r.value, "nccl_version": v.value}


def allreduce_episode_stats(stats: torch.Tensor, group=None, ctx=None) -> torch.Tensor:
    """Sum of the per-rank statistics vectors (envs.STATS_NAMES order); returns a new tensor, `stats` keeps accumulating.
    With a context that owns a communicator (``init_comm``) this is ``acb200_episode_stats`` -- ncclAllReduce issued by the
    library on torch's current stream; otherwise torch.distributed (gloo in the CPU tests, or no process group: a copy)."""
    import ctypes as C
    import torch.distributed as dist
    if ctx is not None and stats.is_cuda and comm_info(ctx)["world"] > 0:
        out = torch.empty_like(stats)
        ctx.set_stream(torch.cuda.current_stream(stats.device).cuda_stream)
        ctx.check(ctx.lib.acb200_episode_stats(ctx.handle, C.c_void_p(stats.data_ptr()), C.c_void_p(out.data_ptr()), stats.numel(), 1))
        return out
    out = stats.detach().clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def make_sharded_env(n_total: int, rank: int, world: int, device=None, **kw):
    """This rank's BatchedRateControlEnv over its slice of a fleet of n_total aircraft."""
    from .envs import BatchedRateControlEnv
    first, count = shard_range(n_total, rank, world)
    return BatchedRateControlEnv(count, id_offset=first, device=rank if device is None else device, **kw)


def bind_host_to_gpu(device_index: int) -> list:
    """Pin this process to the CPU cores NVML reports as local to the GPU, so that pinned host buffers allocated
    afterwards are first-touched on the GPU's own NUMA node (the host-buffer step is PCIe-bound, and on an 8-GPU box
    remote-node staging memory is what the ranks end up fighting over).  Returns the core list ([] = left unchanged:
    NVML not importable, or no affinity information)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(int(device_index))
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cores = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        allowed = sorted(set(cores) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return []
