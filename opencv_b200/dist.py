"""Host-side sharding plumbing for the multi-GPU frame stream (one process per GPU, no data-path collective).

`frame_block` is the contiguous-block rule of the C scheduler (b200_sched_partition); `max_over_ranks` /
`sum_over_ranks` are the only communication the path has: reducing the measured time (max) and the output checksum (sum)
for reporting.  They work on any torch.distributed backend (NCCL on the GPU box, gloo in the CPU tests)."""
import ctypes as C


def frame_block(n_frames, world, rank):
    """(first, count) of rank's contiguous frame block: [rank*n/world, (rank+1)*n/world)."""
    from . import capi
    f, c = C.c_int(), C.c_int()
    capi.lib.b200_sched_partition(int(n_frames), int(world), int(rank), C.byref(f), C.byref(c))
    return f.value, c.value


def _reduce(value, op_name, device=None):
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device or ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=getattr(dist.ReduceOp, op_name))
    return float(t.item())


def max_over_ranks(value, device=None):
    return _reduce(float(value), "MAX", device)


def sum_over_ranks(value, device=None):
    return _reduce(float(value), "SUM", device)


def job_throughput(units_per_rank_per_step, steps, local_seconds, device=None):
    """whole-job units/s: all ranks' units over the slowest rank's time (the bench contract's max-over-ranks rule)"""
    import torch.distributed as dist
    world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
    total_units = sum_over_ranks(units_per_rank_per_step * steps, device) if world > 1 else units_per_rank_per_step * steps
    return total_units / max_over_ranks(local_seconds, device)
