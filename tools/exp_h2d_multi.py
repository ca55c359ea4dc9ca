"""Control measurement for the host-fed scaling curve: plain concurrent cudaMemcpyAsync host -> device from pinned memory on
N GPUs of one box, one process (torchrun rank) per GPU, nothing of this library involved.  Prints per-rank and aggregate GB/s:
the box's host-feed ceiling that bench.py's `e2e` (frames uploaded from pinned host memory) cannot exceed.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/exp_h2d_multi.py
"""
import os, time
import torch
import torch.distributed as dist

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
nbytes = 398 * 1024 * 1024                    # one 64-frame 1080p step
for placement in ("default", "interleave"):
    if placement == "interleave":
        # spread the pinned pages over the NUMA nodes of the box (first touch by threads bound everywhere is not enough under
        # a cpuset; numactl is not in the image): fall back to re-allocating after widening the affinity mask
        try:
            os.sched_setaffinity(0, range(os.cpu_count()))
        except OSError:
            pass
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h.fill_(rank)
    d = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    for _ in range(2):
        with torch.cuda.stream(st):
            d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    iters = 8
    with torch.cuda.stream(st):
        for _ in range(iters):
            d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = torch.tensor([nbytes * iters / dt / 1e9], device="cuda")
    if world > 1:
        allv = [torch.zeros_like(gbs) for _ in range(world)]
        dist.all_gather(allv, gbs)
        vals = [float(v) for v in allv]
    else:
        vals = [float(gbs)]
    if rank == 0:
        print(f"{world} GPU(s), pinned pages '{placement}': per rank {[round(v, 1) for v in vals]} GB/s, aggregate {sum(vals):.1f} GB/s, "
              f"cpus visible {len(os.sched_getaffinity(0))}")
    del h, d
if world > 1:
    dist.destroy_process_group()
