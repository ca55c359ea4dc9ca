"""Control measurement for the host-fed scaling curve: plain concurrent cudaMemcpyAsync between pinned host memory and N GPUs
of one box, one process (torchrun rank) per GPU, nothing of this library involved.  Per rank one 64-frame 1080p step worth of
bytes (398 MB up, and in the two-way modes 100 MB down on a second stream), as one copy or in 24 MB chunks like
b200_hal_resize_warpAffine_batch issues them.  Prints per-rank and aggregate host->device GB/s: the box's host-feed ceiling
that bench.py's `e2e` cannot exceed (its step time is the SLOWEST rank's).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/exp_h2d_multi.py
"""
import os, time
import torch
import torch.distributed as dist

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
up, down, chunk = 64 * 1920 * 1080 * 3, 64 * 960 * 540 * 3, 24 * 1024 * 1024
h_up = torch.empty(up, dtype=torch.uint8).pin_memory(); h_up.fill_(rank)
h_dn = torch.empty(down, dtype=torch.uint8).pin_memory()
d_up = torch.empty(up, dtype=torch.uint8, device="cuda")
d_dn = torch.zeros(down, dtype=torch.uint8, device="cuda")
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()


def one_step(two_way, chunked):
    with torch.cuda.stream(s_up):
        if chunked:
            for o in range(0, up, chunk):
                d_up[o:o + chunk].copy_(h_up[o:o + chunk], non_blocking=True)
        else:
            d_up.copy_(h_up, non_blocking=True)
    if two_way:
        with torch.cuda.stream(s_dn):
            if chunked:
                c4 = chunk // 4
                for o in range(0, down, c4):
                    h_dn[o:o + c4].copy_(d_dn[o:o + c4], non_blocking=True)
            else:
                h_dn.copy_(d_dn, non_blocking=True)


for two_way in (False, True):
    for chunked in (False, True):
        for _ in range(2):
            one_step(two_way, chunked)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        iters = 6
        t0 = time.perf_counter()
        for _ in range(iters):
            one_step(two_way, chunked)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        gbs = torch.tensor([up * iters / dt / 1e9], device="cuda")
        if world > 1:
            allv = [torch.zeros_like(gbs) for _ in range(world)]
            dist.all_gather(allv, gbs)
            vals = [float(v) for v in allv]
        else:
            vals = [float(gbs)]
        if rank == 0:
            slow = min(vals)
            print(f"{world} GPU(s), {'up+down' if two_way else 'up only'}, {'24 MB chunks' if chunked else 'one copy'}: H2D per rank "
                  f"{[round(v, 1) for v in vals]} GB/s, aggregate {sum(vals):.1f}; slowest rank -> {world * 64 / (up / slow / 1e9):.0f} frames/s "
                  f"= {world * 64 * 960 * 540 / (up / slow / 1e9) / 1e6:.0f} Mpix/s ceiling for a step-synchronous job")
if world > 1:
    dist.destroy_process_group()
