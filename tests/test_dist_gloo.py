"""N>1 host logic on CPU: two gloo ranks shard the 8192-frame stream with the C partitioner (contiguous, disjoint,
covering), reduce per-rank times with MAX and units with SUM exactly as bench.py does under NCCL, and -- because the
data path has no collective -- the union of the per-rank oracle outputs equals the single-process output."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, n_frames, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from opencv_b200 import dist as bdist
    from common import oracle, rotation_inverse, INTER_LINEAR, BORDER_CONSTANT
    first, count = bdist.frame_block(n_frames, world, rank)
    # every rank learns every block: they must tile [0, n) in rank order
    blocks = [None] * world
    dist.all_gather_object(blocks, (first, count))
    # per-frame work on the rank's block (CPU stand-in for the GPU pipeline: same arithmetic through the oracle)
    O = oracle()
    M = rotation_inverse(24, 16, 15, 0.9)
    acc = 0
    for f in range(first, first + count):
        src = np.random.default_rng(f).integers(0, 256, (32, 48, 3), dtype=np.uint8)
        mid, out = np.zeros((16, 24, 3), np.uint8), np.zeros((16, 24, 3), np.uint8)
        assert O.resize(src, mid, INTER_LINEAR) == 0
        assert O.warpAffine(mid, out, M, INTER_LINEAR, BORDER_CONSTANT) == 0
        acc += int(out.astype(np.int64).sum()) * (f + 1)
    local_seconds = 0.010 * (rank + 1)                        # rank 1 is the slow one
    total = bdist.sum_over_ranks(acc)
    slowest = bdist.max_over_ranks(local_seconds)
    thr = bdist.job_throughput(units_per_rank_per_step=count, steps=3, local_seconds=local_seconds)
    q.put((rank, blocks, total, slowest, thr))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharding_reductions_and_union():
    world, n_frames, port = 2, 11, 29533 + os.getpid() % 500
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    blocks = res[0][1]
    assert blocks == res[1][1] == [(0, 5), (5, 6)]                     # contiguous-block rule [g*n/G, (g+1)*n/G)
    assert res[0][2] == res[1][2]                                      # SUM agreed on by both ranks
    assert res[0][3] == res[1][3] == pytest.approx(0.020)              # MAX over ranks = the slow rank
    assert res[0][4] == pytest.approx(3 * n_frames / 0.020)            # whole-job units / slowest time
    # single-process run over all frames gives the same checksum: sharding changes nothing
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from common import oracle, rotation_inverse, INTER_LINEAR, BORDER_CONSTANT
    O = oracle()
    M = rotation_inverse(24, 16, 15, 0.9)
    acc = 0
    for f in range(n_frames):
        src = np.random.default_rng(f).integers(0, 256, (32, 48, 3), dtype=np.uint8)
        mid, out = np.zeros((16, 24, 3), np.uint8), np.zeros((16, 24, 3), np.uint8)
        O.resize(src, mid, INTER_LINEAR)
        O.warpAffine(mid, out, M, INTER_LINEAR, BORDER_CONSTANT)
        acc += int(out.astype(np.int64).sum()) * (f + 1)
    assert float(acc) == res[0][2]
