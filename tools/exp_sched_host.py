"""Times b200_sched_run_pipeline_host: pinned 1080p host frames sharded over all visible GPUs."""
import sys, os, time, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import bench
from opencv_b200 import capi
lib = capi.lib
n, w, h, mw, mh = 256, 1920, 1080, 960, 540
src = torch.randint(0, 256, (n, h, w * 3), dtype=torch.uint8).pin_memory()
dst = torch.empty((n, mh, mw * 3), dtype=torch.uint8).pin_memory()
M = bench.pipeline_matrix(); Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
for ndev in range(1, lib.b200_device_count() + 1):
    devs = (C.c_int * ndev)(*range(ndev)); ms = (C.c_float * ndev)()
    best = 1e9
    for rep in range(4):
        t0 = time.perf_counter()
        rc = lib.b200_sched_run_pipeline_host(devs, ndev, 16, src.data_ptr(), h * w * 3, w * 3, w, h, mw, mh, dst.data_ptr(), mh * mw * 3, mw * 3, mw, mh, n, 1, Mc, 1, 0, bv, ms)
        assert rc == 0, capi.last_error()
        best = min(best, time.perf_counter() - t0)
    print(f"{ndev} GPU(s): {n} host frames in {best * 1e3:.1f} ms = {n / best:.0f} frames/s, {n * h * w * 3 / best / 1e9:.1f} GB/s host->device total; per-device ms {[round(x, 1) for x in ms]}")
