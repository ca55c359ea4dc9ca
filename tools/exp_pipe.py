"""Times the batched resize->warpAffine pipeline (env B200_PIPE_FUSED=0: two batched launches, 1: one persistent launch)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import bench
from opencv_b200 import capi
lib = capi.lib
n, w, h, mw, mh, R = int(os.environ.get("NB", 64)), 1920, 1080, 960, 540, 4
st0 = torch.cuda.current_stream(); sp = C.c_void_p(st0.cuda_stream)
s = torch.empty((n * R, h, w * 3), dtype=torch.uint8, device="cuda")
lib.b200_synth_frames_dev(s.data_ptr(), h * w * 3, w * 3, w, h, 3, 0, n * R, 7, sp)
mid = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
d = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
M = bench.pipeline_matrix(); Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
sfs, mfs = h * w * 3, mh * mw * 3
def step(i):
    c = i % R
    assert lib.b200_resize_warpAffine_batch_dev(16, s.data_ptr() + c * n * sfs, sfs, w * 3, w, h, mid.data_ptr(), mfs, mw * 3, mw, mh, d.data_ptr() + c * n * mfs, mfs, mw * 3, mw, mh, n, 1, Mc, 1, 0, bv, sp) == 0, capi.last_error()
l0 = lib.b200_launch_count()
for i in range(10): step(i)
torch.cuda.synchronize()
per_step = (lib.b200_launch_count() - l0) / 10
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st0)
for i in range(200): step(i)
e1.record(st0); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 200 * 1e3
print(f"B200_PIPE_FUSED={os.environ.get('B200_PIPE_FUSED', '-')}: {n} frames/step, {per_step:.0f} launches/step, {t:.1f} us/step, {n / t * 1e6:.0f} frames/s, checksum {int(d.to(torch.int64).sum())}")
