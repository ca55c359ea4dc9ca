"""Experiment: TMA warpAffine tile shape / shared-memory pitch sweep (env B200_EXP_SHAPE="tw,th,pitch")."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import bench
from opencv_b200 import capi
lib = capi.lib
n, mw, mh = int(os.environ.get('EXP_N', 64)), int(os.environ.get('EXP_W', 960)), int(os.environ.get('EXP_H', 540))
R = max(6, int(1.5e9 // (n * mw * mh * 3 * 2)))
R = min(R, 64)
st0 = torch.cuda.current_stream(); sp = C.c_void_p(st0.cuda_stream)
s = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
lib.b200_synth_frames_dev(s.data_ptr(), mh * mw * 3, mw * 3, mw, mh, 3, 0, n * R, 7, sp)
d = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
M = bench.affine_inverse(mw, mh, 15.0, 0.9, mw / 2.0, mh / 2.0); Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
fs = mh * mw * 3
def step(i):
    c = i % R
    assert lib.b200_warpAffine_batch_dev(16, s.data_ptr() + c * n * fs, fs, mw * 3, mw, mh, d.data_ptr() + c * n * fs, fs, mw * 3, mw, mh, n, Mc, 1, 0, bv, sp) == 0
for i in range(12): step(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st0)
for i in range(300): step(i)
e1.record(st0); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 300 * 1e3
print(f"B200_EXP_SHAPE={os.environ.get('B200_EXP_SHAPE', '-'):12s} {n}x{mw}x{mh} warp {t:7.1f} us  checksum {int(d[:n].to(torch.int64).sum())}")
