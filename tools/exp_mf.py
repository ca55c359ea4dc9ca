"""Times the multi-frame gather rows: warpPerspective 4K 8UC4 x 8 frames, remap 720p 8UC3 x 32 frames (CV_32FC2 map)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import numpy as np, torch
from opencv_b200 import capi
lib = capi.lib
st = torch.cuda.current_stream(); sp = C.c_void_p(st.cuda_stream); bv = (C.c_double * 4)(0, 0, 0, 0)
def timeit(fn, iters=30):
    for i in range(4): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(iters): fn(i)
    e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
w, h, nb = 3840, 2160, 8
s = torch.randint(0, 256, (16, h, w * 4), dtype=torch.uint8, device="cuda"); d = torch.empty_like(s)
Hi = np.linalg.inv(np.array([[1.12085816, 0.0203087925, 97], [0.0422569112, 1.05413990, -64], [4.77793919e-05, -3.76190421e-06, 1]]))
Hc = (C.c_double * 9)(*Hi.ravel())
fs = h * w * 4
def fp(i):
    o = (i % 2) * nb * fs
    assert lib.b200_warpPerspective_batch_dev(24, s.data_ptr() + o, fs, w * 4, w, h, d.data_ptr() + o, fs, w * 4, w, h, nb, Hc, 1, 1, bv, sp) == 0
print(f"warpPerspective 4K 8UC4 x8: {timeit(fp):.1f} us")
del s, d
w, h, nb = 1280, 720, 32
s = torch.randint(0, 256, (64, h, w * 3), dtype=torch.uint8, device="cuda"); d = torch.empty_like(s)
xs, ys = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
mxy = torch.from_numpy(np.ascontiguousarray(np.stack([xs * 0.98 + 7.3, ys * 0.97 + 4.1], -1))).cuda()
fs = h * w * 3
def fr(i):
    o = (i % 2) * nb * fs
    assert lib.b200_remap_batch_dev(16, s.data_ptr() + o, fs, w * 3, w, h, d.data_ptr() + o, fs, w * 3, w, h, nb, 13, mxy.data_ptr(), w * 8, None, 0, 1, 4, bv, sp) == 0
print(f"remap 720p 8UC3 x32 (CV_32FC2): {timeit(fr, 40):.1f} us")
