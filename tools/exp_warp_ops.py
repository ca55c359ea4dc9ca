"""Times warpPerspective 4K 8UC4 / 8UC3 and remap 720p 8UC3 (generic gather kernels) on rotating buffers."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import numpy as np, torch
from opencv_b200 import capi
lib = capi.lib
st0 = torch.cuda.current_stream(); sp = C.c_void_p(st0.cuda_stream)
bv = (C.c_double * 4)(0, 0, 0, 0)
def timeit(fn, n, iters=100):
    for i in range(10): fn(i % n)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st0)
    for i in range(iters): fn(i % n)
    e1.record(st0); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
Hi = np.linalg.inv(np.array([[1.12085816, 0.0203087925, 97], [0.0422569112, 1.05413990, -64], [4.77793919e-05, -3.76190421e-06, 1]]))
Hc = (C.c_double * 9)(*Hi.ravel())
for cn in (4, 3, 1):
    n = 12
    s = torch.empty((n, 2160, 3840 * cn), dtype=torch.uint8, device="cuda"); d = torch.empty_like(s)
    lib.b200_synth_frames_dev(s.data_ptr(), 2160 * 3840 * cn, 3840 * cn, 3840, 2160, cn, 0, n, 7, sp)
    def fn(i): assert lib.b200_warpPerspective_dev((cn - 1) << 3, s[i].data_ptr(), 3840 * cn, 3840, 2160, d[i].data_ptr(), 3840 * cn, 3840, 2160, Hc, 1, 1, bv, sp) == 0
    print(f"warpPerspective 4K 8UC{cn} LINEAR REPLICATE {timeit(fn, n):7.2f} us  checksum {int(d[0].to(torch.int64).sum())}")
    del s, d
w, h = 1280, 720
xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
X, Y = (xs - 639.5) / 600, (ys - 359.5) / 600
r = np.sqrt(X * X + Y * Y); th = np.arctan(r)
sc = np.where(r < 1e-8, 1.0, th * (1 - 0.05 * th ** 2 + 0.01 * th ** 4) / np.maximum(r, 1e-300))
mxy = torch.from_numpy(np.ascontiguousarray(np.stack([(600 * X * sc + 639.5).astype(np.float32), (600 * Y * sc + 359.5).astype(np.float32)], -1))).cuda()
for cn in (3, 4):
    n = 64
    s = torch.empty((n, h, w * cn), dtype=torch.uint8, device="cuda"); d = torch.empty_like(s)
    lib.b200_synth_frames_dev(s.data_ptr(), h * w * cn, w * cn, w, h, cn, 0, n, 7, sp)
    def fn(i): assert lib.b200_remap32fc2_dev((cn - 1) << 3, s[i].data_ptr(), w * cn, w, h, d[i].data_ptr(), w * cn, w, h, mxy.data_ptr(), w * 8, 1, 4, bv, sp) == 0
    print(f"remap 720p 8UC{cn} 32FC2 map REFLECT_101        {timeit(fn, n, 200):7.2f} us  checksum {int(d[0].to(torch.int64).sum())}")
    def fnb(i): assert lib.b200_remap_batch_dev((cn - 1) << 3, s.data_ptr() + (i % 2) * 32 * h * w * cn, h * w * cn, w * cn, w, h, d.data_ptr() + (i % 2) * 32 * h * w * cn, h * w * cn, w * cn, w, h, 32, 13, mxy.data_ptr(), w * 8, None, 0, 1, 4, bv, sp) == 0
    print(f"remap 720p 8UC{cn} x32 frames per launch         {timeit(fnb, 2, 40):7.2f} us")
