"""Times a few single-launch rows of the bench (noise calibration between runs / boxes)."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import numpy as np, torch
from opencv_b200 import capi
lib = capi.lib
st = torch.cuda.current_stream(); sp = C.c_void_p(st.cuda_stream)
def timeit(fn, iters=48):
    for i in range(6): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(iters): fn(i)
    e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
s = torch.randint(0, 256, (16, 2160, 3840 * 3), dtype=torch.uint8, device="cuda")
for (dw, dh, interp, nm) in ((1920, 1080, 1, "LINEAR 2x2 box"), (1280, 720, 3, "AREA 3x3"), (1600, 900, 3, "AREA fractional")):
    d = torch.empty((16, dh, dw * 3), dtype=torch.uint8, device="cuda")
    def fn(i):
        assert lib.b200_resize_dev(16, s[i % 16].data_ptr(), 3840 * 3, 3840, 2160, d[i % 16].data_ptr(), dw * 3, dw, dh, 0.0, 0.0, interp, sp) == 0
    print(f"resize 4K->{dw}x{dh} {nm}: {timeit(fn):.2f} us")
del s
nbb, bw_, bh_ = 32, 640, 640
sb = torch.randint(0, 256, (nbb * 2, 1080, 1920 * 3), dtype=torch.uint8, device="cuda")
db = torch.empty((nbb * 2, 3, bh_, bw_), dtype=torch.float32, device="cuda")
sc_, mn_ = (C.c_double * 4)(1 / 255., 1 / 255., 1 / 255., 1.0), (C.c_double * 4)(104.0, 117.0, 123.0, 0.0)
def fb(i):
    o = i % 2
    assert lib.b200_blobFromImages_dev(16, sb.data_ptr() + o * nbb * 1080 * 1920 * 3, 1080 * 1920 * 3, 1920 * 3, 1920, 1080,
                                       db.data_ptr() + o * nbb * 3 * bh_ * bw_ * 4, bw_, bh_, nbb, sc_, mn_, 1, sp) == 0
print(f"blobFromImages 32x1080p->640x640: {timeit(fb, 30):.1f} us")
