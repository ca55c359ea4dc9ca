"""Times the generic resize kernels (linear / area) on rotating 4K buffers."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
from opencv_b200 import capi
lib = capi.lib
st0 = torch.cuda.current_stream(); sp = C.c_void_p(st0.cuda_stream)
def timeit(fn, n=16, iters=200):
    for i in range(20): fn(i % n)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st0)
    for i in range(iters): fn(i % n)
    e1.record(st0); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
n = 16
s = torch.empty((n, 2160, 3840 * 3), dtype=torch.uint8, device="cuda")
lib.b200_synth_frames_dev(s.data_ptr(), 2160 * 3840 * 3, 3840 * 3, 3840, 2160, 3, 0, n, 7, sp)
for (dw, dh, interp, nm) in ((1600, 900, 1, "LINEAR 4K->1600x900 8UC3"), (1600, 900, 3, "AREA 4K->1600x900 8UC3"), (2560, 1440, 1, "LINEAR 4K->1440p 8UC3"), (1600, 900, 2, "CUBIC 4K->1600x900 8UC3")):
    d = torch.empty((n, dh, dw * 3), dtype=torch.uint8, device="cuda")
    def fn(i): assert lib.b200_resize_dev(16, s[i].data_ptr(), 3840 * 3, 3840, 2160, d[i].data_ptr(), dw * 3, dw, dh, 0.0, 0.0, interp, sp) == 0
    print(f"{nm:32s} {timeit(fn):7.2f} us  checksum {int(d[0].to(torch.int64).sum())}")
sf = torch.rand((n, 1080, 1920), dtype=torch.float32, device="cuda"); df = torch.empty((n, 2160, 3840), dtype=torch.float32, device="cuda")
def fn(i): assert lib.b200_resize_dev(5, sf[i].data_ptr(), 1920 * 4, 1920, 1080, df[i].data_ptr(), 3840 * 4, 3840, 2160, 0.0, 0.0, 1, sp) == 0
print(f"{'LINEAR 1080p->4K 32FC1':32s} {timeit(fn):7.2f} us")
s4 = torch.empty((n, 1080, 1920 * 4), dtype=torch.uint8, device="cuda"); d4 = torch.empty((n, 1440, 2560 * 4), dtype=torch.uint8, device="cuda")
lib.b200_synth_frames_dev(s4.data_ptr(), 1080 * 1920 * 4, 1920 * 4, 1920, 1080, 4, 0, n, 7, sp)
def fn(i): assert lib.b200_resize_dev(24, s4[i].data_ptr(), 1920 * 4, 1920, 1080, d4[i].data_ptr(), 2560 * 4, 2560, 1440, 0.0, 0.0, 1, sp) == 0
print(f"{'LINEAR 1080p->1440p 8UC4':32s} {timeit(fn):7.2f} us")
