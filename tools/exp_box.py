import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
from opencv_b200 import capi
lib = capi.lib
n, R, w, h, mw, mh = int(os.environ.get("NB", 64)), 4, 1920, 1080, 960, 540
s = torch.empty((n * R, h, w * 3), dtype=torch.uint8, device="cuda")
lib.b200_synth_frames_dev(s.data_ptr(), h * w * 3, w * 3, w, h, 3, 0, n * R, 7, None)
mid = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
st = torch.cuda.current_stream(); sp = C.c_void_p(st.cuda_stream)
def fn(i):
    assert lib.b200_resize_batch_dev(16, s.data_ptr() + i * n * h * w * 3, h * w * 3, w * 3, w, h, mid.data_ptr() + i * n * mh * mw * 3, mh * mw * 3, mw * 3, mw, mh, n, 0.0, 0.0, 1, sp) == 0
for i in range(8): fn(i % R)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for i in range(200): fn(i % R)
e1.record(st); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 200 * 1e-3
print(f"variant {os.environ.get('B200_BOX_VARIANT','0')} frames {n}: {t*1e6:.1f} us  {n*(w*h*3+mw*mh*3)/t/1e9:.0f} GB/s")
a = torch.empty(1 << 29, dtype=torch.uint8, device="cuda"); b = torch.empty_like(a)
for _ in range(3): b.copy_(a)
torch.cuda.synchronize(); e0.record(st)
for _ in range(20): b.copy_(a)
e1.record(st); torch.cuda.synchronize()
print(f"torch copy 512MiB: {2*(1<<29)/(e0.elapsed_time(e1)/20*1e-3)/1e9:.0f} GB/s")
