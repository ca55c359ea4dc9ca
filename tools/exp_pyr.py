"""Times b200_pyrDown_dev (4K 8UC3 / 32FC1 / 16UC4 / 8UC1 -> half size) and b200_pyrUp_dev (1080p -> 4K), 8 rotating frames."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
from opencv_b200 import capi
lib = capi.lib
st = torch.cuda.current_stream(); sp = C.c_void_p(st.cuda_stream)
for nm, typ, tdt, cn, esz1 in (("8UC3", 16, torch.uint8, 3, 1), ("32FC1", 5, torch.float32, 1, 4), ("16UC4", 26, torch.int16, 4, 2), ("8UC1", 0, torch.uint8, 1, 1)):
    w, h, n = 3840, 2160, 8
    s = (torch.rand((n, h, w * cn), device="cuda") * 200).to(tdt)
    d = torch.empty((n, h // 2, (w // 2) * cn), dtype=tdt, device="cuda")
    esz = esz1 * cn
    def fn(i):
        assert lib.b200_pyrDown_dev(typ, s[i % n].data_ptr(), w * esz, w, h, d[i % n].data_ptr(), (w // 2) * esz, w // 2, h // 2, 4, sp) == 0
    for i in range(8): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(80): fn(i)
    e1.record(st); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 80 * 1e3
    gb = (w * h * esz + (w // 2) * (h // 2) * esz) / t / 1e3
    print(f"pyrDown 4K {nm}: {t:.1f} us, {gb:.0f} GB/s, checksum {int(d.to(torch.float64).sum())}")

for nm, typ, tdt, cn, esz1 in (("8UC3", 16, torch.uint8, 3, 1), ("32FC1", 5, torch.float32, 1, 4), ("16UC4", 26, torch.int16, 4, 2), ("8UC1", 0, torch.uint8, 1, 1)):
    w, h, n = 1920, 1080, 8
    s = (torch.rand((n, h, w * cn), device="cuda") * 200).to(tdt)
    d = torch.empty((n, h * 2, w * 2 * cn), dtype=tdt, device="cuda")
    esz = esz1 * cn
    def fn(i):
        assert lib.b200_pyrUp_dev(typ, s[i % n].data_ptr(), w * esz, w, h, d[i % n].data_ptr(), w * 2 * esz, w * 2, h * 2, 4, sp) == 0
    for i in range(8): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for i in range(80): fn(i)
    e1.record(st); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 80 * 1e3
    gb = (w * h * esz + 4 * w * h * esz) / t / 1e3
    print(f"pyrUp 1080p -> 4K {nm}: {t:.1f} us, {gb:.0f} GB/s, checksum {int(d.to(torch.float64).sum())}")
