import sys, os, ctypes as C, time
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch, numpy as np
from opencv_b200 import capi
import bench
lib = capi.lib
lib.b200_hal_set_min_pixels(0)
SRC_W, SRC_H, MID_W, MID_H, CN = 1920, 1080, 960, 540, 3
hs = torch.randint(0, 255, (SRC_H, SRC_W, CN), dtype=torch.uint8).pin_memory()
hm = torch.empty((MID_H, MID_W, CN), dtype=torch.uint8).pin_memory()
hd = torch.empty((MID_H, MID_W, CN), dtype=torch.uint8).pin_memory()
M = bench.pipeline_matrix(); Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
def r(): assert lib.b200_hal_resize(16, hs.data_ptr(), SRC_W*3, SRC_W, SRC_H, hm.data_ptr(), MID_W*3, MID_W, MID_H, 0.5, 0.5, 1) == 0
def w(): assert lib.b200_hal_warpAffine(16, hm.data_ptr(), MID_W*3, MID_W, MID_H, hd.data_ptr(), MID_W*3, MID_W, MID_H, Mc, 1, 0, bv) == 0
for f, name in ((r, "hal_resize 1080p->540p"), (w, "hal_warpAffine 960x540")):
    for _ in range(5): f()
    t0 = time.perf_counter()
    for _ in range(100): f()
    print(name, f"{(time.perf_counter()-t0)/100*1e6:.1f} us/call")
# raw copies for reference
d = torch.empty((SRC_H, SRC_W, CN), dtype=torch.uint8, device="cuda")
for _ in range(3): d.copy_(hs, non_blocking=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(50): d.copy_(hs, non_blocking=True)
torch.cuda.synchronize(); t = (time.perf_counter()-t0)/50
print(f"H2D 6.2MB pinned: {t*1e6:.1f} us  {6.2208e6/t/1e9:.1f} GB/s")
dm = torch.empty((MID_H, MID_W, CN), dtype=torch.uint8, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(50): hm.copy_(dm, non_blocking=True)
torch.cuda.synchronize(); t = (time.perf_counter()-t0)/50
print(f"D2H 1.5MB pinned: {t*1e6:.1f} us  {1.5552e6/t/1e9:.1f} GB/s")
