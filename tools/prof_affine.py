"""A few 8-frame warpAffine 4K launches (8UC3 then 8UC4) for an ncu capture."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch
import bench
from opencv_b200 import capi
lib = capi.lib
st0 = torch.cuda.current_stream(); sp = C.c_void_p(st0.cuda_stream)
w, h, nb = 3840, 2160, 8
M = bench.affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2); Mc = (C.c_double * 6)(*M.ravel())
bv = (C.c_double * 4)(0, 0, 0, 0)
for cn in (3, 4):
    s = torch.empty((2 * nb, h, w * cn), dtype=torch.uint8, device="cuda"); d = torch.empty_like(s)
    lib.b200_synth_frames_dev(s.data_ptr(), h * w * cn, w * cn, w, h, cn, 0, 2 * nb, 7, sp)
    fs = h * w * cn
    for i in range(3):
        assert lib.b200_warpAffine_batch_dev((cn - 1) << 3, s.data_ptr() + (i % 2) * nb * fs, fs, w * cn, w, h, d.data_ptr() + (i % 2) * nb * fs, fs, w * cn, w, h, nb, Mc, 1, 0, bv, sp) == 0
    torch.cuda.synchronize()
print("done")
