"""Launches each BASELINE op a few times on rotating buffers (for ncu captures: short, one process, one GPU)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import numpy as np, torch
import bench
from opencv_b200 import capi
lib = capi.lib
st = torch.cuda.current_stream(); sp = C.c_void_p(st.cuda_stream)
bv = (C.c_double * 4)(0, 0, 0, 0)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
def synth(n, h, w, cn):
    t = torch.empty((n, h, w * cn), dtype=torch.uint8, device="cuda")
    assert lib.b200_synth_frames_dev(t.data_ptr(), h * w * cn, w * cn, w, h, cn, 0, n, 7, sp) == 0
    return t
if which in ("all", "affine4k"):
    w, h, n = 3840, 2160, 4
    s, d = synth(n, h, w, 3), torch.empty((n, h, w * 3), dtype=torch.uint8, device="cuda")
    M = bench.affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2); Mc = (C.c_double * 6)(*M.ravel())
    for i in range(reps):
        assert lib.b200_warpAffine_dev(16, s[i % n].data_ptr(), w * 3, w, h, d[i % n].data_ptr(), w * 3, w, h, Mc, 1, 0, bv, sp) == 0
if which in ("all", "resize"):
    s = synth(4, 2160, 3840, 3)
    for (dw, dh, interp) in ((1920, 1080, 1), (1600, 900, 3), (1280, 720, 3), (1600, 900, 1)):
        d = torch.empty((4, dh, dw * 3), dtype=torch.uint8, device="cuda")
        for i in range(reps):
            assert lib.b200_resize_dev(16, s[i % 4].data_ptr(), 3840 * 3, 3840, 2160, d[i % 4].data_ptr(), dw * 3, dw, dh, 0.0, 0.0, interp, sp) == 0
    s = torch.rand((4, 1080, 1920), dtype=torch.float32, device="cuda"); d = torch.empty((4, 2160, 3840), dtype=torch.float32, device="cuda")
    for i in range(reps):
        assert lib.b200_resize_dev(5, s[i % 4].data_ptr(), 1920 * 4, 1920, 1080, d[i % 4].data_ptr(), 3840 * 4, 3840, 2160, 0.0, 0.0, 1, sp) == 0
if which in ("all", "persp"):
    s, d = synth(4, 2160, 3840, 4), torch.empty((4, 2160, 3840 * 4), dtype=torch.uint8, device="cuda")
    Hi = np.linalg.inv(np.array([[1.12085816, 0.0203087925, 97], [0.0422569112, 1.05413990, -64], [4.77793919e-05, -3.76190421e-06, 1]]))
    Hc = (C.c_double * 9)(*Hi.ravel())
    for i in range(reps):
        assert lib.b200_warpPerspective_dev(24, s[i % 4].data_ptr(), 3840 * 4, 3840, 2160, d[i % 4].data_ptr(), 3840 * 4, 3840, 2160, Hc, 1, 1, bv, sp) == 0
if which in ("all", "remap"):
    w, h = 1280, 720
    s, d = synth(8, h, w, 3), torch.empty((8, h, w * 3), dtype=torch.uint8, device="cuda")
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    X, Y = (xs - 639.5) / 600, (ys - 359.5) / 600
    r = np.sqrt(X * X + Y * Y); th = np.arctan(r)
    sc = np.where(r < 1e-8, 1.0, th * (1 - 0.05 * th ** 2 + 0.01 * th ** 4) / np.maximum(r, 1e-300))
    mx, my = (600 * X * sc + 639.5).astype(np.float32), (600 * Y * sc + 359.5).astype(np.float32)
    t_mxy = torch.from_numpy(np.ascontiguousarray(np.stack([mx, my], -1))).cuda()
    for i in range(reps):
        assert lib.b200_remap32fc2_dev(16, s[i % 8].data_ptr(), w * 3, w, h, d[i % 8].data_ptr(), w * 3, w, h, t_mxy.data_ptr(), w * 8, 1, 4, bv, sp) == 0
torch.cuda.synchronize()
print("done")
