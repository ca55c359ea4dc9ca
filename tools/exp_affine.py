"""warpAffine INTER_LINEAR 8UC3 / 8UC4 / 8UC1: times the fast path (1 frame and 8 frames per launch, 4K and 1080p) and
checks it against the generic kernel (b200_set_force_generic) on the same buffers."""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import numpy as np, torch
import bench
from opencv_b200 import capi
lib = capi.lib
st0 = torch.cuda.current_stream(); sp = C.c_void_p(st0.cuda_stream)
def timeit(fn, n, iters=60):
    for i in range(6): fn(i % n)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st0)
    for i in range(iters): fn(i % n)
    e1.record(st0); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
PEAK = 6546.2
cases = [(3840, 2160), (1920, 1080)]
borders = [(0, (0, 0, 0, 0)), (0, (7, 9, 11, 13)), (4, (0, 0, 0, 0)), (1, (0,) * 4), (3, (0,) * 4)]
for (w, h) in cases:
    M = bench.affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2); Mc = (C.c_double * 6)(*M.ravel())
    for cn in (3, 4, 1):
        nb = 8
        sets = 3 if w > 2000 else 8
        s = torch.empty((sets * nb, h, w * cn), dtype=torch.uint8, device="cuda"); d = torch.empty_like(s)
        lib.b200_synth_frames_dev(s.data_ptr(), h * w * cn, w * cn, w, h, cn, 0, sets * nb, 7, sp)
        fs = h * w * cn
        typ = (cn - 1) << 3
        ok = True
        for border, bvv in borders:
            bv = (C.c_double * 4)(*bvv)
            g = torch.empty((h, w * cn), dtype=torch.uint8, device="cuda"); f = torch.empty_like(g)
            lib.b200_set_force_generic(1)
            assert lib.b200_warpAffine_dev(typ, s.data_ptr(), w * cn, w, h, g.data_ptr(), w * cn, w, h, Mc, 1, border, bv, sp) == 0
            lib.b200_set_force_generic(0)
            assert lib.b200_warpAffine_dev(typ, s.data_ptr(), w * cn, w, h, f.data_ptr(), w * cn, w, h, Mc, 1, border, bv, sp) == 0, capi.last_error()
            torch.cuda.synchronize()
            nbad = int((g != f).sum())
            if nbad: ok = False; print(f"  MISMATCH {w}x{h} C{cn} border {border} bv {bvv}: {nbad} bytes differ")
        bv = (C.c_double * 4)(0, 0, 0, 0)
        def one(i): assert lib.b200_warpAffine_dev(typ, s.data_ptr() + i * fs, w * cn, w, h, d.data_ptr() + i * fs, w * cn, w, h, Mc, 1, 0, bv, sp) == 0
        def batch(i): assert lib.b200_warpAffine_batch_dev(typ, s.data_ptr() + i * nb * fs, fs, w * cn, w, h, d.data_ptr() + i * nb * fs, fs, w * cn, w, h, nb, Mc, 1, 0, bv, sp) == 0
        t1 = timeit(one, sets * nb); t8 = timeit(batch, sets, 30)
        gb1 = 2 * fs / t1 / 1e3; gb8 = 2 * fs * nb / t8 / 1e3
        print(f"warpAffine {w}x{h} 8UC{cn}: 1 frame {t1:7.2f} us {gb1:6.0f} GB/s ({gb1 / PEAK:.3f})   {nb} frames {t8:7.2f} us {gb8:6.0f} GB/s ({gb8 / PEAK:.3f})   parity vs generic: {'ok' if ok else 'FAIL'}")
        del s, d
