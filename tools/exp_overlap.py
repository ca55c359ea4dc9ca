"""Experiment: does running the DRAM-bound box kernel of sub-batch k+1 next to the issue-bound warp kernel of sub-batch k
(two streams, event dependencies) beat running them back to back?"""
import sys, os, ctypes as C
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import torch, numpy as np
import bench
from opencv_b200 import capi
lib = capi.lib
n, w, h, mw, mh = 64, 1920, 1080, 960, 540
R = 4
st0 = torch.cuda.current_stream()
s = torch.empty((n * R, h, w * 3), dtype=torch.uint8, device="cuda")
lib.b200_synth_frames_dev(s.data_ptr(), h * w * 3, w * 3, w, h, 3, 0, n * R, 7, C.c_void_p(st0.cuda_stream))
mid = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
d = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
M = bench.pipeline_matrix(); Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
sfs, mfs = h * w * 3, mh * mw * 3
S1, S2 = torch.cuda.Stream(), torch.cuda.Stream()
def box(stream, c, f0, nf):
    assert lib.b200_resize_batch_dev(16, s.data_ptr() + (c * n + f0) * sfs, sfs, w * 3, w, h, mid.data_ptr() + f0 * mfs, mfs, mw * 3, mw, mh, nf, 0.0, 0.0, 1, C.c_void_p(stream.cuda_stream)) == 0
def warp(stream, c, f0, nf):
    assert lib.b200_warpAffine_batch_dev(16, mid.data_ptr() + f0 * mfs, mfs, mw * 3, mw, mh, d.data_ptr() + (c * n + f0) * mfs, mfs, mw * 3, mw, mh, nf, Mc, 1, 0, bv, C.c_void_p(stream.cuda_stream)) == 0
def step_seq(i, sub):
    c = i % R
    for f0 in range(0, n, sub):
        box(st0, c, f0, sub); warp(st0, c, f0, sub)
def step_overlap(i, sub):
    c = i % R
    # S1 runs the warps, S2 the boxes; warp(k) needs box(k); box(k+1) may overlap warp(k).  mid is per-frame so no WAR hazard inside a step
    S1.wait_stream(st0); S2.wait_stream(st0)
    evs = []
    for f0 in range(0, n, sub):
        box(S2, c, f0, sub)
        e = torch.cuda.Event(); e.record(S2); evs.append(e)
    for k, f0 in enumerate(range(0, n, sub)):
        S1.wait_event(evs[k]); warp(S1, c, f0, sub)
    st0.wait_stream(S1); st0.wait_stream(S2)
def step_overlap2(i, sub):
    c = i % R
    S1.wait_stream(st0); S2.wait_stream(st0)
    box(S2, c, 0, sub)
    e = torch.cuda.Event(); e.record(S2)
    for k, f0 in enumerate(range(0, n, sub)):
        S1.wait_event(e)
        warp(S1, c, f0, sub)            # launched first: takes its 6 CTAs/SM; the next box squeezes into what is left
        if f0 + sub < n:
            box(S2, c, f0 + sub, sub)
            e = torch.cuda.Event(); e.record(S2)
    st0.wait_stream(S1); st0.wait_stream(S2)
def timeit(fn, sub, iters=100):
    for i in range(5): fn(i, sub)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st0)
    for i in range(iters): fn(i, sub)
    e1.record(st0); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
for sub in (64, 32, 16, 8):
    print(f"sub {sub}: sequential {timeit(step_seq, sub):.1f} us   two-stream {timeit(step_overlap, sub):.1f} us   interleaved {timeit(step_overlap2, sub):.1f} us")
