"""b200_resize_warpAffine_noalloc_dev -- the pipeline entry for callers that may not allocate GPU memory
(a running capture loop, a plugin without a memory budget, code where cudaMalloc would stall the device).
Its intermediate frame lives in a fixed array inside the library.  Proved here: it returns exactly what the ordinary
entry point (caller-provided intermediate buffer) returns -- twice in a row with different frames, because the
array is shared and keeps the previous call's contents --, both equal the oracle's per-frame cv::resize +
cv::warpAffine, and a call leaves the device's free memory where it was.
This file sorts last on purpose: a `-x` run has already executed every other test when it gets here.
It also passes alone in a fresh process (it then performs the per-device set-up itself)."""
import ctypes as C

import numpy as np
import pytest

from common import oracle, assert_same, rotation_inverse, INTER_LINEAR, INTER_AREA, BORDER_CONSTANT, BORDER_REFLECT_101

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _run_both(lib, src, sw, sh, mw, mh, n, M, resize_interp, border, bv):
    """returns (no-alloc output, reference-entry output) for a batch of n 8UC3 frames held in `src`"""
    s_step, d_step = sw * 3, mw * 3
    Mc = (C.c_double * 6)(*M.ravel())
    bvc = (C.c_double * 4)(*bv)
    out_a = torch.zeros((n, mh, d_step), dtype=torch.uint8, device="cuda")
    out_b = torch.zeros((n, mh, d_step), dtype=torch.uint8, device="cuda")
    mid = torch.zeros((n, mh, d_step), dtype=torch.uint8, device="cuda")      # the existing entry needs this from the caller
    rc = lib.b200_resize_warpAffine_noalloc_dev(16, src.data_ptr(), sh * s_step, s_step, sw, sh, mw, mh,
                                                out_a.data_ptr(), mh * d_step, d_step, mw, mh, n, resize_interp,
                                                Mc, INTER_LINEAR, border, bvc, None)
    assert rc == 0
    torch.cuda.synchronize()
    rc = lib.b200_resize_warpAffine_batch_dev(16, src.data_ptr(), sh * s_step, s_step, sw, sh,
                                              mid.data_ptr(), mh * d_step, d_step, mw, mh,
                                              out_b.data_ptr(), mh * d_step, d_step, mw, mh, n, resize_interp,
                                              Mc, INTER_LINEAR, border, bvc, None)
    assert rc == 0
    torch.cuda.synchronize()
    return out_a.cpu().numpy(), out_b.cpu().numpy()


def test_noalloc_pipeline_matches_regular_entry_twice_in_a_row():
    assert torch.cuda.is_available()
    from opencv_b200 import capi
    lib = capi.lib
    # call 1: three 384x216 frames halved (2x2 box) then rotated
    n, sw, sh, mw, mh = 3, 384, 216, 192, 108
    src = torch.zeros((n, sh, sw * 3), dtype=torch.uint8, device="cuda")
    assert lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, 3, 10, n, 77, None) == 0
    M = rotation_inverse(mw, mh, 15.0, 0.9, cx=mw / 2, cy=mh / 2)
    a, b = _run_both(lib, src, sw, sh, mw, mh, n, M, INTER_LINEAR, BORDER_CONSTANT, (0, 0, 0, 0))
    assert a.any(), "output is all zero: nothing ran"
    assert np.array_equal(a, b), f"first call: {(a != b).sum()} bytes differ from the regular entry point"
    # call 2, straight after: other frames, a LARGER intermediate frame (fractional INTER_AREA), another border,
    # so stale contents of the shared array would show
    n, sw, sh, mw, mh = 2, 500, 300, 333, 177
    src = torch.zeros((n, sh, sw * 3), dtype=torch.uint8, device="cuda")
    assert lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, 3, 500, n, 78, None) == 0
    M = rotation_inverse(mw, mh, -40.0, 1.2)
    a, b = _run_both(lib, src, sw, sh, mw, mh, n, M, INTER_AREA, BORDER_REFLECT_101, (5, 6, 7, 0))
    assert np.array_equal(a, b), f"second call: {(a != b).sum()} bytes differ from the regular entry point"
    # ... and both are what the reference computes frame by frame (the regular entry alone proves nothing about results:
    # for so few frames it runs the same two kernels)
    O = oracle()
    src_h = src.cpu().numpy()
    for f in range(n):
        m, o = np.zeros((mh, mw, 3), np.uint8), np.zeros((mh, mw, 3), np.uint8)
        assert O.resize(src_h[f].reshape(sh, sw, 3), m, INTER_AREA) == 0
        assert O.warpAffine(m, o, M, INTER_LINEAR, BORDER_REFLECT_101, (5, 6, 7, 0)) == 0
        assert_same(a[f].reshape(o.shape), o, f"no-alloc entry vs oracle, frame {f}")


def test_noalloc_pipeline_allocates_nothing():
    """what the name promises: once the per-device set-up has run (first call), a call changes neither the device's free
    memory (no cudaMalloc / cudaFree underneath) nor, as far as torch can tell, torch's own pool"""
    from opencv_b200 import capi
    lib = capi.lib
    n, sw, sh, mw, mh = 2, 384, 216, 192, 108
    src = torch.zeros((n, sh, sw * 3), dtype=torch.uint8, device="cuda")
    dst = torch.zeros((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    assert lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, 3, 3, n, 5, None) == 0
    Mc = (C.c_double * 6)(*rotation_inverse(mw, mh, 15.0, 0.9).ravel())
    bv = (C.c_double * 4)(0, 0, 0, 0)

    def call():
        assert lib.b200_resize_warpAffine_noalloc_dev(16, src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, mw, mh, dst.data_ptr(), mh * mw * 3, mw * 3,
                                                      mw, mh, n, INTER_LINEAR, Mc, INTER_LINEAR, BORDER_CONSTANT, bv, None) == 0
        torch.cuda.synchronize()
    call()                                                   # set-up call (symbol address, kernel images, tensor maps)
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for _ in range(3):
        call()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, f"device free memory moved by {free0 - free1} bytes across no-alloc calls"
    assert torch.cuda.memory_reserved() == reserved0


def test_noalloc_pipeline_rejects_oversized_intermediate():
    from opencv_b200 import capi
    lib = capi.lib
    src = torch.zeros((1, 8, 8 * 3), dtype=torch.uint8, device="cuda")
    dst = torch.zeros((1, 8, 8 * 3), dtype=torch.uint8, device="cuda")
    Mc = (C.c_double * 6)(1, 0, 0, 0, 1, 0)
    bv = (C.c_double * 4)(0, 0, 0, 0)
    # 4096x2304x3 bytes is ~28 MB of intermediate frame: over the 8 MiB the library carries -> refused, nothing launched
    before = lib.b200_launch_count()
    rc = lib.b200_resize_warpAffine_noalloc_dev(16, src.data_ptr(), 8 * 24, 24, 8, 8, 4096, 2304,
                                                dst.data_ptr(), 8 * 24, 24, 8, 8, 1, INTER_LINEAR, Mc, INTER_LINEAR,
                                                BORDER_CONSTANT, bv, None)
    assert rc == -2
    assert lib.b200_launch_count() == before
