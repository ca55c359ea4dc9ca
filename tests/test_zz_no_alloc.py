"""b200_resize_warpAffine_noalloc_dev -- the pipeline entry for callers that may not allocate GPU memory
(a running capture loop, a plugin without a memory budget, code where cudaMalloc would stall the device).
Its intermediate frame lives in a fixed array inside the library, so the only thing to prove is that it
returns exactly what the ordinary entry point (caller-provided intermediate buffer) returns -- twice in a
row with different frames, because the array is shared and keeps the previous call's contents.
This file sorts last on purpose: a `-x` run has already executed every other test when it gets here.
It also passes alone in a fresh process (it then performs the per-device set-up itself)."""
import ctypes as C

import numpy as np
import pytest

from common import rotation_inverse, INTER_LINEAR, INTER_AREA, BORDER_CONSTANT, BORDER_REFLECT_101

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
