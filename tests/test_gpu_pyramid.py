"""cv::pyrDown / cv::pyrUp on the GPU (b200_hal_pyrdown / b200_hal_pyrup = the cv_hal_pyrdown / cv_hal_pyrup replacements, and the
device and batch entries) against the
oracle's restatement of imgproc/src/pyramids.cpp (pinned to the reference in tests/test_oracle_pinned.py): bit-exact for every depth,
CV_32F included (the kernel reproduces the reference's vector-body and scalar-tail evaluation orders)."""
import ctypes as C
import os

import numpy as np
import pytest

from common import *  # noqa: F403

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def test_pyrdown_hook_matches_oracle_on_every_cell():
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(0)
    rng = np.random.default_rng(3)
    for (dt, cn, (w, h), border, (dw, dh)) in pyrdown_cases():
        src = pyr_image(rng, dt, h, w, cn)
        got = np.zeros((dh, dw, cn), dt)
        rc = capi.lib.b200_hal_pyrdown(src.ctypes.data, src.strides[0], w, h, got.ctypes.data, got.strides[0], dw, dh, cv_type(src) & 7, cn, border)
        assert rc == 0, capi.last_error()
        want = oracle_pyrDown(src, dw, dh, border)
        assert (got.view(np.uint8) == want.view(np.uint8)).all(), f"{dt.__name__}C{cn} {w}x{h} -> {dw}x{dh} border {border}"


@pytest.mark.parametrize("dt,cn", [(np.uint8, 3), (np.uint8, 1), (np.uint16, 4), (np.float32, 3), (np.float32, 1), (np.int16, 2)])
def test_pyrdown_large_images_device_batch_and_pyramid(dt, cn):
    import opencv_b200 as cvb
    rng = np.random.default_rng(8)
    w, h = 1921, 1079
    src = pyr_image(rng, dt, h, w, cn)
    want = oracle_pyrDown(src, (w + 1) // 2, (h + 1) // 2, BORDER_REFLECT_101)
    tdt = {np.uint8: torch.uint8, np.uint16: torch.uint16, np.int16: torch.int16, np.float32: torch.float32}[dt]
    t = torch.from_numpy(src).cuda()
    got = cvb.pyrDown(t)
    assert (got.cpu().numpy().view(np.uint8) == want.view(np.uint8)).all()
    assert (cvb.pyrDown(src).view(np.uint8) == want.view(np.uint8)).all()                       # numpy -> the hook
    pyr = cvb.buildPyramid(t, 3)                                                                 # three levels, each from the one above
    lvl = src
    for k in range(1, 4):
        lvl = oracle_pyrDown(lvl, (lvl.shape[1] + 1) // 2, (lvl.shape[0] + 1) // 2, BORDER_REFLECT_101)
        assert (pyr[k].cpu().numpy().view(np.uint8) == lvl.view(np.uint8)).all(), f"level {k}"
    # frames of a batch in one launch
    from opencv_b200 import capi
    n = 3
    frames = np.stack([pyr_image(rng, dt, 240, 322, cn) for _ in range(n)])
    tf = torch.from_numpy(frames).cuda()
    out = torch.zeros((n, 120, 161, cn), dtype=tdt, device="cuda")
    esz = frames.itemsize * cn
    rc = capi.lib.b200_pyrDown_batch_dev(cv_type(frames[0]), tf.data_ptr(), 240 * 322 * esz, 322 * esz, 322, 240, out.data_ptr(), 120 * 161 * esz, 161 * esz,
                                         161, 120, n, BORDER_REPLICATE, None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    for f in range(n):
        assert (out[f].cpu().numpy().view(np.uint8) == oracle_pyrDown(frames[f], 161, 120, BORDER_REPLICATE).view(np.uint8)).all(), f"frame {f}"


def test_pyrdown_declines_what_the_reference_handles_elsewhere():
    from opencv_b200 import capi
    src = np.zeros((64, 64, 1), np.float64)
    dst = np.zeros((32, 32, 1), np.float64)
    assert capi.lib.b200_hal_pyrdown(src.ctypes.data, src.strides[0], 64, 64, dst.ctypes.data, dst.strides[0], 32, 32, 6, 1, BORDER_REFLECT_101) == 1   # CV_64F
    s8 = np.zeros((64, 64, 1), np.uint8)
    d8 = np.zeros((32, 32, 1), np.uint8)
    assert capi.lib.b200_hal_pyrdown(s8.ctypes.data, 64, 64, 64, d8.ctypes.data, 32, 32, 32, 0, 1, BORDER_CONSTANT) == 1      # rejected by cv::pyrDown itself
    d9 = np.zeros((40, 40, 1), np.uint8)
    assert capi.lib.b200_hal_pyrdown(s8.ctypes.data, 64, 64, 64, d9.ctypes.data, 40, 40, 40, 0, 1, BORDER_REFLECT_101) == 1   # size outside +-2


HAL_LIB = os.path.join(ROOT, "oracle", "_ref", "libcvref_b200hal.so")


@pytest.mark.skipif(not (os.path.exists(HAL_LIB) and os.path.exists(cvref_path())), reason="oracle/_ref flavours not built")
def test_reference_cv_pyrdown_runs_on_the_gpu_through_the_hal():
    """the reference's own cv::pyrDown, built with hal/cuda_b200 registered, reaches b200_hal_pyrdown at pyramids.cpp:1292"""
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(1 << 16)
    stock, hal = C.CDLL(cvref_path()), C.CDLL(HAL_LIB)
    rng = np.random.default_rng(21)
    l0 = capi.lib.b200_launch_count()
    n = 0
    for dt, cn in ((np.uint8, 3), (np.uint16, 1), (np.float32, 4), (np.int16, 2)):
        src = pyr_image(rng, dt, 721, 1283, cn)
        for border in (BORDER_REFLECT_101, BORDER_REPLICATE):
            a, b = np.zeros((361, 642, cn), dt), np.zeros((361, 642, cn), dt)
            for L, d in ((hal, a), (stock, b)):
                assert L.cvref_pyrDown(cv_type(src), src.ctypes.data_as(C.c_void_p), C.c_size_t(src.strides[0]), 1283, 721, d.ctypes.data_as(C.c_void_p),
                                       C.c_size_t(d.strides[0]), 0, 0, border, None) == 0
            assert (a.view(np.uint8) == b.view(np.uint8)).all(), f"cv::pyrDown {dt.__name__}C{cn} border {border}"
            n += 1
    assert capi.lib.b200_launch_count() - l0 == n
    capi.lib.b200_hal_set_min_pixels(0)


# ---------------------------------------------------------------- cv::pyrUp ----
def test_pyrup_hook_matches_oracle_on_every_cell():
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(0)
    rng = np.random.default_rng(5)
    n = 0
    for (dt, cn, (w, h), (dw, dh)) in pyrup_cases():
        src = pyrup_image(rng, dt, h, w, cn)
        got = np.zeros((dh, dw, cn), dt)
        rc = capi.lib.b200_hal_pyrup(src.ctypes.data, src.strides[0], w, h, got.ctypes.data, got.strides[0], dw, dh, cv_type(src) & 7, cn, BORDER_REFLECT_101)
        assert rc == 0, capi.last_error()
        want = oracle_pyrUp(src, dw, dh)
        assert (got.view(np.uint8) == want.view(np.uint8)).all(), f"{dt.__name__}C{cn} {w}x{h} -> {dw}x{dh}"
        n += 1
    assert n > 400


@pytest.mark.parametrize("dt,cn", [(np.uint8, 3), (np.uint8, 1), (np.uint16, 4), (np.float32, 3), (np.float32, 1), (np.int16, 2), (np.float64, 1)])
def test_pyrup_large_images_device_and_batch(dt, cn):
    import opencv_b200 as cvb
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(0)
    rng = np.random.default_rng(13)
    w, h = 963, 541                                        # odd sizes: unaligned row starts, partial tiles on both axes
    src = pyrup_image(rng, dt, h, w, cn)
    want = oracle_pyrUp(src, 2 * w, 2 * h)
    got = cvb.pyrUp(torch.from_numpy(src).cuda())
    assert (got.cpu().numpy().view(np.uint8) == want.view(np.uint8)).all()
    assert (cvb.pyrUp(src).view(np.uint8) == want.view(np.uint8)).all()                          # numpy -> the hook
    want_odd = oracle_pyrUp(src, 2 * w + 1, 2 * h - 1)
    assert (cvb.pyrUp(src, (2 * w + 1, 2 * h - 1)).view(np.uint8) == want_odd.view(np.uint8)).all()
    # a batch of padded frames in one launch
    frames = [pyrup_image(rng, dt, 120, 161, cn) for _ in range(3)]
    esz = src.dtype.itemsize * cn
    tf = torch.from_numpy(np.stack(frames).view(np.uint8)).cuda()
    out = torch.zeros((3, 240, 322 * esz), dtype=torch.uint8, device="cuda")
    rc = capi.lib.b200_pyrUp_batch_dev(cv_type(frames[0]), tf.data_ptr(), 120 * 161 * esz, 161 * esz, 161, 120, out.data_ptr(), 240 * 322 * esz, 322 * esz,
                                       322, 240, 3, BORDER_REFLECT_101, None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    for f in range(3):
        assert (out[f].cpu().numpy().reshape(-1) == oracle_pyrUp(frames[f], 322, 240).view(np.uint8).reshape(-1)).all(), f"frame {f}"


def test_pyrup_declines_what_the_reference_handles_elsewhere():
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(0)
    s8, d8 = np.zeros((32, 32, 1), np.uint8), np.zeros((64, 64, 1), np.uint8)
    assert capi.lib.b200_hal_pyrup(s8.ctypes.data, 32, 32, 32, d8.ctypes.data, 64, 64, 64, 0, 1, BORDER_REPLICATE) == 1       # cv::pyrUp asserts BORDER_DEFAULT
    assert capi.lib.b200_hal_pyrup(s8.ctypes.data, 32, 32, 32, d8.ctypes.data, 64, 64, 64, 1, 1, BORDER_REFLECT_101) == 1     # CV_8S: cv::pyrUp raises
    d9 = np.zeros((70, 70, 1), np.uint8)
    assert capi.lib.b200_hal_pyrup(s8.ctypes.data, 32, 32, 32, d9.ctypes.data, 70, 70, 70, 0, 1, BORDER_REFLECT_101) == 1     # size outside 2n +- 1
    d3 = np.zeros((2, 3, 1), np.uint8)
    assert capi.lib.b200_hal_pyrup(s8.ctypes.data, 32, 1, 1, d3.ctypes.data, 3, 3, 2, 0, 1, BORDER_REFLECT_101) == 1          # 1 -> 3 columns


@pytest.mark.skipif(not (os.path.exists(HAL_LIB) and os.path.exists(cvref_path())), reason="oracle/_ref flavours not built")
def test_reference_cv_pyrup_runs_on_the_gpu_through_the_hal():
    """the reference's own cv::pyrUp, built with hal/cuda_b200 registered, reaches b200_hal_pyrup at pyramids.cpp:1391"""
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(1 << 16)
    stock, hal = C.CDLL(cvref_path()), C.CDLL(HAL_LIB)
    rng = np.random.default_rng(29)
    l0 = capi.lib.b200_launch_count()
    n = 0
    for dt, cn in ((np.uint8, 3), (np.uint16, 1), (np.float32, 4), (np.int16, 2), (np.float64, 1)):
        src = pyrup_image(rng, dt, 361, 643, cn)
        a, b = np.zeros((722, 1286, cn), dt), np.zeros((722, 1286, cn), dt)
        for L, d in ((hal, a), (stock, b)):
            assert L.cvref_pyrUp(cv_type(src), src.ctypes.data_as(C.c_void_p), C.c_size_t(src.strides[0]), 643, 361, d.ctypes.data_as(C.c_void_p),
                                 C.c_size_t(d.strides[0]), 0, 0, BORDER_REFLECT_101) == 0
        assert (a.view(np.uint8) == b.view(np.uint8)).all(), f"cv::pyrUp {dt.__name__}C{cn}"
        n += 1
    assert capi.lib.b200_launch_count() - l0 == n
    capi.lib.b200_hal_set_min_pixels(0)
