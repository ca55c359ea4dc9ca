"""Drop-in proof at the real boundary: the REFERENCE's own cv::warpAffine / cv::warpPerspective / cv::remap /
cv::resize (oracle/_ref/libcvref_b200hal.so = the reference sources compiled with hal/cuda_b200 registered through
custom_hal.hpp, see oracle/build_ref.sh) call into libb200geom.so at their CALL_HAL sites
(imgwarp.cpp:1608-1622, :2255, :2749; resize.cpp:3648); results must equal the stock reference build
(libcvref.so) bit for bit, and the library's launch counter must show that the GPU did the work."""
import ctypes as C
import os

import numpy as np
import pytest

from common import *  # noqa: F403

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

HAL_LIB = os.path.join(ROOT, "oracle", "_ref", "libcvref_b200hal.so")
needs = pytest.mark.skipif(not (os.path.exists(HAL_LIB) and os.path.exists(cvref_path())),
                           reason="oracle/_ref flavours not built (bash oracle/build_ref.sh in the build container)")

_vp, _sz, _i = C.c_void_p, C.c_size_t, C.c_int
_IMG = [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i]


class PublicApi:
    """cv::warpAffine / cv::warpPerspective / cv::resize / cv::remap of one reference flavour"""

    def __init__(self, path):
        self.lib = C.CDLL(path)
        self.lib.cvref_cv_warpAffine.argtypes = _IMG + [_vp, _i, _i, _vp]
        self.lib.cvref_cv_warpPerspective.argtypes = _IMG + [_vp, _i, _i, _vp]
        self.lib.cvref_cv_resize.argtypes = _IMG + [C.c_double, C.c_double, _i, _vp]
        self.hooks = HookLib(self.lib, {"warpAffine": "cvref_hal_warpAffine", "warpPerspective": "cvref_hal_warpPerspective",
                                        "resize": "cvref_hal_resize", "remap32f": "cvref_remap32f",
                                        "remap32fc2": "cvref_remap32fc2", "remap16s": "cvref_remap16s"})

    def warp(self, name, src, dst, M, flags, border, bv):
        Mc = (C.c_double * M.size)(*M.ravel())
        bvc = (C.c_double * 4)(*bv)
        h, w = src.shape[:2]
        dh, dw = dst.shape[:2]
        return getattr(self.lib, "cvref_cv_" + name)(cv_type(src), src.ctypes.data, src.strides[0], w, h,
                                                     dst.ctypes.data, dst.strides[0], dw, dh, Mc, flags, border, bvc)

    def resize(self, src, dst, interp):
        h, w = src.shape[:2]
        dh, dw = dst.shape[:2]
        return self.lib.cvref_cv_resize(cv_type(src), src.ctypes.data, src.strides[0], w, h, dst.ctypes.data,
                                        dst.strides[0], dw, dh, 0.0, 0.0, interp, None)


@needs
def test_reference_public_api_runs_on_the_gpu_through_the_hal_and_matches_stock():
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(1 << 16)            # the shipped default gate
    stock, hal = PublicApi(cvref_path()), PublicApi(HAL_LIB)
    rng = np.random.default_rng(99)
    launches0 = capi.lib.b200_launch_count()
    n_calls = 0
    for dtype, cn in ((np.uint8, 3), (np.uint8, 4), (np.uint16, 1), (np.float32, 3)):
        src = rand_image(rng, 480, 640, cn, dtype)
        shape = (500, 700) if cn == 1 else (500, 700, cn)
        Mf = np.array([[0.9, 0.2, 30.0], [-0.2, 0.95, 40.0]])              # forward matrix: cv::warpAffine inverts it
        Hf = np.array([[1.05, 0.02, 12.0], [0.03, 0.98, -7.0], [2e-5, -1e-5, 1.0]])
        for border in (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP):
            for flags in (INTER_NEAREST, INTER_LINEAR):
                for name, M in (("warpAffine", Mf), ("warpPerspective", Hf)):
                    a, b = np.zeros(shape, dtype), np.zeros(shape, dtype)
                    assert hal.warp(name, src, a, M, flags, border, (3, 4, 5, 6)) == 0
                    assert stock.warp(name, src, b, M, flags, border, (3, 4, 5, 6)) == 0
                    assert_same(a, b, f"cv::{name} {dtype.__name__}C{cn} flags{flags} border{border}")
                    n_calls += 1
        mx = rng.uniform(-20, 660, (500, 700)).astype(np.float32)
        my = rng.uniform(-20, 500, (500, 700)).astype(np.float32)
        m1, m2 = convert_maps_fixed(mx, my)
        for interp in (INTER_NEAREST, INTER_LINEAR):
            for f in (lambda L, d: L.hooks.remap32f(src, d, mx, my, interp, BORDER_REFLECT_101),
                      lambda L, d: L.hooks.remap16s(src, d, m1, m2, interp, BORDER_REFLECT_101)):
                a, b = np.zeros(shape, dtype), np.zeros(shape, dtype)
                assert f(hal, a) == 0 and f(stock, b) == 0
                assert_same(a, b, f"cv::remap {dtype.__name__}C{cn} interp{interp}")
                n_calls += 1
        for (dh, dw) in ((240, 320), (300, 500), (960, 1280), (160, 640)):
            exact = (INTER_NEAREST_EXACT,) + ((INTER_LINEAR_EXACT,) if dtype in (np.uint8, np.uint16, np.int16) else ())
            for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA, INTER_CUBIC, INTER_LANCZOS4) + exact:
                a = np.zeros((dh, dw) if cn == 1 else (dh, dw, cn), dtype)
                b = np.zeros_like(a)
                assert hal.resize(src, a, interp) == 0 and stock.resize(src, b, interp) == 0
                assert_same(a, b, f"cv::resize {dtype.__name__}C{cn} ->{dw}x{dh} interp{interp}")
                n_calls += 1
    assert capi.lib.b200_launch_count() - launches0 == n_calls, "every call must have been served by a b200 kernel"
    # below the pixel gate (the re-entrancy guard for the CPU fallback's per-tile remap, imgwarp.cpp:2040-2044) and for an
    # unsupported interpolation the hooks decline and the reference's CPU path answers: still identical, no launch
    small = rand_image(rng, 60, 80, 3, np.uint8)
    l1 = capi.lib.b200_launch_count()
    a, b = np.zeros((64, 90, 3), np.uint8), np.zeros((64, 90, 3), np.uint8)
    assert hal.warp("warpAffine", small, a, Mf, INTER_LINEAR, BORDER_CONSTANT, (0,) * 4) == 0
    assert stock.warp("warpAffine", small, b, Mf, INTER_LINEAR, BORDER_CONSTANT, (0,) * 4) == 0
    assert_same(a, b, "gated small image")
    big = rand_image(rng, 480, 640, 3, np.uint8)
    a, b = np.zeros((500, 700, 3), np.uint8), np.zeros((500, 700, 3), np.uint8)
    src64 = rng.random((480, 640)).astype(np.float64)                                    # 64F: served except with BORDER_TRANSPARENT
    a64, b64 = np.zeros((500, 700), np.float64), np.zeros((500, 700), np.float64)
    assert hal.warp("warpAffine", src64, a64, Mf, INTER_CUBIC, 5, (0,) * 4) == 0
    assert stock.warp("warpAffine", src64, b64, Mf, INTER_CUBIC, 5, (0,) * 4) == 0
    assert_same(a64, b64, "64F with BORDER_TRANSPARENT falls back")
    assert capi.lib.b200_launch_count() == l1
    assert hal.warp("warpAffine", src64, a64, Mf, INTER_CUBIC, BORDER_CONSTANT, (0,) * 4) == 0
    assert stock.warp("warpAffine", src64, b64, Mf, INTER_CUBIC, BORDER_CONSTANT, (0,) * 4) == 0
    assert_same(a64, b64, "64F cubic on the GPU")
    assert capi.lib.b200_launch_count() == l1 + 1
    l1 += 1
    assert hal.warp("warpAffine", src64, a64, Mf, INTER_LINEAR, BORDER_REFLECT_101, (0,) * 4) == 0
    assert stock.warp("warpAffine", src64, b64, Mf, INTER_LINEAR, BORDER_REFLECT_101, (0,) * 4) == 0
    assert_same(a64, b64, "64F bilinear on the GPU")
    assert capi.lib.b200_launch_count() == l1 + 1
    l1 += 1
    init = rand_image(rng, 500, 700, 3, np.uint8)                                         # BORDER_TRANSPARENT: dst contents survive
    a, b = init.copy(), init.copy()
    assert hal.warp("warpAffine", big, a, Mf, INTER_LINEAR, 5, (0,) * 4) == 0
    assert stock.warp("warpAffine", big, b, Mf, INTER_LINEAR, 5, (0,) * 4) == 0
    assert_same(a, b, "cv::warpAffine BORDER_TRANSPARENT")
    assert capi.lib.b200_launch_count() == l1 + 1
    l1 += 1
    for interp in (INTER_CUBIC, INTER_LANCZOS4):                                          # served by the GPU
        a[:] = 0; b[:] = 0
        assert hal.warp("warpAffine", big, a, Mf, interp, BORDER_REFLECT_101, (0,) * 4) == 0
        assert stock.warp("warpAffine", big, b, Mf, interp, BORDER_REFLECT_101, (0,) * 4) == 0
        assert_same(a, b, f"cv::warpAffine interp {interp}")
    assert capi.lib.b200_launch_count() == l1 + 2
    capi.lib.b200_hal_set_min_pixels(0)


@needs
def test_default_policy_leaves_the_cells_the_cpu_wins_to_the_cpu():
    """b200_hal_set_policy(0), the C default: INTER_NEAREST cells, light resizes and large INTER_LINEAR remaps return NOT_IMPLEMENTED
    (the reference's CPU path answers, results identical, no launch); the cells the GPU wins are still served"""
    from opencv_b200 import capi
    stock, hal = PublicApi(cvref_path()), PublicApi(HAL_LIB)
    rng = np.random.default_rng(5)
    old = capi.lib.b200_hal_set_policy(0)
    try:
        capi.lib.b200_hal_set_min_pixels(1 << 16)
        src = rand_image(rng, 720, 1280, 3, np.uint8)
        Mf = np.array([[0.9, 0.2, 30.0], [-0.2, 0.95, 40.0]])
        served, declined = 0, 0
        for name, call, gpu in (
                ("warpAffine NEAREST", lambda L, d: L.warp("warpAffine", src, d, Mf, INTER_NEAREST, BORDER_CONSTANT, (0,) * 4), False),
                ("warpAffine LINEAR", lambda L, d: L.warp("warpAffine", src, d, Mf, INTER_LINEAR, BORDER_CONSTANT, (0,) * 4), True),
                ("warpAffine CUBIC", lambda L, d: L.warp("warpAffine", src, d, Mf, INTER_CUBIC, BORDER_REPLICATE, (0,) * 4), True)):
            a, b = np.zeros((720, 1280, 3), np.uint8), np.zeros((720, 1280, 3), np.uint8)
            l0 = capi.lib.b200_launch_count()
            assert call(hal, a) == 0 and call(stock, b) == 0
            assert_same(a, b, name)
            assert (capi.lib.b200_launch_count() - l0 == 1) == gpu, name
        for (dw, dh, interp, gpu) in ((640, 360, INTER_LINEAR, False), (640, 360, INTER_AREA, False), (1000, 500, INTER_LINEAR, False), (1000, 500, INTER_NEAREST, False),
                                      (1000, 500, INTER_AREA, True), (1000, 500, INTER_CUBIC, True), (1000, 500, INTER_LANCZOS4, True)):
            a, b = np.zeros((dh, dw, 3), np.uint8), np.zeros((dh, dw, 3), np.uint8)
            l0 = capi.lib.b200_launch_count()
            assert hal.resize(src, a, interp) == 0 and stock.resize(src, b, interp) == 0
            assert_same(a, b, f"resize {dw}x{dh} interp {interp}")
            assert (capi.lib.b200_launch_count() - l0 == 1) == gpu, f"resize {dw}x{dh} interp {interp}"
        big = rand_image(rng, 1080, 1920, 1, np.uint8)
        for (w, h, gpu) in ((640, 480, True), (1920, 1080, False)):                 # INTER_LINEAR remap: small served, > 2^20 pixels left to the CPU
            mx = rng.uniform(0, 1900, (h, w)).astype(np.float32); my = rng.uniform(0, 1000, (h, w)).astype(np.float32)
            a, b = np.zeros((h, w), np.uint8), np.zeros((h, w), np.uint8)
            l0 = capi.lib.b200_launch_count()
            assert hal.hooks.remap32f(big, a, mx, my, INTER_LINEAR, BORDER_REFLECT_101) == 0 and stock.hooks.remap32f(big, b, mx, my, INTER_LINEAR, BORDER_REFLECT_101) == 0
            assert_same(a, b, f"remap {w}x{h}")
            assert (capi.lib.b200_launch_count() - l0 == 1) == gpu, f"remap {w}x{h}"
        # pyramids: pyrUp always, pyrDown only where the host loses (8UC3)
        L_s, L_h = C.CDLL(cvref_path()), C.CDLL(HAL_LIB)
        for (fn, cn, dsz, gpu) in (("cvref_pyrDown", 3, (640, 360), True), ("cvref_pyrDown", 1, (640, 360), False), ("cvref_pyrDown", 4, (640, 360), False),
                                   ("cvref_pyrUp", 1, (2560, 1440), True), ("cvref_pyrUp", 3, (2560, 1440), True)):
            img = rand_image(rng, 720, 1280, cn, np.uint8).reshape(720, 1280, cn)
            a, b = np.zeros((dsz[1], dsz[0], cn), np.uint8), np.zeros((dsz[1], dsz[0], cn), np.uint8)
            l0 = capi.lib.b200_launch_count()
            for L, d in ((L_h, a), (L_s, b)):
                args = [cv_type(img), img.ctypes.data_as(C.c_void_p), C.c_size_t(img.strides[0]), 1280, 720, d.ctypes.data_as(C.c_void_p), C.c_size_t(d.strides[0]), 0, 0, BORDER_REFLECT_101]
                assert getattr(L, fn)(*(args + ([None] if fn == "cvref_pyrDown" else []))) == 0
            assert_same(a, b, f"{fn} cn{cn}")
            assert (capi.lib.b200_launch_count() - l0 == 1) == gpu, f"{fn} cn{cn}"
    finally:
        capi.lib.b200_hal_set_policy(old)
        capi.lib.b200_hal_set_min_pixels(0)


@needs
def test_reference_cv_warppolar_reaches_the_remap_hook():
    """cv::warpPolar (imgwarp.cpp:3163-3281) builds its maps on the host and calls cv::remap: with hal/cuda_b200 linked the image comes
    from b200_hal_remap32f, outliers filled or kept"""
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(1 << 16)
    stock, hal = C.CDLL(cvref_path()), C.CDLL(HAL_LIB)
    rng = np.random.default_rng(43)
    src = rand_image(rng, 480, 640, 3, np.uint8)
    try:
        for flags in (INTER_LINEAR | WARP_FILL_OUTLIERS, INTER_LINEAR, INTER_LINEAR | WARP_FILL_OUTLIERS | WARP_POLAR_LOG, INTER_CUBIC | WARP_FILL_OUTLIERS):
            a, b = np.full((600, 400, 3), 9, np.uint8), np.full((600, 400, 3), 9, np.uint8)
            l0 = capi.lib.b200_launch_count()
            for L, d in ((hal, a), (stock, b)):
                L.cvref_warpPolar.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_float, C.c_double, C.c_int]
                assert L.cvref_warpPolar(cv_type(src), src.ctypes.data, src.strides[0], 640, 480, d.ctypes.data, d.strides[0], 400, 600, 320.0, 240.0, 230.0, flags) == 0
            assert_same(a, b, f"cv::warpPolar flags {flags}")
            assert capi.lib.b200_launch_count() - l0 == 1, f"flags {flags}"
    finally:
        capi.lib.b200_hal_set_min_pixels(0)
