"""The HAL error contract on a host WITHOUT a usable GPU (SURVEY 8(b) "Errors"; imgproc/src/hal_replacement.hpp:1608-1614:
CALL_HAL throws on every return code other than 0 / 1).  Run in a child process with CUDA_VISIBLE_DEVICES empty so that it
means the same thing in the build container and on a GPU box: every b200_hal_* hook must DECLINE (1), never fail (-1),
and the reference's own public cv::resize / cv::warpAffine / cv::warpPerspective compiled with hal/cuda_b200 registered
(oracle/_ref/libcvref_b200hal.so) must then answer from their CPU path with stock output."""
import os
import subprocess
import sys
import textwrap

import pytest

from common import ROOT, cvref_path

HAL_LIB = os.path.join(ROOT, "oracle", "_ref", "libcvref_b200hal.so")

CHILD = textwrap.dedent("""
    import ctypes as C, os, sys
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, ROOT)
    from common import *
    from opencv_b200 import capi
    assert capi.lib.b200_device_count() == 0, "child must not see a GPU"
    capi.lib.b200_hal_set_min_pixels(0)
    hooks = HookLib(capi.lib, {"warpAffine": "b200_hal_warpAffine", "warpPerspective": "b200_hal_warpPerspective",
                               "resize": "b200_hal_resize", "remap32f": "b200_hal_remap32f",
                               "remap32fc2": "b200_hal_remap32fc2", "remap16s": "b200_hal_remap16s"})
    rng = np.random.default_rng(5)
    src = rand_image(rng, 300, 400, 3, np.uint8)
    dst = np.full((320, 420, 3), 77, np.uint8)
    M = np.array([[0.9, 0.2, 30.0], [-0.2, 0.95, 40.0]])
    H = np.array([[1.05, 0.02, 12.0], [0.03, 0.98, -7.0], [2e-5, -1e-5, 1.0]])
    mx = rng.uniform(-20, 420, (320, 420)).astype(np.float32)
    my = rng.uniform(-20, 320, (320, 420)).astype(np.float32)
    m1, m2 = convert_maps_fixed(mx, my)
    mxy = np.ascontiguousarray(np.stack([mx, my], axis=-1))
    for _ in range(2):                                   # the second round takes the remembered "no device" path
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4):
            assert hooks.warpAffine(src, dst, M, interp, BORDER_CONSTANT) == 1, capi.last_error()
            assert hooks.warpPerspective(src, dst, H, interp, BORDER_REPLICATE) == 1, capi.last_error()
            assert hooks.remap32f(src, dst, mx, my, interp, BORDER_REFLECT_101) == 1, capi.last_error()
            assert hooks.remap32fc2(src, dst, mxy, interp, BORDER_REFLECT_101) == 1, capi.last_error()
            assert hooks.remap16s(src, dst, m1, m2, interp, BORDER_WRAP) == 1, capi.last_error()
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA, INTER_CUBIC, INTER_LANCZOS4, INTER_LINEAR_EXACT):
            out = np.full((150, 200, 3), 77, np.uint8)
            assert hooks.resize(src, out, interp) == 1, capi.last_error()
            assert (out == 77).all()
    assert (dst == 77).all(), "a declining hook must not touch dst"
    assert capi.lib.b200_launch_count() == 0
    if HAVE_REF:
        _vp, _sz, _i = C.c_void_p, C.c_size_t, C.c_int
        _IMG = [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i]
        def api(path):
            L = C.CDLL(path)
            L.cvref_cv_warpAffine.argtypes = _IMG + [_vp, _i, _i, _vp]
            L.cvref_cv_warpPerspective.argtypes = _IMG + [_vp, _i, _i, _vp]
            L.cvref_cv_resize.argtypes = _IMG + [C.c_double, C.c_double, _i, _vp]
            return L
        stock, hal = api(STOCK), api(HAL)
        def warp(L, name, Mx, d):
            Mc = (C.c_double * Mx.size)(*Mx.ravel()); bv = (C.c_double * 4)(1, 2, 3, 4)
            return getattr(L, "cvref_cv_" + name)(cv_type(src), src.ctypes.data, src.strides[0], 400, 300,
                                                  d.ctypes.data, d.strides[0], 420, 320, Mc, INTER_LINEAR, BORDER_CONSTANT, bv)
        for name, Mx in (("warpAffine", M), ("warpPerspective", H)):
            a, b = np.zeros_like(dst), np.zeros_like(dst)
            assert warp(hal, name, Mx, a) == 0, "cv::%s threw through the HAL" % name
            assert warp(stock, name, Mx, b) == 0
            assert_same(a, b, "cv::%s without a GPU" % name)
        for interp in (INTER_LINEAR, INTER_AREA, INTER_CUBIC):
            a, b = np.zeros((150, 200, 3), np.uint8), np.zeros((150, 200, 3), np.uint8)
            for L, d in ((hal, a), (stock, b)):
                assert L.cvref_cv_resize(cv_type(src), src.ctypes.data, src.strides[0], 400, 300, d.ctypes.data,
                                         d.strides[0], 200, 150, 0.0, 0.0, interp, None) == 0
            assert_same(a, b, "cv::resize without a GPU")
    print("NO_DEVICE_OK")
""")


def test_hooks_decline_and_opencv_falls_back_without_a_gpu():
    have_ref = os.path.exists(HAL_LIB) and os.path.exists(cvref_path())
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    head = f"ROOT = {ROOT!r}\nHAVE_REF = {have_ref!r}\nSTOCK = {cvref_path()!r}\nHAL = {HAL_LIB!r}\n"
    r = subprocess.run([sys.executable, "-c", head + CHILD], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "NO_DEVICE_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
    # said once, not on every call
    assert r.stderr.count("b200 HAL:") == 1, r.stderr[-2000:]
