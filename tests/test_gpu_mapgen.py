"""b200_fisheye_initUndistortRectifyMap_dev against the oracle's restatement of cv::fisheye::initUndistortRectifyMap
(calib3d/src/fisheye.cpp:530-631; the restatement is pinned to the reference bit for bit in tests/test_oracle_pinned.py).
The kernel reproduces every fp64 step of the reference, including the repeated-addition row accumulators; the one thing
it cannot share with the host is atan (CUDA math library vs the host libm, both within an ulp of the true value but not of
each other).  Tolerance, written here: at most 1e-5 of the entries may differ, and then by one float ulp (CV_32FC1 maps)
or by one unit of 1/32 pixel (CV_16SC2 + CV_16UC1 maps); infinities and their fixed-point images must agree exactly."""
import ctypes as C

import numpy as np
import pytest

from common import *  # noqa: F403

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _gpu_map(case, m1type, iR):
    from opencv_b200 import capi
    _name, K, D, _R, _P, (w, h) = case
    if m1type == 11:
        m1 = torch.zeros((h, w, 2), dtype=torch.int16, device="cuda"); m2 = torch.zeros((h, w), dtype=torch.int16, device="cuda")
    else:
        m1 = torch.zeros((h, w), dtype=torch.float32, device="cuda"); m2 = torch.zeros((h, w), dtype=torch.float32, device="cuda")
    rc = capi.lib.b200_fisheye_initUndistortRectifyMap_dev((C.c_double * 9)(*np.asarray(K).ravel()), (C.c_double * 4)(*np.asarray(D).ravel()),
                                                           (C.c_double * 9)(*np.asarray(iR).ravel()), w, h, m1type, m1.data_ptr(),
                                                           m1.stride(0) * m1.element_size(), m2.data_ptr(), m2.stride(0) * m2.element_size(), None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    a1, a2 = m1.cpu().numpy(), m2.cpu().numpy()
    return a1, (a2.view(np.uint16) if m1type == 11 else a2)


@pytest.mark.parametrize("m1type", [11, 5])
def test_fisheye_map_on_device_equals_oracle(m1type):
    total = bad = 0
    for case in fisheye_cases():
        _name, K, D, R, P, (w, h) = case
        RR = np.eye(3) if R is None else R
        PP = np.eye(3) if P is None else P
        iR = np.linalg.inv(PP @ RR).ravel()
        o1, o2 = oracle_fisheye_map(case, m1type, iR)
        g1, g2 = _gpu_map(case, m1type, iR)
        if m1type == 5:
            for g, o in ((g1, o1), (g2, o2)):
                assert (np.isinf(g) == np.isinf(o)).all() and (g[np.isinf(o)] == o[np.isinf(o)]).all(), case[0]
                fin = np.isfinite(o)
                d = np.abs(g[fin].view(np.int32).astype(np.int64) - o[fin].view(np.int32).astype(np.int64))
                assert d.max(initial=0) <= 1, f"{case[0]}: {d.max()} float ulps apart"
                total += d.size; bad += int((d != 0).sum())
        else:
            # back to 1/32 pixel units: (xy << 5) + fraction
            gx = g1[..., 0].astype(np.int64) * 32 + (g2 & 31); gy = g1[..., 1].astype(np.int64) * 32 + (g2 >> 5)
            ox = o1[..., 0].astype(np.int64) * 32 + (o2 & 31); oy = o1[..., 1].astype(np.int64) * 32 + (o2 >> 5)
            for g, o in ((gx, ox), (gy, oy)):
                d = np.abs(g - o)
                d = np.minimum(d, 65536 * 32 - d)                    # the (short) cast wraps (fisheye.cpp:615-616)
                assert d.max() <= 1, f"{case[0]}: {d.max()} units of 1/32 px apart"
                total += d.size; bad += int((d != 0).sum())
    assert bad <= 1e-5 * total, f"{bad} of {total} map entries differ from the oracle"


def test_remap_through_a_device_built_fisheye_map():
    """BASELINE config 4 end to end on the device: build the map, remap a frame with it; against the oracle's map + remap"""
    import opencv_b200 as cvb
    case = fisheye_cases()[0]
    _name, K, D, _R, _P, (w, h) = case
    rng = np.random.default_rng(4)
    src = rand_image(rng, h, w, 3, np.uint8)
    o1, o2 = oracle_fisheye_map(case, 11, np.eye(3).ravel())
    want = np.zeros_like(src)
    assert oracle().remap16s(src, want, o1, o2, INTER_LINEAR, BORDER_REFLECT_101) == 0
    m1, m2 = cvb.fisheye_initUndistortRectifyMap(K, D, None, None, (w, h), 11)
    got = cvb.remap(torch.from_numpy(src).cuda(), m1, m2, interpolation=INTER_LINEAR, borderMode=BORDER_REFLECT_101).cpu().numpy()
    assert (got != want).sum() <= 3 * 16, "a handful of pixels at most may move with a last-bit map difference"


def test_fisheye_undistort_image_entry():
    """b200_fisheye_undistortImage_dev = map + remap on one stream; against the oracle's map + remap16s (BORDER_CONSTANT)"""
    from opencv_b200 import capi
    case = fisheye_cases()[1]
    _name, K, D, _R, P, (w, h) = case
    rng = np.random.default_rng(8)
    src = rand_image(rng, h, w, 4, np.uint8)
    iKnew = np.linalg.inv(P).ravel()
    o1, o2 = oracle_fisheye_map((_name, K, D, None, None, (w, h)), 11, iKnew)
    want = np.zeros_like(src)
    assert oracle().remap16s(src, want, o1, o2, INTER_LINEAR, BORDER_CONSTANT) == 0
    t = torch.from_numpy(src).cuda()
    out = torch.zeros_like(t)
    m1 = torch.zeros((h, w, 2), dtype=torch.int16, device="cuda"); m2 = torch.zeros((h, w), dtype=torch.int16, device="cuda")
    rc = capi.lib.b200_fisheye_undistortImage_dev(24, t.data_ptr(), w * 4, w, h, out.data_ptr(), w * 4, w, h, (C.c_double * 9)(*K.ravel()),
                                                  (C.c_double * 4)(*D.ravel()), (C.c_double * 9)(*iKnew), m1.data_ptr(), w * 4, m2.data_ptr(), w * 2, None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    assert (out.cpu().numpy() != want).sum() <= 4 * 16


@pytest.mark.parametrize("m1type", [11, 5, 13])
def test_undistort_map_on_device_is_bit_exact(m1type):
    """b200_initUndistortRectifyMap_dev: pure IEEE arithmetic (fused where the reference's AVX2 body fuses) -> bit-identical to the
    oracle's restatement of the AVX2 dispatch variant, vector chunks and scalar tail alike"""
    from opencv_b200 import capi
    for case in undistort_cases():
        _n, K, dist, R, P, (w, h) = case
        RR = np.eye(3) if R is None else R
        Ar = K if P is None else P
        iR = np.linalg.inv(Ar @ RR).ravel()
        tilt = tilt_matrix(dist[12], dist[13]).ravel() if len(dist) >= 14 else np.eye(3).ravel()
        o1, o2 = oracle_undistort_map(case, m1type, iR, tilt)
        if m1type == 11:
            m1 = torch.zeros((h, w, 2), dtype=torch.int16, device="cuda"); m2 = torch.zeros((h, w), dtype=torch.int16, device="cuda")
        elif m1type == 5:
            m1 = torch.zeros((h, w), dtype=torch.float32, device="cuda"); m2 = torch.zeros((h, w), dtype=torch.float32, device="cuda")
        else:
            m1 = torch.zeros((h, w, 2), dtype=torch.float32, device="cuda"); m2 = None
        rc = capi.lib.b200_initUndistortRectifyMap_dev((C.c_double * 9)(*K.ravel()), (C.c_double * 12)(*dist12_of(dist)), (C.c_double * 9)(*iR),
                                                       (C.c_double * 9)(*tilt), w, h, m1type, m1.data_ptr(), m1.stride(0) * m1.element_size(),
                                                       m2.data_ptr() if m2 is not None else None, m2.stride(0) * m2.element_size() if m2 is not None else 0, None)
        assert rc == 0, capi.last_error()
        torch.cuda.synchronize()
        assert (m1.cpu().numpy().view(np.uint8) == o1.view(np.uint8)).all(), f"{case[0]} map1 (m1type {m1type})"
        if m2 is not None:
            assert (m2.cpu().numpy().view(np.uint8) == o2.view(np.uint8)).all(), f"{case[0]} map2 (m1type {m1type})"


def test_undistort_entry_and_python_mirror():
    """cv::undistort on the device: stripe-wise maps (each stripe inverts its own camera matrix, as undistort.dispatch.cpp:311-325 does)
    + remap, against the oracle's restatement of cv::undistort -- image and both maps, through the C entry and the Python mirror"""
    import opencv_b200 as cvb
    from opencv_b200 import capi
    rng = np.random.default_rng(12)
    for case in undistort_image_cases():
        _n, K, dist, newK, (w, h) = case
        src = rand_image(rng, h, w, 3, np.uint8)
        want, o1, o2, tilt = oracle_undistort(case, src)
        t = torch.from_numpy(src).cuda()
        got, maps = cvb.undistort(t, K, dist, newK)
        assert_same(got.cpu().numpy(), want, f"cv::undistort (python mirror) {case[0]}")
        assert (maps[0].cpu().numpy() == o1).all() and (maps[1].cpu().numpy().view(np.uint16) == o2).all(), case[0] + " maps"
        out = torch.zeros_like(t)
        m1, m2 = torch.zeros_like(maps[0]), torch.zeros_like(maps[1])
        rc = capi.lib.b200_undistort_dev(16, t.data_ptr(), w * 3, w, h, out.data_ptr(), w * 3, (C.c_double * 9)(*K.ravel()),
                                         (C.c_double * 12)(*dist12_of(dist)), (C.c_double * 9)(*newK.ravel()) if newK is not None else None,
                                         (C.c_double * 9)(*tilt), m1.data_ptr(), w * 4, m2.data_ptr(), w * 2, None)
        assert rc == 0, capi.last_error()
        torch.cuda.synchronize()
        assert_same(out.cpu().numpy(), want, f"b200_undistort_dev {case[0]}")


def test_undistort_map_python_mirror_defaults():
    """opencv_b200.initUndistortRectifyMap: newCameraMatrix=None centres the principal point (undistort.dispatch.cpp:108), R and the
    tilt go through b200_undistort_params; against the oracle fed the oracle's own matrices"""
    import opencv_b200 as cvb
    for case in undistort_cases()[:3]:
        _n, K, dist, R, P, (w, h) = case
        newK = P
        if newK is None:
            newK = K.copy(); newK[0, 2] = (w - 1) * 0.5; newK[1, 2] = (h - 1) * 0.5
        tx, ty = (dist[12], dist[13]) if len(dist) >= 14 else (0.0, 0.0)
        iR, tilt = oracle_undistort_params(newK, R, tx, ty)
        o1, o2 = oracle_undistort_map(case, 11, iR, tilt)
        m1, m2 = cvb.initUndistortRectifyMap(K, dist, R, P, (w, h), 11)
        assert (m1.cpu().numpy() == o1).all() and (m2.cpu().numpy().view(np.uint16) == o2).all(), case[0]


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_stitching_warper_maps_and_warp(kind):
    """b200_stitch_buildMaps_dev against the oracle's restatement of the warpers' map loops: the plane projector bit for bit; the
    spherical / cylindrical ones within 4e-6 relative (+1e-4 px absolute) -- sinf / cosf come from the CUDA math library, the
    reference's from the host libm.  b200_stitch_warp_dev (maps + remap): bit-identical to the oracle's remap through the SAME maps."""
    from opencv_b200 import capi
    rng = np.random.default_rng(31)
    for case in stitch_cases():
        K, R, T, scale, (w, h) = case
        Kd, Rd = K.astype(np.float64), R.astype(np.float64)
        kr = np.ascontiguousarray((K @ R.T).astype(np.float32)).ravel()          # setCameraParams: K * R^t in float (warpers.cpp:152)
        tt = T if kind == 0 else np.zeros(3, np.float32)
        # a ROI like the warper would detect: forward-map the corners on the host (the test needs only SOME window of the plane)
        tl, (mw, mh) = {0: ((-300, -250), (600, 480)), 1: ((-380, 900), (610, 470)), 2: ((-380, -250), (610, 470))}[kind]
        ox, oy = oracle_stitch_maps(kind, kr, tt, scale, tl, mw, mh)
        xm = torch.zeros((mh, mw), dtype=torch.float32, device="cuda"); ym = torch.zeros_like(xm)
        rc = capi.lib.b200_stitch_buildMaps_dev(kind, kr.ctypes.data, tt.ctypes.data, scale, tl[0], tl[1], mw, mh, xm.data_ptr(), mw * 4, ym.data_ptr(), mw * 4, None)
        assert rc == 0, capi.last_error()
        torch.cuda.synchronize()
        gx, gy = xm.cpu().numpy(), ym.cpu().numpy()
        if kind == 0:
            assert (gx.view(np.uint32) == ox.view(np.uint32)).all() and (gy.view(np.uint32) == oy.view(np.uint32)).all()
        else:
            assert np.allclose(gx, ox, rtol=4e-6, atol=1e-4) and np.allclose(gy, oy, rtol=4e-6, atol=1e-4)
        src = rand_image(rng, h, w, 3, np.uint8)
        t = torch.from_numpy(src).cuda()
        out = torch.zeros((mh, mw, 3), dtype=torch.uint8, device="cuda")
        rc = capi.lib.b200_stitch_warp_dev(kind, kr.ctypes.data, tt.ctypes.data, scale, tl[0], tl[1], 16, t.data_ptr(), w * 3, w, h,
                                           out.data_ptr(), mw * 3, mw, mh, xm.data_ptr(), mw * 4, ym.data_ptr(), mw * 4, INTER_LINEAR, BORDER_REFLECT, None)
        assert rc == 0, capi.last_error()
        torch.cuda.synchronize()
        want = np.zeros((mh, mw, 3), np.uint8)
        gx2, gy2 = xm.cpu().numpy(), ym.cpu().numpy()
        assert oracle().remap32f(src, want, gx2, gy2, INTER_LINEAR, BORDER_REFLECT) == 0
        assert_same(out.cpu().numpy(), want, f"stitch warp kind {kind}")


def test_warppolar_forward_maps_bit_exact_and_image_equals_oracle():
    """b200_warpPolar_dev / b200_warpPolar_buildMaps_dev against the oracle's restatement of cv::warpPolar (imgwarp.cpp:3163-3224,
    pinned to the reference in tests/test_oracle_pinned.py): the maps bit for bit (the transcendental tables are host work in the
    reference and here), the image as any remap"""
    import opencv_b200 as cvb
    from opencv_b200 import capi
    rng = np.random.default_rng(41)
    n = 0
    for (dt, cn, (sw, sh), (dw, dh), (cx, cy), rad, flags) in warppolar_cases():
        src = rand_image(rng, sh, sw, cn, dt)
        want = np.full((dh, dw) if cn == 1 else (dh, dw, cn), 7, dt)
        assert oracle_warpPolar(src, want, cx, cy, rad, flags) == 0
        wmx, wmy = oracle_warpPolar_maps(dw, dh, cx, cy, rad, flags & WARP_POLAR_LOG)
        dst0 = torch.full(want.shape, 7, dtype=torch.from_numpy(want).dtype, device="cuda")
        got, (mx, my) = cvb.warpPolar(torch.from_numpy(src).cuda(), (dw, dh), (cx, cy), rad, flags, dst=dst0)
        assert (mx.cpu().numpy().view(np.uint32) == wmx.view(np.uint32)).all() and (my.cpu().numpy().view(np.uint32) == wmy.view(np.uint32)).all(), \
            f"maps {dw}x{dh} centre {cx},{cy} radius {rad} flags {flags}"
        assert_same(got.cpu().numpy(), want, f"warpPolar {dt.__name__}C{cn} {sw}x{sh}->{dw}x{dh} flags {flags}")
        n += 1
    assert n >= 75
    # a second geometry on the same thread replaces the cached tables; the first one again rebuilds them
    for (dw, dh, rad) in ((100, 200, 75.0), (64, 181, 120.0), (100, 200, 75.0)):
        mx = torch.zeros((dh, dw), dtype=torch.float32, device="cuda"); my = torch.zeros_like(mx)
        assert capi.lib.b200_warpPolar_buildMaps_dev(dw, dh, 80.0, 60.0, rad, WARP_POLAR_LOG, mx.data_ptr(), dw * 4, my.data_ptr(), dw * 4, None) == 0
        torch.cuda.synchronize()
        wmx, wmy = oracle_warpPolar_maps(dw, dh, 80.0, 60.0, rad, True)
        assert (mx.cpu().numpy().view(np.uint32) == wmx.view(np.uint32)).all() and (my.cpu().numpy().view(np.uint32) == wmy.view(np.uint32)).all()
    # dsize defaults (imgwarp.cpp:3168-3176) and the inverse direction
    t = torch.zeros((64, 64), dtype=torch.uint8, device="cuda")
    assert cvb.warpPolar(t, None, (32, 32), 31.0, INTER_LINEAR)[0].shape == (97, 31)
    assert cvb.warpPolar(t, (40, 0), (32, 32), 31.0, INTER_LINEAR)[0].shape == (126, 40)
    mx = torch.zeros((8, 8), dtype=torch.float32, device="cuda")
    assert capi.lib.b200_warpPolar_buildMaps_dev(8, 8, 1.0, 1.0, 4.0, WARP_INVERSE_MAP, mx.data_ptr(), 32, mx.data_ptr(), 32, None) == 1
