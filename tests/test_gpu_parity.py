"""GPU parity tests proper: the CUDA path, called through the C ABI (b200_hal_* with host buffers and
b200_*_dev with device buffers), against the pinned oracle on the same seeded inputs, the committed golden
fixtures from the reference, and -- at BASELINE.json's full sizes -- the oracle again plus size-independent
properties.  Bit-exact for 8U/16U/16S and all INTER_NEAREST; 32F is also required value-identical
(tolerance 0; north_star allows 1e-5 relative, we do not need it)."""
import ctypes as C

import numpy as np
import pytest

from common import *  # noqa: F403
from golden_io import load_golden, run_record

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def G():
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    from opencv_b200 import capi
    capi.lib.b200_hal_set_min_pixels(0)          # let the hooks take tiny test images
    names = {k: "b200_hal_" + k for k in SIGS}
    return HookLib(capi.lib, names)


O = oracle()


def test_library_is_the_cuda_extension(G):
    from opencv_b200 import capi
    before = capi.lib.b200_launch_count()
    src = np.arange(64 * 64, dtype=np.uint8).reshape(64, 64)
    dst = np.zeros((32, 32), np.uint8)
    assert G.resize(src, dst, INTER_AREA) == 0
    assert capi.lib.b200_launch_count() == before + 1     # a kernel of ours ran


def test_golden_fixtures_from_reference(G):
    recs = load_golden()
    for i, rec in enumerate(recs):
        got = run_record(G, rec)
        assert_same(got, rec["dst"], f"golden #{i} {rec['op']} {rec['params'][:2]}")


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.int16])
@pytest.mark.parametrize("cn", [1, 3, 4])
def test_warp_remap_matrix(G, dtype, cn):
    rng = np.random.default_rng(1000 + cn)
    sh, sw, dh, dw = 61, 83, 67, 90
    src = rand_image(rng, sh, sw, cn, dtype)
    dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
    bv = (10.4, 20.5, 300.7, -4)
    mx = rng.uniform(-30, sw + 30, (dh, dw)).astype(np.float32)
    my = rng.uniform(-30, sh + 30, (dh, dw)).astype(np.float32)
    mxy = np.ascontiguousarray(np.stack([mx, my], -1))
    m1, m2 = convert_maps_fixed(mx, my)
    for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4):
        for border in (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101):
            M = rotation_inverse(sw, sh, rng.uniform(-180, 180), rng.uniform(0.5, 1.6))
            H = rand_homography_inverse(rng, dw, dh, 0.3)
            calls = [
                ("warpAffine", lambda L, d: L.warpAffine(src, d, M, interp, border, bv)),
                ("warpPerspective", lambda L, d: L.warpPerspective(src, d, H, interp, border, bv)),
                ("remap32f", lambda L, d: L.remap32f(src, d, mx, my, interp, border, bv)),
                ("remap32fc2", lambda L, d: L.remap32fc2(src, d, mxy, interp, border, bv)),
                ("remap16s", lambda L, d: L.remap16s(src, d, m1, m2, interp, border, bv)),
            ]
            if interp == INTER_NEAREST:
                calls.append(("remap16s/no map2", lambda L, d: L.remap16s(src, d, m1, None, interp, border, bv)))
            for name, f in calls:
                a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
                assert f(G, a) == 0, name
                assert f(O, b) == 0
                assert_same(a, b, f"{name} {dtype.__name__} cn{cn} i{interp} b{border}")


@pytest.mark.parametrize("interp", [INTER_CUBIC, INTER_LANCZOS4])
def test_cubic_lanczos_edge_cases(G, interp):
    """sources narrower than the 4x4 / 8x8 tap window, 2 channels, relative maps, a 1080p frame"""
    rng = np.random.default_rng(31 + interp)
    for (sh, sw), dtype, cn in [((3, 5), np.uint8, 1), ((2, 9), np.float32, 2), ((7, 7), np.uint16, 3), ((1, 1), np.uint8, 4), ((8, 8), np.int16, 2)]:
        src = rand_image(rng, sh, sw, cn, dtype) if cn != 2 else np.ascontiguousarray(rand_image(rng, sh, sw, 3, dtype)[..., :2])
        dshape = (19, 23) if cn == 1 else (19, 23, cn)
        for border in (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101):
            M = rotation_inverse(sw, sh, 33, 0.4)
            a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
            assert G.warpAffine(src, a, M, interp, border, (5, 6, 7, 8)) == 0 and O.warpAffine(src, b, M, interp, border, (5, 6, 7, 8)) == 0
            assert_same(a, b, f"tiny warpAffine {sh}x{sw} {dtype.__name__} cn{cn} i{interp} b{border}")
    src = rand_image(rng, 64, 80, 3, np.uint8)
    mx = rng.uniform(-6, 6, (64, 80)).astype(np.float32); my = rng.uniform(-6, 6, (64, 80)).astype(np.float32)
    a, b = np.zeros_like(src), np.zeros_like(src)
    assert G.remap32f(src, a, mx, my, interp | WARP_RELATIVE_MAP, BORDER_REFLECT_101) == 0
    assert O.remap32f(src, b, mx, my, interp | WARP_RELATIVE_MAP, BORDER_REFLECT_101) == 0
    assert_same(a, b, f"relative remap i{interp}")
    for dtype in (np.uint8, np.float32):
        src = rand_image(rng, 1080, 1920, 3, dtype)
        M = rotation_inverse(1920, 1080, 15, 0.9)
        a, b = np.zeros_like(src), np.zeros_like(src)
        assert G.warpAffine(src, a, M, interp, BORDER_CONSTANT) == 0 and O.warpAffine(src, b, M, interp, BORDER_CONSTANT) == 0
        assert_same(a, b, f"1080p warpAffine {dtype.__name__} i{interp}")


def test_convert_maps_on_device_equals_oracle_and_feeds_remap(G):
    from opencv_b200 import capi
    O.lib.orc_convertMaps.argtypes = CONVERT_MAPS_SIG
    for case in convert_maps_cases(np.random.default_rng(3)):
        a1, a2 = convert_maps_outputs(case)
        assert call_convert_maps(O.lib.orc_convertMaps, case, a1, a2) == 0
        t1, m1, t2, m2, dt, nn = case
        dm1 = torch.from_numpy(m1).cuda(); dm2 = torch.from_numpy(m2.view(np.int16) if m2 is not None and m2.dtype == np.uint16 else m2).cuda() if m2 is not None else None
        g1, g2 = convert_maps_outputs(case)
        o1 = torch.from_numpy(g1).cuda(); o2 = torch.from_numpy(g2.view(np.int16) if g2.dtype == np.uint16 else g2).cuda()
        rc = capi.lib.b200_convertMaps_dev(t1, dm1.data_ptr(), m1.strides[0], t2, dm2.data_ptr() if dm2 is not None else None,
                                           m2.strides[0] if m2 is not None else 0, dt, o1.data_ptr(), g1.strides[0],
                                           o2.data_ptr(), g2.strides[0], m1.shape[1], m1.shape[0], nn, None)
        assert rc == 0, (capi.last_error(), case[0], case[4], case[5])
        torch.cuda.synchronize()
        r1 = o1.cpu().numpy(); r2 = o2.cpu().numpy().view(g2.dtype)
        assert_same(r1, a1, f"convertMaps map1 {t1}->{dt} nn{nn}")
        assert_same(r2, a2, f"convertMaps map2 {t1}->{dt} nn{nn}")
    # rejected by the reference -> declined
    z = torch.zeros((4, 4, 2), dtype=torch.int16, device="cuda"); f = torch.zeros((4, 4, 2), dtype=torch.float32, device="cuda")
    assert capi.lib.b200_convertMaps_dev(CV_16SC2, z.data_ptr(), 16, 0, None, 0, CV_32FC2, f.data_ptr(), 32, None, 0, 4, 4, 0, None) == 1
    assert capi.lib.b200_convertMaps_dev(CV_16SC2, z.data_ptr(), 16, 0, None, 0, CV_32FC1, f.data_ptr(), 16, None, 0, 4, 4, 1, None) == 1
    # maps converted on the device drive remap exactly like maps converted by the reference's rule on the host
    rng = np.random.default_rng(8)
    src = rand_image(rng, 200, 300, 3, np.uint8)
    mx = rng.uniform(-10, 310, (180, 260)).astype(np.float32); my = rng.uniform(-10, 210, (180, 260)).astype(np.float32)
    dmx, dmy = torch.from_numpy(mx).cuda(), torch.from_numpy(my).cuda()
    f1 = torch.empty((180, 260, 2), dtype=torch.int16, device="cuda"); f2 = torch.empty((180, 260), dtype=torch.int16, device="cuda")
    assert capi.lib.b200_convertMaps_dev(CV_32FC1, dmx.data_ptr(), 260 * 4, CV_32FC1, dmy.data_ptr(), 260 * 4, CV_16SC2,
                                         f1.data_ptr(), 260 * 4, f2.data_ptr(), 260 * 2, 260, 180, 0, None) == 0
    dsrc = torch.from_numpy(src).cuda(); ddst = torch.zeros((180, 260, 3), dtype=torch.uint8, device="cuda")
    bv = (C.c_double * 4)(1, 2, 3, 4)
    assert capi.lib.b200_remap16s_dev(16, dsrc.data_ptr(), 900, 300, 200, ddst.data_ptr(), 780, 260, 180, f1.data_ptr(), 260 * 4,
                                      f2.data_ptr(), 260 * 2, INTER_LINEAR, BORDER_CONSTANT, bv, None) == 0
    torch.cuda.synchronize()
    want = np.zeros((180, 260, 3), np.uint8)
    assert O.remap32f(src, want, mx, my, INTER_LINEAR, BORDER_CONSTANT, (1, 2, 3, 4)) == 0
    assert_same(ddst.cpu().numpy(), want, "remap through device-converted maps")


BORDER_TRANSPARENT = 5


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.int16])
@pytest.mark.parametrize("cn", [1, 3, 4])
def test_border_transparent_keeps_destination(G, dtype, cn):
    """BORDER_TRANSPARENT: pixels mapped outside keep the destination; the last row / column is blended from the taps
    that exist (imgwarp.cpp:726-755); bicubic / Lanczos4 use REFLECT_101 taps when the centre sample is inside."""
    rng = np.random.default_rng(500 + cn)
    sh, sw, dh, dw = 61, 83, 67, 90
    src = rand_image(rng, sh, sw, cn, dtype)
    mx = rng.uniform(-3, sw + 3, (dh, dw)).astype(np.float32); my = rng.uniform(-3, sh + 3, (dh, dw)).astype(np.float32)
    mx[::3, ::2] = np.round(mx[::3, ::2]); my[::2, ::3] = np.round(my[::2, ::3]); mx[5, :] = sw - 1; my[:, 7] = sh - 1
    m1, m2 = convert_maps_fixed(mx, my)
    for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4):
        mats = [rotation_inverse(sw, sh, 21, 0.8), np.array([[1, 0, 0.5], [0, 1, 0.25]]), np.array([[1, 0, 2.0], [0, 1, 1.0]])]
        calls = [(f"warpAffine{i}", lambda L, d, M=M: L.warpAffine(src, d, M, interp, BORDER_TRANSPARENT)) for i, M in enumerate(mats)]
        calls.append(("warpPerspective", lambda L, d: L.warpPerspective(src, d, rand_homography_inverse(np.random.default_rng(1), dw, dh, 0.3), interp, BORDER_TRANSPARENT)))
        calls.append(("remap32f", lambda L, d: L.remap32f(src, d, mx, my, interp, BORDER_TRANSPARENT)))
        calls.append(("remap16s", lambda L, d: L.remap16s(src, d, m1, m2, interp, BORDER_TRANSPARENT)))
        for name, f in calls:
            init = rand_image(rng, dh, dw, cn, dtype)
            a, b = init.copy(), init.copy()
            assert f(G, a) == 0 and f(O, b) == 0
            assert_same(a, b, f"transparent {name} {dtype.__name__} cn{cn} i{interp}")
            assert (a == init).any() and (a != init).any()      # some pixels kept, some written
    # a strided destination ROI keeps its padding and its untouched pixels
    big = rand_image(rng, dh + 4, dw + 6, cn, dtype)
    ref = big.copy()
    roi_a, roi_b = big[2:2 + dh, 3:3 + dw], ref[2:2 + dh, 3:3 + dw]
    M = rotation_inverse(sw, sh, 21, 0.8)
    assert G.warpAffine(src, roi_a, M, INTER_LINEAR, BORDER_TRANSPARENT) == 0 and O.warpAffine(src, roi_b, M, INTER_LINEAR, BORDER_TRANSPARENT) == 0
    assert_same(big, ref, "transparent into a ROI")


RESIZE_SIZES = [((97, 131), (113, 150)), ((97, 131), (48, 65)), ((96, 132), (48, 66)), ((96, 132), (32, 44)),
                ((96, 132), (24, 66)), ((96, 132), (48, 33)), ((97, 131), (40, 50)), ((97, 131), (200, 300)),
                ((97, 131), (97, 60)), ((97, 131), (150, 70)), ((97, 131), (50, 200)), ((64, 64), (16, 16)),
                ((1, 50), (1, 20)), ((50, 1), (20, 1)), ((5, 5), (11, 13)), ((100, 100), (7, 9)),
                ((300, 10), (7, 4))]


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.int16])
@pytest.mark.parametrize("cn", [1, 2, 3, 4])
def test_resize_matrix(G, dtype, cn):
    rng = np.random.default_rng(2000 + cn)
    for (sh, sw), (dh, dw) in RESIZE_SIZES:
        src = rand_image(rng, sh, sw, cn, dtype)
        dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        exact = (INTER_NEAREST_EXACT,) + ((INTER_LINEAR_EXACT,) if dtype in (np.uint8, np.uint16, np.int16) else ())
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA, INTER_CUBIC, INTER_LANCZOS4) + exact:
            a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
            assert G.resize(src, a, interp) == 0 and O.resize(src, b, interp) == 0
            assert_same(a, b, f"resize {dtype.__name__} cn{cn} {sh}x{sw}->{dh}x{dw} i{interp}")
    for (sh, sw), f in [((27, 35), 0.5), ((35, 27), 0.5), ((31, 37), 1 / 3.), ((29, 35), 0.25)]:
        src = rand_image(rng, sh, sw, cn, dtype)
        dw, dh = int(np.rint(sw * f)), int(np.rint(sh * f))
        dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA):
            a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
            assert G.resize(src, a, interp, (f, f)) == 0 and O.resize(src, b, interp, (f, f)) == 0
            assert_same(a, b, f"ragged resize {dtype.__name__} cn{cn} {sh}x{sw} f={f:.3f} i{interp}")


def test_lanczos4_through_the_device_resident_entry(G):
    """b200_resize_dev INTER_LANCZOS4: the host-built coefficient tables are cached on the device per geometry; two geometries,
    two calls each (the second call of a geometry re-uses the cached table), several depths"""
    import opencv_b200 as cvb
    rng = np.random.default_rng(31)
    for dtype, cn in ((np.uint8, 3), (np.uint16, 1), (np.float32, 4)):
        for (sh, sw), (dh, dw) in (((90, 120), (135, 77)), ((64, 200), (50, 333))):
            for _rep in range(2):
                src = rand_image(rng, sh, sw, cn, dtype)
                want = np.zeros((dh, dw) if cn == 1 else (dh, dw, cn), dtype)
                assert O.resize(src, want, INTER_LANCZOS4) == 0
                got = cvb.resize(torch.from_numpy(src).cuda(), (dw, dh), interpolation=INTER_LANCZOS4).cpu().numpy()
                assert_same(got, want, f"device LANCZOS4 {dtype.__name__} cn{cn} {sw}x{sh}->{dw}x{dh}")


def test_64f_warp_remap_resize(G):
    """CV_64F INTER_LINEAR / CUBIC / LANCZOS4 in the warp family, the same and INTER_AREA in resize: float table weights / coefficients, double products and sums (imgwarp.cpp:1546, resize.cpp linear_tab);
    every border but TRANSPARENT, three map formats, up- and downscale"""
    rng = np.random.default_rng(64)
    for cn in (1, 3, 4):
        src = (rng.random((97, 131, cn)) * 2000 - 1000).squeeze()
        shape = (110, 140) if cn == 1 else (110, 140, cn)
        M = rotation_inverse(131, 97, 20, 0.9)
        Hm = rand_homography_inverse(rng, 131, 97)
        mx = rng.uniform(-10, 140, (110, 140)).astype(np.float32); my = rng.uniform(-10, 105, (110, 140)).astype(np.float32)
        m1, m2 = convert_maps_fixed(mx, my)
        for border in (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101):
            for interp in (INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4):
                for name, call in (("warpAffine", lambda L, d: L.warpAffine(src, d, M, interp, border, (1.5, 2.5, 3.5, 4.5))),
                                   ("warpPerspective", lambda L, d: L.warpPerspective(src, d, Hm, interp, border, (1.5, 2.5, 3.5, 4.5))),
                                   ("remap32f", lambda L, d: L.remap32f(src, d, mx, my, interp, border, (1.5, 2.5, 3.5, 4.5))),
                                   ("remap16s", lambda L, d: L.remap16s(src, d, m1, m2, interp, border, (1.5, 2.5, 3.5, 4.5)))):
                    a, b = np.zeros(shape), np.zeros(shape)
                    assert call(G, a) == 0 and call(O, b) == 0
                    assert (a == b).all(), f"64F {name} interp{interp} cn{cn} border{border}: {(a != b).sum()} values differ"
        for (dh, dw) in ((150, 200), (50, 77), (97, 262), (48, 65), (32, 43), (24, 33), (97, 131)):   # incl. /2 and /3 with ragged edges, /4
            for interp in (INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4, INTER_AREA):
                a = np.zeros((dh, dw) if cn == 1 else (dh, dw, cn)); b = np.zeros_like(a)
                if (dh, dw) == (97, 131):
                    a[:] = 1; b[:] = 1          # equal sizes never reach the hook in cv::resize; both sides still agree
                assert G.resize(src, a, interp) == 0 and O.resize(src, b, interp) == 0
                assert (a == b).all(), f"64F resize interp{interp} cn{cn} ->{dw}x{dh}"
    a = np.zeros((40, 50))
    assert G.warpAffine(np.zeros((40, 50)), a, np.eye(2, 3), INTER_LINEAR, 5) == 1                    # not with BORDER_TRANSPARENT


def test_nearest_any_depth(G):
    # NEAREST is a pure element copy for every depth (imgwarp.cpp:268-369, resize.cpp:1026-1172)
    rng = np.random.default_rng(7)
    for dt in (np.int8, np.int32, np.float64):
        src = rng.integers(-100, 100, (40, 50, 3)).astype(dt)
        a = np.zeros((33, 47, 3), dt)
        # no numpy->cv type for these in tests.common; call through the package API instead
        import opencv_b200 as cvb
        out = cvb.resize(src, (47, 33), interpolation=INTER_NEAREST)
        sx = np.minimum(np.floor(np.arange(47) * (1.0 / (47 / 50))).astype(int), 49)
        sy = np.minimum(np.floor(np.arange(33) * (1.0 / (33 / 40))).astype(int), 39)
        assert (out == src[sy][:, sx]).all()
    src5 = np.zeros((12, 12, 5), np.uint8)
    import opencv_b200 as cvb
    with pytest.raises(NotImplementedError):
        cvb.resize(src5, (6, 6), interpolation=INTER_LINEAR)         # 2x2 decimation of > 4 channels: hook declines -> OpenCV falls back


@pytest.mark.parametrize("dt", [np.int8, np.int32])
def test_linear_exact_on_signed_8_and_32_bit_data(G, dt):
    """INTER_LINEAR_EXACT for CV_8S (fixedpoint32 like CV_16S, resize.cpp:66-71) and CV_32S (fixedpoint64 with its magnitude-rounded
    128-bit product); 16S runs in test_resize_matrix.  Extreme values exercise the saturating operators."""
    rng = np.random.default_rng(88)
    info = np.iinfo(dt)
    for cn in (1, 3, 4):
        for (sh, sw), (dh, dw) in (((48, 64), (47, 33)), ((21, 37), (96, 128)), ((9, 130), (130, 9)), ((300, 400), (170, 555))):
            src = rng.integers(info.min, info.max + 1, (sh, sw, cn)).astype(dt)
            src[::5, ::3] = info.min; src[1::7, 1::4] = info.max
            if cn == 1:
                src = src.reshape(sh, sw)
            a = np.zeros((dh, dw) if cn == 1 else (dh, dw, cn), dt); b = np.zeros_like(a)
            assert G.resize(src, a, INTER_LINEAR_EXACT) == 0 and O.resize(src, b, INTER_LINEAR_EXACT) == 0
            assert_same(a, b, f"LINEAR_EXACT {dt.__name__} cn{cn} {sh}x{sw}->{dh}x{dw}")


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
def test_warps_with_more_than_four_channels(G, dtype):
    """cn > 4 through the warp hooks (the reference's Imgproc_Warp.multichannel runs up to 14 channels): groups of 4 channels, each
    identical to the oracle on that group with the same border value (channel c takes borderValue[c & 3], imgwarp.cpp:278-279)"""
    rng = np.random.default_rng(55)
    sh, sw, dh, dw = 97, 131, 110, 140
    M = rotation_inverse(sw, sh, 20, 0.9)
    Hm = rand_homography_inverse(rng, sw, sh)
    mx = rng.uniform(-10, sw + 10, (dh, dw)).astype(np.float32); my = rng.uniform(-10, sh + 10, (dh, dw)).astype(np.float32)
    bv = (1.5, 22, 3, 47)
    for cn in (5, 6, 9, 14):
        src = np.stack([rand_image(rng, sh, sw, 1, dtype) for _ in range(cn)], -1)
        for interp, border in ((INTER_NEAREST, BORDER_CONSTANT), (INTER_LINEAR, BORDER_CONSTANT), (INTER_LINEAR, BORDER_REFLECT_101), (INTER_AREA, BORDER_WRAP)):
            for name, call in (("warpAffine", lambda L, s_, d_: L.warpAffine(s_, d_, M, interp, border, bv)),
                               ("warpPerspective", lambda L, s_, d_: L.warpPerspective(s_, d_, Hm, interp, border, bv)),
                               ("remap32f", lambda L, s_, d_: L.remap32f(s_, d_, mx, my, interp, border, bv))):
                got = np.zeros((dh, dw, cn), dtype)
                assert call(G, src, got) == 0
                for c0 in range(0, cn, 4):
                    k = min(4, cn - c0)
                    sg = np.ascontiguousarray(src[..., c0:c0 + k]); sg = sg.reshape(sh, sw) if k == 1 else sg
                    want = np.zeros((dh, dw) if k == 1 else (dh, dw, k), dtype)
                    assert call(O, sg, want) == 0
                    assert_same(got[..., c0:c0 + k].reshape(want.shape), want, f"{name} cn{cn} channels {c0}..{c0 + k - 1} interp{interp} border{border}")
    dst = np.full((dh, dw, 5), 9, dtype)
    assert G.warpAffine(src[..., :5].copy(), dst, M, INTER_LINEAR, 5, bv) == 1 and (dst == 9).all()      # BORDER_TRANSPARENT with > 4 channels: CPU path
    assert G.remap32f(src[..., :5].copy(), dst, mx, my, INTER_CUBIC, BORDER_CONSTANT, bv) == 1                # the reference asserts cn <= 4 for CUBIC / LANCZOS4


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.int16, np.float32, np.float64])
def test_resize_with_more_than_four_channels(G, dtype):
    """cn > 4 through the resize hook: groups of 4 channels, each identical to the oracle on that group, in every mode where the
    reference treats channels alike whatever their number (pinned in tests/test_oracle_pinned.py); the other cells are declined"""
    rng = np.random.default_rng(56)
    n = 0
    for cn in (5, 6, 9, 14):
        for (sh, sw, dh, dw) in ((61, 83, 70, 97), (64, 96, 32, 48), (90, 120, 30, 40), (90, 120, 31, 47), (60, 80, 120, 160)):
            src = (rng.random((sh, sw, cn)) * 1000 - 300) if dtype == np.float64 else np.stack([rand_image(rng, sh, sw, 1, dtype) for _ in range(cn)], -1)
            for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4, INTER_AREA, INTER_LINEAR_EXACT, INTER_NEAREST_EXACT):
                if interp == INTER_LINEAR_EXACT and dtype in (np.float32, np.float64):
                    continue
                got = np.full((dh, dw, cn), 7, dtype)
                rc = G.resize(src, got, interp)
                if resize_cn_gt4_declined(dtype, interp, sw, sh, dw, dh):
                    assert rc == 1 and (got == 7).all(), f"cn{cn} interp{interp} {sw}x{sh}->{dw}x{dh} should be left to the CPU"
                    continue
                assert rc == 0, f"cn{cn} interp{interp} {sw}x{sh}->{dw}x{dh}"
                for c0 in range(0, cn, 4):
                    k = min(4, cn - c0)
                    sg = np.ascontiguousarray(src[..., c0:c0 + k]); sg = sg.reshape(sh, sw) if k == 1 else sg
                    want = np.zeros((dh, dw) if k == 1 else (dh, dw, k), dtype)
                    assert O.resize(sg, want, interp) == 0
                    assert_same(np.ascontiguousarray(got[..., c0:c0 + k]).reshape(want.shape), want, f"resize cn{cn} channels {c0}..{c0 + k - 1} interp{interp} {sw}x{sh}->{dw}x{dh}")
                n += 1
    assert n > 50


def test_unsupported_returns_not_implemented_without_touching_dst(G):
    src = np.zeros((40, 40, 3), np.uint8)
    dst = np.full((40, 40, 3), 9, np.uint8)
    M = np.eye(2, 3)
    assert G.warpAffine(src, dst, M, 7, BORDER_CONSTANT) == 1          # not an interpolation mode (INTER_MAX)
    assert G.warpAffine(src, dst, M, INTER_LINEAR, 6) == 1              # BORDER_ISOLATED: not a border mode of the warps
    assert G.resize(np.zeros((40, 40), np.int32), np.zeros((17, 23), np.int32), INTER_LINEAR) == 1         # 32S: only NEAREST and LINEAR_EXACT
    assert (dst == 9).all()
    from opencv_b200 import capi
    old = capi.lib.b200_hal_set_min_pixels(1 << 16)
    try:
        assert G.warpAffine(src, dst, M, INTER_LINEAR, BORDER_CONSTANT) == 1   # below the re-entrancy gate
    finally:
        capi.lib.b200_hal_set_min_pixels(0)
    assert old == 0


def test_strided_roi_views(G):
    # dst and src as ROIs of larger buffers (non-contiguous step), as cv::Mat ROIs reach the hooks
    rng = np.random.default_rng(11)
    big = rand_image(rng, 120, 160, 3, np.uint8)
    src = big[10:100, 20:140]
    out_big = np.full((130, 170, 3), 77, np.uint8)
    ref_big = out_big.copy()
    M = rotation_inverse(120, 90, 20, 1.1)
    assert G.warpAffine(src, out_big[5:105, 7:157], M, INTER_LINEAR, BORDER_REFLECT_101) == 0
    assert O.warpAffine(src, ref_big[5:105, 7:157], M, INTER_LINEAR, BORDER_REFLECT_101) == 0
    assert_same(out_big, ref_big, "ROI warpAffine (and nothing outside the ROI written)")


# ------------------------------------------------------------------ BASELINE configs, full size ----
def _c1(w, h):
    rng = np.random.default_rng(12345)
    src = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    return src, rotation_inverse(w, h, 15.0, 0.9, cx=w / 2, cy=h / 2)


@pytest.mark.parametrize("w,h", [(1920, 1080), (3840, 2160)])
def test_config1_warpaffine_full_size(G, w, h):
    src, M = _c1(w, h)
    a, b = np.zeros_like(src), np.zeros_like(src)
    assert G.warpAffine(src, a, M, INTER_LINEAR, BORDER_CONSTANT) == 0
    assert O.warpAffine(src, b, M, INTER_LINEAR, BORDER_CONSTANT) == 0
    assert_same(a, b, f"C1 {w}x{h}")


def test_config2_resize_full_size(G):
    rng = np.random.default_rng(2)
    src = rng.integers(0, 256, (2160, 3840, 3), dtype=np.uint8)
    for interp in (INTER_LINEAR, INTER_AREA):
        a, b = np.zeros((1080, 1920, 3), np.uint8), np.zeros((1080, 1920, 3), np.uint8)
        assert G.resize(src, a, interp) == 0 and O.resize(src, b, interp) == 0
        assert_same(a, b, f"C2a interp {interp}")
        s = src.astype(np.uint16)                      # property: exact 2x == (a+b+c+d+2)>>2
        box = ((s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2).astype(np.uint8)
        assert (a == box).all()
    for (dw, dh) in ((1600, 900), (1280, 720)):        # C2c: fractional and 3x3 area, generic linear
        for interp in (INTER_AREA, INTER_LINEAR):
            a, b = np.zeros((dh, dw, 3), np.uint8), np.zeros((dh, dw, 3), np.uint8)
            assert G.resize(src, a, interp) == 0 and O.resize(src, b, interp) == 0
            assert_same(a, b, f"C2c {dw}x{dh} interp {interp}")
    f = rng.random((1080, 1920), dtype=np.float32)     # C2b
    a, b = np.zeros((2160, 3840), np.float32), np.zeros((2160, 3840), np.float32)
    assert G.resize(f, a, INTER_LINEAR) == 0 and O.resize(f, b, INTER_LINEAR) == 0
    assert_same(a, b, "C2b 32FC1 2x upscale")
    for interp in (INTER_CUBIC, INTER_LANCZOS4):       # full-size runs of the 4- and 8-tap modes (vector body + scalar tail widths)
        a, b = np.zeros((1080, 1919, 3), np.uint8), np.zeros((1080, 1919, 3), np.uint8)
        assert G.resize(src, a, interp) == 0 and O.resize(src, b, interp) == 0
        assert_same(a, b, f"4K -> 1919x1080 interp {interp}")
        a, b = np.zeros((2160, 3838), np.float32), np.zeros((2160, 3838), np.float32)
        assert G.resize(f, a, interp) == 0 and O.resize(f, b, interp) == 0
        assert_same(a, b, f"32FC1 upscale interp {interp}")


def test_config3_warpperspective_full_size(G):
    rng = np.random.default_rng(3)
    w, h = 3840, 2160
    src = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
    # inverse of SURVEY 8(d) C3's forward homography (numpy inverse; both sides get the same matrix)
    Hf = np.array([[1.12085816, 0.0203087925, 97], [0.0422569112, 1.05413990, -64],
                   [4.77793919e-05, -3.76190421e-06, 1]])
    Hi = np.linalg.inv(Hf)
    a, b = np.zeros_like(src), np.zeros_like(src)
    assert G.warpPerspective(src, a, Hi, INTER_LINEAR, BORDER_REPLICATE) == 0
    assert O.warpPerspective(src, b, Hi, INTER_LINEAR, BORDER_REPLICATE) == 0
    assert_same(a, b, "C3")


def test_config4_remap_fisheye_three_map_formats(G):
    rng = np.random.default_rng(4)
    w, h = 1280, 720
    src = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    mx, my = fisheye_maps(w, h)
    mxy = np.ascontiguousarray(np.stack([mx, my], -1))
    m1, m2 = convert_maps_fixed(mx, my)
    outs = []
    for name, f in (("32f", lambda L, d: L.remap32f(src, d, mx, my, INTER_LINEAR, BORDER_REFLECT_101)),
                    ("32fc2", lambda L, d: L.remap32fc2(src, d, mxy, INTER_LINEAR, BORDER_REFLECT_101)),
                    ("16s", lambda L, d: L.remap16s(src, d, m1, m2, INTER_LINEAR, BORDER_REFLECT_101))):
        a, b = np.zeros_like(src), np.zeros_like(src)
        assert f(G, a) == 0 and f(O, b) == 0
        assert_same(a, b, "C4 " + name)
        outs.append(a)
    assert_same(outs[0], outs[1], "planar vs interleaved float maps")
    assert_same(outs[0], outs[2], "float maps vs convertMaps fixed-point maps")


# ------------------------------------------------------------------ device-resident + batched ----
def _t(a):
    return torch.from_numpy(a).cuda()


def test_device_resident_api_matches_host_hooks(G):
    import opencv_b200 as cvb
    rng = np.random.default_rng(21)
    src = rand_image(rng, 200, 300, 3, np.uint8)
    M = rotation_inverse(300, 200, 33, 0.8)
    Mf = np.linalg.inv(np.vstack([M, [0, 0, 1]]))[:2]           # forward matrix for the public API
    host = cvb.warpAffine(src, M, (320, 210), flags=cvb.INTER_LINEAR | cvb.WARP_INVERSE_MAP, borderMode=cvb.BORDER_WRAP)
    dev = cvb.warpAffine(_t(src), M, (320, 210), flags=cvb.INTER_LINEAR | cvb.WARP_INVERSE_MAP, borderMode=cvb.BORDER_WRAP)
    assert_same(dev.cpu().numpy(), host, "device vs host warpAffine")
    ref = np.zeros_like(host)
    assert O.warpAffine(src, ref, M, INTER_LINEAR, BORDER_WRAP) == 0
    assert_same(host, ref, "host hook vs oracle")
    fwd = cvb.warpAffine(src, Mf, (320, 210))                   # exercises the host-side inversion
    assert fwd.shape == (210, 320, 3)
    for interp in (cvb.INTER_NEAREST, cvb.INTER_LINEAR, cvb.INTER_AREA):
        r_h = cvb.resize(src, (123, 77), interpolation=interp)
        r_d = cvb.resize(_t(src), (123, 77), interpolation=interp)
        assert_same(r_d.cpu().numpy(), r_h, f"device vs host resize {interp}")
    assert_same(cvb.resize(src, None, 0.5, 0.5, cvb.INTER_AREA), cvb.resize(src, (150, 100), interpolation=cvb.INTER_AREA), "fx/fy vs dsize")
    mx = rng.uniform(-5, 305, (100, 140)).astype(np.float32)
    my = rng.uniform(-5, 205, (100, 140)).astype(np.float32)
    assert_same(cvb.remap(_t(src), _t(mx), _t(my), cvb.INTER_LINEAR, cvb.BORDER_REFLECT_101).cpu().numpy(),
                cvb.remap(src, mx, my, cvb.INTER_LINEAR, cvb.BORDER_REFLECT_101), "device vs host remap")
    Hm = rand_homography_inverse(rng, 300, 200, 0.2)
    assert_same(cvb.warpPerspective(_t(src), Hm, flags=cvb.INTER_LINEAR | cvb.WARP_INVERSE_MAP).cpu().numpy(),
                cvb.warpPerspective(src, Hm, flags=cvb.INTER_LINEAR | cvb.WARP_INVERSE_MAP), "device vs host warpPerspective")


def test_batched_pipeline_equals_per_frame_calls(G):
    from opencv_b200 import capi
    n, sw, sh, mw, mh = 5, 192, 108, 96, 54
    step_s, step_m = 640, 320                                     # padded rows
    src = torch.zeros((n, sh, step_s), dtype=torch.uint8, device="cuda")
    assert capi.lib.b200_synth_frames_dev(src.data_ptr(), sh * step_s, step_s, sw, sh, 3, 100, n, 2026, None) == 0
    mid = torch.zeros((n, mh, step_m), dtype=torch.uint8, device="cuda")
    dst = torch.zeros((n, mh, step_m), dtype=torch.uint8, device="cuda")
    M = rotation_inverse(mw, mh, 15, 0.9, cx=mw / 2, cy=mh / 2)
    Mc = (C.c_double * 6)(*M.ravel())
    bv = (C.c_double * 4)(0, 0, 0, 0)
    rc = capi.lib.b200_resize_warpAffine_batch_dev(16, src.data_ptr(), sh * step_s, step_s, sw, sh,
                                                   mid.data_ptr(), mh * step_m, step_m, mw, mh,
                                                   dst.data_ptr(), mh * step_m, step_m, mw, mh,
                                                   n, INTER_LINEAR, Mc, INTER_LINEAR, BORDER_CONSTANT, bv, None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    src_h, dst_h = src.cpu().numpy(), dst.cpu().numpy()
    # synthetic frames: deterministic, frame-indexed (a second generation of frame 102 alone matches)
    one = torch.zeros((1, sh, step_s), dtype=torch.uint8, device="cuda")
    assert capi.lib.b200_synth_frames_dev(one.data_ptr(), sh * step_s, step_s, sw, sh, 3, 102, 1, 2026, None) == 0
    assert (one.cpu().numpy()[0] == src_h[2]).all()
    for f in range(n):
        frame = np.ascontiguousarray(src_h[f, :, :sw * 3]).reshape(sh, sw, 3)
        m = np.zeros((mh, mw, 3), np.uint8)
        o = np.zeros((mh, mw, 3), np.uint8)
        assert O.resize(frame, m, INTER_LINEAR) == 0
        assert O.warpAffine(m, o, M, INTER_LINEAR, BORDER_CONSTANT) == 0
        got = dst_h[f, :, :mw * 3].reshape(mh, mw, 3)
        assert_same(np.ascontiguousarray(got), o, f"pipeline frame {f}")


@pytest.mark.parametrize("dtype,cn", [(np.uint8, 3), (np.uint8, 1), (np.uint16, 4), (np.float32, 1)])
@pytest.mark.parametrize("chunk", [0, 1, 4])
def test_host_frame_stream_equals_per_frame_oracle(G, dtype, cn, chunk):
    """b200_hal_resize_warpAffine_batch: n host frames, chunks that do not divide n, more chunks than staging slots."""
    import opencv_b200 as cvb
    rng = np.random.default_rng(77 + cn)
    n, sw, sh, mw, mh, dw, dh = 11, 256, 144, 128, 72, 120, 80
    frames = np.stack([rand_image(rng, sh, sw, cn, dtype) for _ in range(n)])
    if cn == 1:
        frames = frames.reshape(n, sh, sw)
    M = rotation_inverse(mw, mh, -20, 1.1, cx=mw / 2, cy=mh / 2)
    for border in (BORDER_CONSTANT, BORDER_REFLECT_101):
        got = cvb.resize_warpAffine_frames(frames, (mw, mh), M, (dw, dh), flags=INTER_LINEAR | WARP_INVERSE_MAP,
                                           borderMode=border, borderValue=(9, 8, 7, 6), chunk_frames=chunk)
        for f in range(n):
            m = np.zeros((mh, mw) + (() if cn == 1 else (cn,)), dtype)
            o = np.zeros((dh, dw) + (() if cn == 1 else (cn,)), dtype)
            assert O.resize(frames[f], m, INTER_LINEAR) == 0
            assert O.warpAffine(m, o, M, INTER_LINEAR, border, (9, 8, 7, 6)) == 0
            assert_same(got[f], o, f"host stream frame {f} border {border}")


def test_host_frame_stream_honours_strides_and_leaves_padding_alone(G):
    """ROI-style frames: rows and frames padded on both sides; bytes between dst rows / frames are not written."""
    from opencv_b200 import capi
    rng = np.random.default_rng(5)
    n, sw, sh, mw, mh = 7, 200, 100, 100, 50
    s_step, s_fs = 200 * 3 + 40, (200 * 3 + 40) * 100 + 1000          # 2-D copy path (step not within the block rule)
    d_step, d_fs = 100 * 3 + 20, (100 * 3 + 20) * 50 + 333
    sbuf = rng.integers(0, 256, n * s_fs, dtype=np.uint8)
    dbuf = np.full(n * d_fs, 0xA5, np.uint8)
    M = rotation_inverse(mw, mh, 7, 0.95, cx=mw / 2, cy=mh / 2)
    Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
    for chunk in (0, 2):
        dbuf[:] = 0xA5
        rc = capi.lib.b200_hal_resize_warpAffine_batch(16, sbuf.ctypes.data, s_fs, s_step, sw, sh, mw, mh,
                                                       dbuf.ctypes.data, d_fs, d_step, mw, mh, n, INTER_AREA, Mc,
                                                       INTER_LINEAR, BORDER_REPLICATE, bv, chunk)
        assert rc == 0, capi.last_error()
        written = np.zeros(n * d_fs, bool)
        for f in range(n):
            frame = np.lib.stride_tricks.as_strided(sbuf[f * s_fs:], (sh, sw, 3), (s_step, 3, 1))
            m = np.zeros((mh, mw, 3), np.uint8); o = np.zeros((mh, mw, 3), np.uint8)
            assert O.resize(np.ascontiguousarray(frame), m, INTER_AREA) == 0
            assert O.warpAffine(m, o, M, INTER_LINEAR, BORDER_REPLICATE) == 0
            got = np.lib.stride_tricks.as_strided(dbuf[f * d_fs:], (mh, mw, 3), (d_step, 3, 1))
            assert_same(np.ascontiguousarray(got), o, f"strided host stream frame {f}")
            for y in range(mh):
                written[f * d_fs + y * d_step: f * d_fs + y * d_step + mw * 3] = True
        assert (dbuf[~written] == 0xA5).all(), "bytes outside the destination rows were overwritten"
    # argument checks: frames overlapping their stride, zero frames
    assert capi.lib.b200_hal_resize_warpAffine_batch(16, sbuf.ctypes.data, 10, s_step, sw, sh, mw, mh, dbuf.ctypes.data,
                                                     d_fs, d_step, mw, mh, n, INTER_AREA, Mc, INTER_LINEAR, 1, bv, 0) == 1
    assert capi.lib.b200_hal_resize_warpAffine_batch(16, sbuf.ctypes.data, s_fs, s_step, sw, sh, mw, mh, dbuf.ctypes.data,
                                                     d_fs, d_step, mw, mh, 0, INTER_AREA, Mc, INTER_LINEAR, 1, bv, 0) == 0


@pytest.mark.parametrize("chunk", [0, 3])
def test_host_frame_streams_of_the_single_operations(G, chunk):
    """b200_hal_resize_batch / warpAffine_batch / warpPerspective_batch / remap_batch (BASELINE configs 2-4 as frame streams): every
    frame equals the per-frame oracle; one map in three formats; 16U and 32F frames as well"""
    import opencv_b200 as cvb
    from opencv_b200 import capi
    rng = np.random.default_rng(123)
    n, w, h, dw, dh = 10, 200, 120, 170, 110
    bvv = (5, 6, 7, 8)
    for dtype, cn in ((np.uint8, 3), (np.uint16, 1), (np.float32, 4)):
        frames = np.stack([rand_image(rng, h, w, cn, dtype) for _ in range(n)])
        if cn == 1:
            frames = frames.reshape(n, h, w)
        oshape = (dh, dw) + (() if cn == 1 else (cn,))
        Hm = rand_homography_inverse(rng, w, h)
        got = cvb.warpPerspective_frames(frames, Hm, (dw, dh), flags=INTER_LINEAR | WARP_INVERSE_MAP, borderMode=BORDER_REPLICATE, chunk_frames=chunk)
        for f in range(n):
            o = np.zeros(oshape, dtype)
            assert O.warpPerspective(frames[f], o, Hm, INTER_LINEAR, BORDER_REPLICATE) == 0
            assert_same(got[f], o, f"warpPerspective stream frame {f}")
        mx = rng.uniform(-5, w + 5, (dh, dw)).astype(np.float32); my = rng.uniform(-5, h + 5, (dh, dw)).astype(np.float32)
        m1, m2 = convert_maps_fixed(mx, my)
        for (a1, a2, call) in ((mx, my, lambda d, f: O.remap32f(frames[f], d, mx, my, INTER_LINEAR, BORDER_REFLECT_101, bvv)),
                               (np.ascontiguousarray(np.stack([mx, my], -1)), None, lambda d, f: O.remap32f(frames[f], d, mx, my, INTER_LINEAR, BORDER_REFLECT_101, bvv)),
                               (m1, m2, lambda d, f: O.remap16s(frames[f], d, m1, m2, INTER_LINEAR, BORDER_REFLECT_101, bvv))):
            got = cvb.remap_frames(frames, a1, a2, INTER_LINEAR, BORDER_REFLECT_101, bvv, chunk_frames=chunk)
            for f in range(n):
                o = np.zeros(oshape, dtype)
                assert call(o, f) == 0
                assert_same(got[f], o, f"remap stream frame {f}")
        # resize and warpAffine streams through the C entries
        esz = frames.itemsize * cn
        M = rotation_inverse(w, h, 12, 0.93)
        Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(*bvv)
        out = np.zeros((n,) + oshape, dtype)
        assert capi.lib.b200_hal_warpAffine_batch(cv_type(frames[0]), frames.ctypes.data, frames.strides[0], frames.strides[1], w, h, out.ctypes.data,
                                                  out.strides[0], out.strides[1], dw, dh, n, Mc, INTER_LINEAR, BORDER_CONSTANT, bv, chunk) == 0, capi.last_error()
        for f in range(n):
            o = np.zeros(oshape, dtype)
            assert O.warpAffine(frames[f], o, M, INTER_LINEAR, BORDER_CONSTANT, bvv) == 0
            assert_same(out[f], o, f"warpAffine stream frame {f}")
        out[:] = 0
        assert capi.lib.b200_hal_resize_batch(cv_type(frames[0]), frames.ctypes.data, frames.strides[0], frames.strides[1], w, h, out.ctypes.data,
                                              out.strides[0], out.strides[1], dw, dh, n, 0.0, 0.0, INTER_LINEAR, chunk) == 0, capi.last_error()
        for f in range(n):
            o = np.zeros(oshape, dtype)
            assert O.resize(frames[f], o, INTER_LINEAR) == 0
            assert_same(out[f], o, f"resize stream frame {f}")
    assert capi.lib.b200_hal_warpAffine_batch(16, frames.ctypes.data, 1, 1, w, h, out.ctypes.data, 1, 1, dw, dh, n, Mc, INTER_LINEAR, 5, bv, 0) == 1   # TRANSPARENT


def test_batched_perspective_and_remap_equal_per_frame_calls(G):
    from opencv_b200 import capi
    rng = np.random.default_rng(41)
    n, sh, sw, dh, dw = 5, 70, 96, 64, 88
    frames = np.stack([rand_image(rng, sh, sw, 3, np.uint8) for _ in range(n)])
    dsrc = torch.from_numpy(frames).cuda()
    bv = (C.c_double * 4)(3, 4, 5, 6)
    H = rand_homography_inverse(rng, dw, dh, 0.25)
    Hc = (C.c_double * 9)(*H.ravel())
    out = torch.zeros((n, dh, dw, 3), dtype=torch.uint8, device="cuda")
    assert capi.lib.b200_warpPerspective_batch_dev(16, dsrc.data_ptr(), sh * sw * 3, sw * 3, sw, sh, out.data_ptr(), dh * dw * 3, dw * 3, dw, dh,
                                                   n, Hc, INTER_LINEAR, BORDER_REPLICATE, bv, None) == 0
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for f in range(n):
        want = np.zeros((dh, dw, 3), np.uint8)
        assert O.warpPerspective(frames[f], want, H, INTER_LINEAR, BORDER_REPLICATE, (3, 4, 5, 6)) == 0
        assert_same(got[f], want, f"batched warpPerspective frame {f}")
    mx = rng.uniform(-5, sw + 5, (dh, dw)).astype(np.float32); my = rng.uniform(-5, sh + 5, (dh, dw)).astype(np.float32)
    mxy = np.ascontiguousarray(np.stack([mx, my], -1))
    m1, m2 = convert_maps_fixed(mx, my)
    dmx, dmy, dmxy = torch.from_numpy(mx).cuda(), torch.from_numpy(my).cuda(), torch.from_numpy(mxy).cuda()
    dm1, dm2 = torch.from_numpy(m1).cuda(), torch.from_numpy(m2.view(np.int16)).cuda()
    for kind, (mt, p1, s1, p2, s2) in {"32f": (5, dmx.data_ptr(), dw * 4, dmy.data_ptr(), dw * 4), "32fc2": (13, dmxy.data_ptr(), dw * 8, None, 0),
                                       "16s": (11, dm1.data_ptr(), dw * 4, dm2.data_ptr(), dw * 2)}.items():
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC):
            out.zero_()
            assert capi.lib.b200_remap_batch_dev(16, dsrc.data_ptr(), sh * sw * 3, sw * 3, sw, sh, out.data_ptr(), dh * dw * 3, dw * 3, dw, dh, n,
                                                 mt, p1, s1, p2, s2, interp, BORDER_REFLECT_101, bv, None) == 0
            torch.cuda.synchronize()
            got = out.cpu().numpy()
            for f in range(n):
                want = np.zeros((dh, dw, 3), np.uint8)
                assert O.remap32f(frames[f], want, mx, my, interp, BORDER_REFLECT_101, (3, 4, 5, 6)) == 0 if kind != "16s" else \
                    O.remap16s(frames[f], want, m1, m2, interp, BORDER_REFLECT_101, (3, 4, 5, 6)) == 0
                assert_same(got[f], want, f"batched remap {kind} i{interp} frame {f}")
    assert capi.lib.b200_remap_batch_dev(16, dsrc.data_ptr(), sh * sw * 3, sw * 3, sw, sh, out.data_ptr(), dh * dw * 3, dw * 3, dw, dh, n,
                                         2, dm2.data_ptr(), dw * 2, None, 0, INTER_LINEAR, BORDER_REFLECT_101, bv, None) == 1


@pytest.mark.parametrize("cn,border", [(3, BORDER_CONSTANT), (3, BORDER_REFLECT_101), (4, BORDER_REPLICATE), (1, BORDER_WRAP)])
def test_single_launch_pipeline_equals_per_frame_oracle(G, cn, border):
    """enough frames for the persistent one-launch pipeline (ring of half-size frames in L2, work queue, per-frame
    dependency counters): every frame must equal cv::resize + cv::warpAffine of that frame, and it must be ONE launch."""
    from opencv_b200 import capi
    n, sw, sh, mw, mh, dw, dh = 45, 256, 144, 128, 72, 128, 72
    src = torch.zeros((n, sh, sw * cn), dtype=torch.uint8, device="cuda")
    assert capi.lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * cn, sw * cn, sw, sh, cn, 7, n, 99, None) == 0
    mid = torch.zeros((n, mh, mw * cn), dtype=torch.uint8, device="cuda")
    dst = torch.zeros((n, dh, dw * cn), dtype=torch.uint8, device="cuda")
    M = rotation_inverse(mw, mh, -12, 1.15, cx=mw / 2, cy=mh / 2)
    Mc = (C.c_double * 6)(*M.ravel())
    bvv = (11, 22, 33, 44)
    bv = (C.c_double * 4)(*bvv)
    typ = (cn - 1) << 3
    for rep in range(2):                                      # a second call re-uses the scratch (counters are reset per call)
        l0 = capi.lib.b200_launch_count()
        rc = capi.lib.b200_resize_warpAffine_batch_dev(typ, src.data_ptr(), sh * sw * cn, sw * cn, sw, sh, mid.data_ptr(), mh * mw * cn, mw * cn, mw, mh,
                                                       dst.data_ptr(), dh * dw * cn, dw * cn, dw, dh, n, INTER_AREA, Mc, INTER_LINEAR, border, bv, None)
        assert rc == 0, capi.last_error()
        torch.cuda.synchronize()
        assert capi.lib.b200_launch_count() - l0 == 1, "expected the one-launch pipeline"
        src_h, dst_h = src.cpu().numpy(), dst.cpu().numpy()
        for f in range(n):
            frame = src_h[f].reshape((sh, sw, cn) if cn > 1 else (sh, sw))
            m = np.zeros((mh, mw, cn) if cn > 1 else (mh, mw), np.uint8)
            o = np.zeros((dh, dw, cn) if cn > 1 else (dh, dw), np.uint8)
            assert O.resize(frame, m, INTER_AREA) == 0 and O.warpAffine(m, o, M, INTER_LINEAR, border, bvv) == 0
            assert_same(dst_h[f].reshape(o.shape), o, f"one-launch pipeline cn{cn} frame {f} call {rep}")
        dst.zero_()


def test_headline_geometry_one_launch_pipeline_against_oracle(G):
    """the geometry bench.py times (BASELINE configs[4]): 64 x 1920x1080 8UC3 -> 960x540 box -> warpAffine rot 15 / scale 0.9 in
    ONE launch of the persistent pipeline; frames 0, 21, 42 and 63 against the oracle's resize + warpAffine"""
    from opencv_b200 import capi
    n, sw, sh, mw, mh = 64, 1920, 1080, 960, 540
    src = torch.empty((n, sh, sw * 3), dtype=torch.uint8, device="cuda")
    assert capi.lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, 3, 0, n, 2026, None) == 0
    mid = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    dst = torch.zeros((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    M = rotation_inverse(mw, mh, 15.0, 0.9, cx=mw / 2.0, cy=mh / 2.0)
    Mc = (C.c_double * 6)(*M.ravel())
    bv = (C.c_double * 4)(0, 0, 0, 0)
    l0 = capi.lib.b200_launch_count()
    rc = capi.lib.b200_resize_warpAffine_batch_dev(16, src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, mid.data_ptr(), mh * mw * 3, mw * 3, mw, mh,
                                                   dst.data_ptr(), mh * mw * 3, mw * 3, mw, mh, n, INTER_LINEAR, Mc, INTER_LINEAR, BORDER_CONSTANT, bv, None)
    assert rc == 0, capi.last_error()
    torch.cuda.synchronize()
    assert capi.lib.b200_launch_count() - l0 == 1, "the headline batch must run as the one-launch pipeline"
    for f in (0, 21, 42, 63):
        frame = src[f].cpu().numpy().reshape(sh, sw, 3)
        m, o = np.zeros((mh, mw, 3), np.uint8), np.zeros((mh, mw, 3), np.uint8)
        assert O.resize(frame, m, INTER_LINEAR) == 0 and O.warpAffine(m, o, M, INTER_LINEAR, BORDER_CONSTANT) == 0
        assert_same(dst[f].cpu().numpy().reshape(o.shape), o, f"headline geometry, frame {f}")


def test_one_launch_pipeline_soak_against_the_two_kernel_path(G):
    """The persistent pipeline hands frames between CTAs through L2 with hand-rolled release / acquire counters and
    generic -> async proxy fences (fast_kernels.cu).  compute-sanitizer is not available on this GPU pool, so the protocol is
    exercised the statistical way: 40 runs of 64 x 1080p frames (2560 frames, ~390k hand-overs of strips and tiles, different
    data every run) must equal, byte for byte, what the two plain batched kernels (b200_resize_batch_dev then
    b200_warpAffine_batch_dev, no cross-CTA protocol at all) produce from the same frames."""
    from opencv_b200 import capi
    lib = capi.lib
    n, sw, sh, mw, mh = 64, 1920, 1080, 960, 540
    src = torch.empty((n, sh, sw * 3), dtype=torch.uint8, device="cuda")
    mid = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    mid2 = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    a = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    b = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    bv = (C.c_double * 4)(3, 5, 7, 0)
    for run in range(40):
        assert lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, 3, run * n, n, 1000 + run, None) == 0
        M = rotation_inverse(mw, mh, 3.0 + 4.0 * run, 0.8 + 0.01 * run, cx=mw / 2.0, cy=mh / 2.0)
        Mc = (C.c_double * 6)(*M.ravel())
        border = (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP)[run % 4]
        l0 = lib.b200_launch_count()
        assert lib.b200_resize_warpAffine_batch_dev(16, src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, mid.data_ptr(), mh * mw * 3, mw * 3, mw, mh,
                                                    a.data_ptr(), mh * mw * 3, mw * 3, mw, mh, n, INTER_LINEAR, Mc, INTER_LINEAR, border, bv, None) == 0
        assert lib.b200_launch_count() - l0 == 1
        assert lib.b200_resize_batch_dev(16, src.data_ptr(), sh * sw * 3, sw * 3, sw, sh, mid2.data_ptr(), mh * mw * 3, mw * 3, mw, mh, n, 0.0, 0.0, INTER_LINEAR, None) == 0
        assert lib.b200_warpAffine_batch_dev(16, mid2.data_ptr(), mh * mw * 3, mw * 3, mw, mh, b.data_ptr(), mh * mw * 3, mw * 3, mw, mh, n, Mc,
                                             INTER_LINEAR, border, bv, None) == 0
        assert torch.equal(a, b), f"run {run}: the one-launch pipeline differs from the two batched kernels"


def test_blob_from_images_fused_kernel_equals_oracle(G):
    from opencv_b200 import capi
    O.lib.orc_blobFromImages.argtypes = BLOB_SIG
    rng = np.random.default_rng(2)
    for cn in (1, 3, 4):
        for (sh, sw), (bh, bw) in BLOB_CASES:
            for swap in (0, 1):
                n = 4
                src = rng.integers(0, 256, (n, sh, sw, cn), dtype=np.uint8)
                sc, mn = blob_params(cn, swap)
                want = np.zeros((n, cn, bh, bw), np.float32)
                assert O.lib.orc_blobFromImages((cn - 1) << 3, src.ctypes.data, sh * sw * cn, sw * cn, sw, sh, want.ctypes.data, bw, bh, n, sc, mn, swap) == 0
                dsrc = torch.from_numpy(src).cuda()
                dblob = torch.full((n, cn, bh, bw), -1.0, dtype=torch.float32, device="cuda")
                rc = capi.lib.b200_blobFromImages_dev((cn - 1) << 3, dsrc.data_ptr(), sh * sw * cn, sw * cn, sw, sh, dblob.data_ptr(), bw, bh, n, sc, mn, swap, None)
                if (sh, sw) == (121, 161):
                    assert rc in (0, 1)          # an odd source at "2x": area-fast with ragged edges is declined
                    if rc == 1:
                        continue
                assert rc == 0, capi.last_error()
                torch.cuda.synchronize()
                assert_same(dblob.cpu().numpy(), want, f"blobFromImages cn{cn} {sh}x{sw}->{bh}x{bw} swap{swap}")
    assert capi.lib.b200_blobFromImages_dev(2, dsrc.data_ptr(), 1, 1, 4, 4, dblob.data_ptr(), 2, 2, 1, sc, mn, 0, None) == 1   # 16U: declined


def test_blob_from_images_crop_and_letterbox_equal_oracle(G):
    """DNN_PMODE_CROP_CENTER / DNN_PMODE_LETTERBOX in the fused kernel (dnn_utils.cpp:303-337), and the Python mirror's crop=True"""
    import opencv_b200 as cvb
    from opencv_b200 import capi
    O.lib.orc_blobFromImagesWithParams.argtypes = BLOB_PAD_SIG
    rng = np.random.default_rng(6)
    bv = (C.c_double * 4)(114.0, 3.7, 250.2, 300.0)
    for cn in (1, 3, 4):
        for (sh, sw), (bh, bw) in BLOB_PAD_CASES:
            for mode in (1, 2):
                swap = int(cn == 3)
                n = 3
                src = rng.integers(0, 256, (n, sh, sw, cn), dtype=np.uint8)
                sc, mn = blob_params(cn, swap)
                want = np.zeros((n, cn, bh, bw), np.float32)
                rco = O.lib.orc_blobFromImagesWithParams((cn - 1) << 3, src.ctypes.data, sh * sw * cn, sw * cn, sw, sh, want.ctypes.data, bw, bh, n, sc, mn, swap, mode, bv)
                dsrc = torch.from_numpy(src).cuda()
                dblob = torch.full((n, cn, bh, bw), -1.0, dtype=torch.float32, device="cuda")
                rc = capi.lib.b200_blobFromImagesWithParams_dev((cn - 1) << 3, dsrc.data_ptr(), sh * sw * cn, sw * cn, sw, sh, dblob.data_ptr(), bw, bh, n,
                                                                sc, mn, swap, mode, bv, None)
                if rc == 1:
                    continue                     # a resize cell the fused kernel leaves to the caller (ragged integer area)
                assert rc == 0 and rco == 0, capi.last_error()
                torch.cuda.synchronize()
                assert_same(dblob.cpu().numpy(), want, f"blobFromImagesWithParams cn{cn} {sh}x{sw}->{bh}x{bw} mode{mode}")
    src = rng.integers(0, 256, (2, 480, 640, 3), dtype=np.uint8)
    want = np.zeros((2, 3, 224, 224), np.float32)
    sc, mn = (C.c_double * 4)(1 / 255., 1 / 255., 1 / 255., 1 / 255.), (C.c_double * 4)(10, 20, 30, 0)
    assert O.lib.orc_blobFromImagesWithParams(16, src.ctypes.data, 480 * 640 * 3, 640 * 3, 640, 480, want.ctypes.data, 224, 224, 2, sc, mn, 1, 1, None) == 0
    got = cvb.blobFromImages(torch.from_numpy(src).cuda(), 1 / 255., (224, 224), (10, 20, 30), swapRB=True, crop=True)
    assert_same(got.cpu().numpy(), want, "blobFromImages(crop=True)")


def test_single_launch_pipeline_fuzz_against_per_frame_device_calls(G):
    """random geometries (odd widths -> tail groups, partial strips and tiles, strong scales that change the tile plan,
    all borders): the batched entry must equal b200_resize_dev + b200_warpAffine_dev per frame, byte for byte"""
    from opencv_b200 import capi
    rng = np.random.default_rng(2027)
    fused = 0
    for it in range(14):
        cn = int(rng.choice([1, 3, 4]))
        mw, mh = int(rng.integers(17, 260)), int(rng.integers(9, 200))
        if cn == 3:
            mw = (mw + 15) // 16 * 16 + int(rng.choice([0, 16]))      # TMA needs 16-byte row pitches: cn*mw % 16 == 0
        elif cn == 1:
            mw = (mw + 15) // 16 * 16
        else:
            mw = (mw + 3) // 4 * 4
        sw, sh = 2 * mw, 2 * mh
        dw = mw if cn != 4 else (int(rng.integers(8, 80)) * 4)
        dw = dw if (dw * cn) % 16 == 0 else mw
        dh = int(rng.integers(8, 220))
        n = int(rng.integers(33, 60))
        border = int(rng.choice([BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101]))
        M = rotation_inverse(mw, mh, rng.uniform(-180, 180), rng.uniform(0.4, 2.5), cx=mw * rng.uniform(0.2, 0.8), cy=mh * rng.uniform(0.2, 0.8))
        Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(9, 99, 199, 29)
        typ = (cn - 1) << 3
        src = torch.zeros((n, sh, sw * cn), dtype=torch.uint8, device="cuda")
        assert capi.lib.b200_synth_frames_dev(src.data_ptr(), sh * sw * cn, sw * cn, sw, sh, cn, it, n, 5, None) == 0
        mid = torch.zeros((n, mh, mw * cn), dtype=torch.uint8, device="cuda")
        dst = torch.zeros((n, dh, dw * cn), dtype=torch.uint8, device="cuda")
        l0 = capi.lib.b200_launch_count()
        rc = capi.lib.b200_resize_warpAffine_batch_dev(typ, src.data_ptr(), sh * sw * cn, sw * cn, sw, sh, mid.data_ptr(), mh * mw * cn, mw * cn, mw, mh,
                                                       dst.data_ptr(), dh * dw * cn, dw * cn, dw, dh, n, INTER_LINEAR, Mc, INTER_LINEAR, border, bv, None)
        assert rc == 0, capi.last_error()
        fused += capi.lib.b200_launch_count() - l0 == 1
        m1 = torch.zeros((mh, mw * cn), dtype=torch.uint8, device="cuda")
        ref = torch.zeros_like(dst)
        for f in range(n):
            assert capi.lib.b200_resize_dev(typ, src[f].data_ptr(), sw * cn, sw, sh, m1.data_ptr(), mw * cn, mw, mh, 0.0, 0.0, INTER_LINEAR, None) == 0
            assert capi.lib.b200_warpAffine_dev(typ, m1.data_ptr(), mw * cn, mw, mh, ref[f].data_ptr(), dw * cn, dw, dh, Mc, INTER_LINEAR, border, bv, None) == 0
        torch.cuda.synchronize()
        bad = int((dst != ref).sum().item())
        assert bad == 0, f"case {it}: cn{cn} {sw}x{sh}->{mw}x{mh}->{dw}x{dh} n{n} border{border}: {bad} bytes differ"
    assert fused >= 8, f"only {fused} of 14 cases took the one-launch path"


def test_host_frame_scheduler_over_all_visible_gpus(G):
    """b200_sched_run_pipeline_host: host frames sharded over every visible GPU == the single-GPU host stream call"""
    from opencv_b200 import capi
    rng = np.random.default_rng(12)
    n, sw, sh, mw, mh = 23, 256, 144, 128, 72
    frames = rng.integers(0, 256, (n, sh, sw, 3), dtype=np.uint8)
    M = rotation_inverse(mw, mh, 10, 0.95, cx=mw / 2, cy=mh / 2)
    Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)
    ndev = capi.lib.b200_device_count()
    devs = (C.c_int * ndev)(*range(ndev))
    ms = (C.c_float * ndev)()
    out = np.zeros((n, mh, mw, 3), np.uint8)
    rc = capi.lib.b200_sched_run_pipeline_host(devs, ndev, 16, frames.ctypes.data, sh * sw * 3, sw * 3, sw, sh, mw, mh,
                                               out.ctypes.data, mh * mw * 3, mw * 3, mw, mh, n, INTER_LINEAR, Mc, INTER_LINEAR,
                                               BORDER_CONSTANT, bv, ms)
    assert rc == 0, capi.last_error()
    import opencv_b200 as cvb
    want = cvb.resize_warpAffine_frames(frames, (mw, mh), M, flags=INTER_LINEAR | WARP_INVERSE_MAP)
    assert_same(out, want, f"host frame scheduler over {ndev} GPU(s)")
    assert all(ms[g] > 0 for g in range(ndev))


def test_scheduler_single_gpu_checksum_is_partition_invariant(G):
    from opencv_b200 import capi
    M = rotation_inverse(96, 54, 15, 0.9, cx=48, cy=27)
    Mc = (C.c_double * 6)(*M.ravel())

    def run(devs, chunk):
        d = (C.c_int * len(devs))(*devs)
        ms = (C.c_float * len(devs))()
        cs = (C.c_ulonglong * len(devs))()
        rc = capi.lib.b200_sched_run_pipeline(d, len(devs), 23, chunk, 192, 108, 3, 96, 54, 96, 54, Mc, 2026, ms, cs)
        assert rc == 0, capi.last_error()
        return sum(cs) & (2 ** 64 - 1), list(ms)

    c1, _ = run([0], 23)
    c2, _ = run([0], 4)                      # different chunking
    c3, _ = run([0, 0, 0], 5)                # three "devices" (same GPU): contiguous blocks 0-6, 7-14, 15-22
    assert c1 == c2 == c3 and c1 != 0


def test_scheduler_reruns_the_block_of_a_failed_device(G):
    """a device that cannot be used (here: an ordinal that does not exist) does not fail the job: its contiguous frame block is
    run again on a device whose own block went through (frames are independent), and b200_last_error() says so"""
    from opencv_b200 import capi
    M = rotation_inverse(96, 54, 15, 0.9, cx=48, cy=27)
    Mc = (C.c_double * 6)(*M.ravel()); bv = (C.c_double * 4)(0, 0, 0, 0)

    def run(devs):
        d = (C.c_int * len(devs))(*devs)
        ms = (C.c_float * len(devs))()
        cs = (C.c_ulonglong * len(devs))()
        rc = capi.lib.b200_sched_run_pipeline(d, len(devs), 23, 5, 192, 108, 3, 96, 54, 96, 54, Mc, 2026, ms, cs)
        return rc, list(cs)

    rc0, want = run([0, 0, 0])
    assert rc0 == 0
    rc1, got = run([0, 99, 0])
    assert rc1 == 0, capi.last_error()
    assert got == want, "the re-run block must give the checksum its own device would have given"
    assert "re-run on device 0" in capi.last_error()
    rc2, _ = run([98, 99])
    assert rc2 != 0                                                              # nobody left to take the frames
    # the host-frame scheduler
    rng = np.random.default_rng(13)
    n, sw, sh, mw, mh = 11, 256, 144, 128, 72
    frames = rng.integers(0, 256, (n, sh, sw, 3), dtype=np.uint8)
    out = np.zeros((n, mh, mw, 3), np.uint8)
    devs = (C.c_int * 2)(99, 0)
    ms = (C.c_float * 2)()
    rc = capi.lib.b200_sched_run_pipeline_host(devs, 2, 16, frames.ctypes.data, sh * sw * 3, sw * 3, sw, sh, mw, mh,
                                               out.ctypes.data, mh * mw * 3, mw * 3, mw, mh, n, INTER_LINEAR, Mc, INTER_LINEAR, BORDER_CONSTANT, bv, ms)
    assert rc == 0, capi.last_error()
    import opencv_b200 as cvb
    want = cvb.resize_warpAffine_frames(frames, (mw, mh), M, flags=INTER_LINEAR | WARP_INVERSE_MAP)
    assert_same(out, want, "host frame scheduler with a failed device")
    torch.cuda.set_device(0)


# ------------------------------------------------------------------ hot-cell kernels vs generic ----
@pytest.mark.parametrize("cn", [1, 3, 4])
def test_tma_warpaffine_fast_path_equals_generic_and_oracle(G, cn):
    """the TMA-staged warpAffine kernel (8U, INTER_LINEAR) must produce the generic kernel's bytes and the
    oracle's: every border, non-zero border value, partial tiles, large rotations / scales / shifts."""
    from opencv_b200 import capi
    rng = np.random.default_rng(300 + cn)
    cases = [((90, 130), (97, 150), 15.0, 0.9), ((200, 333), (131, 257), -73.0, 1.3), ((64, 64), (64, 64), 45.0, 0.55),
             ((50, 70), (300, 400), 170.0, 2.5), ((300, 400), (70, 90), 5.0, 0.3), ((17, 23), (129, 65), 90.0, 1.0),
             ((128, 256), (128, 256), 0.0, 1.0)]
    for (sh, sw), (dh, dw), ang, sc in cases:
        src = rand_image(rng, sh, sw, cn, np.uint8)
        dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        M = rotation_inverse(sw, sh, ang, sc)
        M[:, 2] += rng.uniform(-20, 20, 2)
        for border, bv in ((BORDER_CONSTANT, (0, 0, 0, 0)), (BORDER_CONSTANT, (7, 99, 200, 31)), (BORDER_REPLICATE, (0,) * 4),
                           (BORDER_REFLECT, (0,) * 4), (BORDER_WRAP, (0,) * 4), (BORDER_REFLECT_101, (0,) * 4)):
            fast, gen, ref = np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8)
            assert G.warpAffine(src, fast, M, INTER_LINEAR, border, bv) == 0
            capi.lib.b200_set_force_generic(1)
            try:
                assert G.warpAffine(src, gen, M, INTER_LINEAR, border, bv) == 0
            finally:
                capi.lib.b200_set_force_generic(0)
            assert O.warpAffine(src, ref, M, INTER_LINEAR, border, bv) == 0
            what = f"cn{cn} {sh}x{sw}->{dh}x{dw} rot{ang} s{sc} border{border} bv{bv}"
            assert_same(fast, ref, "fast vs oracle " + what)
            assert_same(gen, ref, "generic vs oracle " + what)


@pytest.mark.parametrize("cn", [1, 3, 4])
def test_box2x2_fast_path_equals_generic_and_oracle(G, cn):
    from opencv_b200 import capi
    rng = np.random.default_rng(400 + cn)
    for (sh, sw) in ((2, 2), (64, 96), (270, 480), (102, 34), (1080, 1920), (66, 4098)):
        src = rand_image(rng, sh, sw, cn, np.uint8)
        dshape = (sh // 2, sw // 2) if cn == 1 else (sh // 2, sw // 2, cn)
        for interp in (INTER_LINEAR, INTER_AREA):
            fast, gen, ref = np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8)
            assert G.resize(src, fast, interp) == 0
            capi.lib.b200_set_force_generic(1)
            try:
                assert G.resize(src, gen, interp) == 0
            finally:
                capi.lib.b200_set_force_generic(0)
            assert O.resize(src, ref, interp) == 0
            assert_same(fast, ref, f"box fast vs oracle cn{cn} {sh}x{sw} i{interp}")
            assert_same(gen, ref, f"box generic vs oracle cn{cn} {sh}x{sw} i{interp}")


def test_relative_maps(G):
    # WARP_RELATIVE_MAP reaches the remap hooks in `interpolation` (imgwarp.cpp:1608-1626)
    rng = np.random.default_rng(78)
    for dtype, cn in ((np.uint8, 3), (np.uint16, 1), (np.float32, 4)):
        src = rand_image(rng, 90, 120, cn, dtype)
        dshape = (90, 120) if cn == 1 else (90, 120, cn)
        mx = rng.uniform(-9, 9, (90, 120)).astype(np.float32)
        my = rng.uniform(-9, 9, (90, 120)).astype(np.float32)
        mxy = np.ascontiguousarray(np.stack([mx, my], -1))
        m1, m2 = convert_maps_fixed(mx, my)
        for interp in (INTER_NEAREST, INTER_LINEAR):
            for name, f in (("32f", lambda L, d: L.remap32f(src, d, mx, my, interp | WARP_RELATIVE_MAP, BORDER_REFLECT_101)),
                            ("32fc2", lambda L, d: L.remap32fc2(src, d, mxy, interp | WARP_RELATIVE_MAP, BORDER_CONSTANT, (4, 5, 6, 7))),
                            ("16s", lambda L, d: L.remap16s(src, d, m1, m2, interp | WARP_RELATIVE_MAP, BORDER_WRAP))):
                a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
                assert f(G, a) == 0 and f(O, b) == 0
                assert_same(a, b, f"relative remap {name} {dtype.__name__} i{interp}")


@pytest.mark.parametrize("cn", [1, 3, 4])
def test_integer_area_fast_path_equals_generic_and_oracle(G, cn):
    """kx x ky integer INTER_AREA for 8U (word-load kernel) vs the generic kernel vs the oracle; widths that are and are not
    multiples of 4 (tail group), kx != ky, ky = 1 and large ky."""
    from opencv_b200 import capi
    rng = np.random.default_rng(500 + cn)
    for (kx, ky), (dh, dw) in (((3, 3), (40, 64)), ((3, 3), (31, 53)), ((2, 4), (20, 36)), ((4, 4), (25, 41)), ((4, 2), (33, 8)),
                               ((3, 1), (17, 30)), ((2, 3), (9, 22)), ((4, 7), (11, 19)), ((3, 3), (720, 1280))):
        src = rand_image(rng, dh * ky, dw * kx, cn, np.uint8)
        src[rng.random(src.shape) < 0.3] = 255
        dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        fast, gen, ref = np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8)
        assert G.resize(src, fast, INTER_AREA) == 0
        capi.lib.b200_set_force_generic(1)
        try:
            assert G.resize(src, gen, INTER_AREA) == 0
        finally:
            capi.lib.b200_set_force_generic(0)
        assert O.resize(src, ref, INTER_AREA) == 0
        assert_same(fast, ref, f"area {kx}x{ky} cn{cn} ->{dh}x{dw} fast vs oracle")
        assert_same(gen, ref, f"area {kx}x{ky} cn{cn} ->{dh}x{dw} generic vs oracle")


def test_hooks_are_thread_safe(G):
    """OpenCV calls the hooks concurrently from arbitrary threads (user threads and parallel_for_ workers):
    8 Python threads (ctypes releases the GIL) hammer different ops at once, each must get its own exact result."""
    import threading
    rng = np.random.default_rng(91)
    jobs = []
    for i in range(8):
        cn = (1, 3, 4)[i % 3]
        src = rand_image(rng, 200 + 8 * i, 300 + 4 * i, cn, np.uint8)
        M = rotation_inverse(src.shape[1], src.shape[0], 10.0 + 7 * i, 0.8 + 0.05 * i)
        dshape = (220, 310) if cn == 1 else (220, 310, cn)
        ref_w, ref_r = np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8)
        assert O.warpAffine(src, ref_w, M, INTER_LINEAR, BORDER_REFLECT_101) == 0
        assert O.resize(src, ref_r, INTER_LINEAR) == 0
        jobs.append((src, M, dshape, ref_w, ref_r))
    errors = []

    def work(k):
        src, M, dshape, ref_w, ref_r = jobs[k]
        try:
            for _ in range(12):
                a, b = np.zeros(dshape, np.uint8), np.zeros(dshape, np.uint8)
                assert G.warpAffine(src, a, M, INTER_LINEAR, BORDER_REFLECT_101) == 0
                assert G.resize(src, b, INTER_LINEAR) == 0
                assert np.array_equal(a, ref_w) and np.array_equal(b, ref_r)
        except Exception as e:      # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=work, args=(k,)) for k in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_scheduler_across_two_gpus_matches_one_gpu(G):
    """contiguous frame blocks over 2 real devices (one host thread + stream each, no collective) give the single-GPU checksum"""
    from opencv_b200 import capi
    if capi.lib.b200_device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    M = rotation_inverse(96, 54, 15, 0.9, cx=48, cy=27)
    Mc = (C.c_double * 6)(*M.ravel())

    def run(devs):
        d = (C.c_int * len(devs))(*devs)
        ms = (C.c_float * len(devs))()
        cs = (C.c_ulonglong * len(devs))()
        rc = capi.lib.b200_sched_run_pipeline(d, len(devs), 50, 8, 192, 108, 3, 96, 54, 96, 54, Mc, 2026, ms, cs)
        assert rc == 0, capi.last_error()
        assert all(m > 0 for m in ms)
        return sum(cs) & (2 ** 64 - 1)

    assert run([0]) == run([0, 1]) == run([1, 0]) != 0


def test_device_tensor_on_a_gpu_that_is_not_current(G):
    """a tensor on cuda:1 while cuda:0 is current: the host mirror makes the tensor's device current for the call (the library
    keys interpolation tables and plans on cudaGetDevice())"""
    import opencv_b200 as cvb
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    O = oracle()
    rng = np.random.default_rng(3)
    src = rand_image(rng, 300, 400, 3, np.uint8)
    M = rotation_inverse(400, 300, 20.0, 0.9)
    torch.cuda.set_device(0)
    t = torch.from_numpy(src).to("cuda:1")
    for interp in (INTER_LINEAR, INTER_CUBIC):
        ref = np.zeros_like(src)
        assert O.warpAffine(src, ref, M, interp, BORDER_REFLECT_101) == 0
        out = cvb.warpAffine(t, M, (400, 300), flags=interp | cvb.WARP_INVERSE_MAP, borderMode=cvb.BORDER_REFLECT_101)
        assert out.device == t.device and torch.cuda.current_device() == 0
        assert_same(out.cpu().numpy(), ref, f"warpAffine on cuda:1, interp {interp}")
    ref = np.zeros((150, 200, 3), np.uint8)
    assert O.resize(src, ref, INTER_AREA) == 0
    assert_same(cvb.resize(t, (200, 150), interpolation=cvb.INTER_AREA).cpu().numpy(), ref, "resize on cuda:1")
