"""Pins the parity oracle (oracle/geom_oracle.c) before anything trusts it:

  1. the reference's own known-answer tests for this path
     (modules/imgproc/test/test_imgwarp.cpp:1241-1314, :1441-1479, :1494-1507, :1564-1588;
      test_imgwarp_strict.cpp:1101-1152 coordinate scheme),
  2. committed golden fixtures produced by the reference itself (tests/golden/make_golden.py),
  3. when oracle/_ref/libcvref.so (the reference's sources compiled here) is present: a randomised
     op x depth x channels x interpolation x border matrix, bit for bit.
CPU only.
"""
import ctypes as C

import numpy as np
import pytest

from common import *  # noqa: F403
from golden_io import load_golden, run_record

O = oracle()


def test_bilinear_table_entry0_quirk():
    # reference imgwarp.cpp:179-206: 32768 saturates to 32767 and the fix-up lands on cell 3
    O.lib.orc_bilinear_tab_i.restype = C.POINTER(C.c_short)
    t = np.ctypeslib.as_array(O.lib.orc_bilinear_tab_i(), shape=(1024, 4))
    assert t[0].tolist() == [32767, 0, 0, 1]
    assert (t[1:].astype(np.int64).sum(1) == 32768).all()
    fy, fx = np.divmod(np.arange(1, 1024), 32)
    closed = np.stack([(32 - fy) * (32 - fx), (32 - fy) * fx, fy * (32 - fx), fy * fx], 1) * 32
    assert (t[1:] == closed).all()


def test_kat_resize_area_regression():
    # test_imgwarp.cpp:1241-1276 (16x16 16UC1 -> 5x5 at scale 0.3, INTER_AREA; tolerance 1)
    inp = np.array([
        90, 94, 80, 3, 231, 2, 186, 245, 188, 165, 10, 19, 201, 169, 8, 228,
        86, 5, 203, 120, 136, 185, 24, 94, 81, 150, 163, 137, 88, 105, 132, 132,
        236, 48, 250, 218, 19, 52, 54, 221, 159, 112, 45, 11, 152, 153, 112, 134,
        78, 133, 136, 83, 65, 76, 82, 250, 9, 235, 148, 26, 236, 179, 200, 50,
        99, 51, 103, 142, 201, 65, 176, 33, 49, 226, 177, 109, 46, 21, 67, 130,
        54, 125, 107, 154, 145, 51, 199, 189, 161, 142, 231, 240, 139, 162, 240, 22,
        231, 86, 79, 106, 92, 47, 146, 156, 36, 207, 71, 33, 2, 244, 221, 71,
        44, 127, 71, 177, 75, 126, 68, 119, 200, 129, 191, 251, 6, 236, 247, 6,
        133, 175, 56, 239, 147, 221, 243, 154, 242, 82, 106, 99, 77, 158, 60, 229,
        2, 42, 24, 174, 27, 198, 14, 204, 246, 251, 141, 31, 114, 163, 29, 147,
        121, 53, 74, 31, 147, 189, 42, 98, 202, 17, 228, 123, 209, 40, 77, 49,
        112, 203, 30, 12, 205, 25, 19, 106, 145, 185, 163, 201, 237, 223, 247, 38,
        33, 105, 243, 117, 92, 179, 204, 248, 160, 90, 73, 126, 2, 41, 213, 204,
        6, 124, 195, 201, 230, 187, 210, 167, 48, 79, 123, 159, 145, 218, 105, 209,
        240, 152, 136, 235, 235, 164, 157, 9, 152, 38, 27, 209, 120, 77, 238, 196,
        240, 233, 10, 241, 90, 67, 12, 79, 0, 43, 58, 27, 83, 199, 190, 182], np.uint16).reshape(16, 16)
    exp = np.array([120, 100, 151, 101, 130, 106, 115, 141, 130, 127, 91, 136, 170, 114, 140,
                    104, 122, 131, 147, 133, 161, 163, 70, 107, 182], np.uint16).reshape(5, 5)
    dst = np.zeros((5, 5), np.uint16)
    assert O.resize(inp, dst, INTER_AREA, (0.3, 0.3)) == 0
    assert np.abs(dst.astype(int) - exp).max() <= 1


def test_kat_area_half_and_quarter_round():
    # test_imgwarp.cpp:1278-1314
    i = np.arange(32 * 32)
    inp = (i % 2 + 253 + i // (16 * 32)).astype(np.uint8).reshape(32, 32)
    d = np.zeros((16, 16), np.uint8)
    assert O.resize(inp, d, INTER_AREA, (0.5, 0.5)) == 0
    assert (d.ravel() == (254 + np.arange(256) // 128).astype(np.uint8)).all()
    d = np.zeros((8, 8), np.uint8)
    assert O.resize(inp, d, INTER_AREA, (0.25, 0.25)) == 0
    assert (d == 254).all()


@pytest.mark.parametrize("dtype,cn", [(np.uint8, 1), (np.uint8, 4), (np.uint16, 1), (np.uint16, 4),
                                      (np.int16, 1), (np.int16, 3), (np.int16, 4),
                                      (np.float32, 1), (np.float32, 4)])
def test_kat_area_half_formula(dtype, cn):
    # test_imgwarp.cpp:1441-1479 Resize.Area_half: ints (a+b+c+d+2)>>2 exactly, 32F sum*0.25 within 7e-5
    rng = np.random.default_rng(17)
    src = rand_image(rng, 200, 200, cn, dtype)
    dst = np.zeros((100, 100) if cn == 1 else (100, 100, cn), dtype)
    assert O.resize(src, dst, INTER_AREA) == 0
    s = src.reshape(200, 200, cn)
    if dtype == np.float32:
        ref = (s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2]) * np.float32(0.25)
        assert np.abs(dst.reshape(ref.shape) - ref).max() <= 7e-5 * 1000
    else:
        w = s.astype(np.int64)
        ref = (w[0::2, 0::2] + w[0::2, 1::2] + w[1::2, 0::2] + w[1::2, 1::2] + 2) >> 2
        assert (dst.reshape(ref.shape) == ref).all()


def test_kat_nearest_regression_15075():
    # test_imgwarp.cpp:1494-1507: 5-channel 12x12 -> 11x11 NN picks exactly one source pixel
    Cn = 5
    src = np.zeros((12, 12, Cn), np.uint8)
    src[5, 5, :] = 1
    dst = np.zeros((11, 11, Cn), np.uint8)
    assert O.resize(src, dst, INTER_NEAREST) == 0
    assert int(dst.sum()) == Cn


def test_kat_warp_regression_28554():
    # test_imgwarp.cpp:1564-1588: 16S NN translate by 64 (forward matrix I|64 -> inverse I|-64)
    src = np.ones((128, 128), np.int16)
    dst = np.full((256, 256), -5, np.int16)
    M = np.array([[1, 0, -64.0], [0, 1, -64.0]])
    assert O.warpAffine(src, dst, M, INTER_NEAREST, BORDER_CONSTANT, (0, 0, 0, 0)) == 0
    ref = np.zeros((256, 256), np.int16)
    ref[64:192, 64:192] = 1
    assert (dst == ref).all()


def test_kat_warpaffine_equals_remap_of_fixed_point_scheme():
    # test_imgwarp_strict.cpp:1101-1152: warpAffine(8U) == remap16s over the AB_BITS=10 map, error level 0
    rng = np.random.default_rng(5)
    src = rand_image(rng, 60, 80, 3, np.uint8)
    M = rotation_inverse(80, 60, 33.0, 1.2)
    dh, dw = 70, 90
    x = np.arange(dw, dtype=np.float64)
    y = np.arange(dh, dtype=np.float64)[:, None]
    rnd = lambda v: np.rint(v).astype(np.int64)  # noqa: E731
    X = rnd(M[0, 0] * x * 1024) + rnd((M[0, 1] * y + M[0, 2]) * 1024) + 16
    Y = rnd(M[1, 0] * x * 1024) + rnd((M[1, 1] * y + M[1, 2]) * 1024) + 16
    X >>= 5
    Y >>= 5
    m1 = np.stack([np.clip(X >> 5, -32768, 32767), np.clip(Y >> 5, -32768, 32767)], -1).astype(np.int16)
    m2 = ((Y & 31) * 32 + (X & 31)).astype(np.uint16)
    for border in (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP):
        a = np.zeros((dh, dw, 3), np.uint8)
        b = np.zeros((dh, dw, 3), np.uint8)
        assert O.warpAffine(src, a, M, INTER_LINEAR, border, (9, 8, 7, 0)) == 0
        assert O.remap16s(src, b, np.ascontiguousarray(m1), m2, INTER_LINEAR, border, (9, 8, 7, 0)) == 0
        assert_same(a, b, f"border {border}")


def test_golden_fixtures_from_reference():
    recs = load_golden()
    assert len(recs) > 200
    for i, rec in enumerate(recs):
        got = run_record(O, rec)
        assert_same(got, rec["dst"], f"golden #{i} {rec['op']} {rec['params'][:2]}")


def test_kat_resize_bitexact_nearest8u():
    # test_resize_bitexact.cpp:191-243 Resize_Bitexact.Nearest8U: literal matrices, and their transposes
    cases = [([[0, 1, 2, 3, 4, 5]], [[1, 3, 5]]),
             ([[0, 1, 2, 3, 4]], [[2]]),
             ([[0, 1, 2, 3, 4]], [[0, 2, 4]]),
             ([[0, 1, 2, 3, 4]], [[1, 3]]),
             ([[0, 1, 2, 3, 4], [5, 6, 7, 8, 9], [10, 11, 12, 13, 14]],
              [[0, 1, 1, 2, 3, 3, 4], [0, 1, 1, 2, 3, 3, 4], [5, 6, 6, 7, 8, 8, 9], [10, 11, 11, 12, 13, 13, 14], [10, 11, 11, 12, 13, 13, 14]]),
             ([[0, 1, 2], [3, 4, 5]], [[0, 0, 1, 1, 2, 2], [0, 0, 1, 1, 2, 2], [3, 3, 4, 4, 5, 5], [3, 3, 4, 4, 5, 5]])]
    for s_, d_ in cases:
        for tr in (False, True):
            src, exp = np.array(s_, np.uint8), np.array(d_, np.uint8)
            if tr:
                src, exp = np.ascontiguousarray(src.T), np.ascontiguousarray(exp.T)
            out = np.zeros_like(exp)
            assert O.resize(src, out, INTER_NEAREST_EXACT) == 0
            assert (out == exp).all(), (s_, tr)


def test_kat_resize_bitexact_linear8u_fixed_point_reference():
    # test_resize_bitexact.cpp:21-153 Resize_Bitexact.Linear8U: in-test 8.8 fixed-point reference (fixedShiftU8 = 8):
    # x coefficients from softdouble, H pass in 8.8, V pass in 16.16, round half up, edges replicate
    rng = np.random.default_rng(0x1234)
    for (sh, sw), (dh, dw), cn in (((48, 64), (31, 53), 1), ((48, 64), (96, 70), 3), ((37, 41), (12, 90), 4), ((20, 30), (23, 7), 2)):
        src = rand_image(rng, sh, sw, cn, np.uint8)
        out = np.zeros((dh, dw) if cn == 1 else (dh, dw, cn), np.uint8)
        assert O.resize(src, out, INTER_LINEAR_EXACT) == 0
        S = src.reshape(sh, sw, cn).astype(np.int64)

        def coefs(n_dst, n_src):
            idx, c = np.zeros(n_dst, np.int64), np.zeros((n_dst, 2), np.int64)
            for d in range(n_dst):
                f = (n_src / n_dst) * (d + 0.5) - 0.5 if True else 0
                f = (1.0 / (n_dst / n_src)) * (d + 0.5) - 0.5
                i = int(np.floor(f))
                if i < 0 or n_src <= 1:
                    idx[d], c[d] = 0, (256, 0)
                elif i >= n_src - 1:
                    idx[d], c[d] = n_src - 1, (256, 0)
                else:
                    c1 = int(np.rint((f - i) * 256))
                    idx[d], c[d] = i, (256 - c1, c1)
            return idx, c
        ix, cx = coefs(dw, sw)
        iy, cy = coefs(dh, sh)
        x1 = np.minimum(ix + 1, sw - 1)
        y1 = np.minimum(iy + 1, sh - 1)
        Hh = S[:, ix, :] * cx[None, :, 0, None] + S[:, x1, :] * cx[None, :, 1, None]          # 8.8
        V = Hh[iy] * cy[:, 0, None, None] + Hh[y1] * cy[:, 1, None, None]                    # 16.16
        ref = ((V + 32768) >> 16).astype(np.uint8)
        assert (out.reshape(ref.shape) == ref).all(), ((sh, sw), (dh, dw), cn)


needs_ref = pytest.mark.skipif(cvref() is None, reason="oracle/_ref not built (needs /root/reference)")


@needs_ref
@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.int16])
@pytest.mark.parametrize("cn", [1, 3, 4])
def test_oracle_vs_reference_warp_remap(dtype, cn):
    R = cvref()
    rng = np.random.default_rng(100 + cn)
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
                assert f(O, a) == 0 and f(R, b) == 0
                assert_same(a, b, f"{name} i{interp} b{border}")


@needs_ref
def test_convert_maps_restatement_vs_reference():
    R = cvref()
    rng = np.random.default_rng(3)
    mx = rng.uniform(-50, 700, (40, 50)).astype(np.float32)
    my = rng.uniform(-50, 700, (40, 50)).astype(np.float32)
    o1 = np.zeros((40, 50, 2), np.int16)
    o2 = np.zeros((40, 50), np.uint16)
    vp = C.c_void_p
    R.lib.cvref_convertMaps.argtypes = [vp, C.c_size_t, vp, C.c_size_t, C.c_int, C.c_int, vp, C.c_size_t, vp, C.c_size_t, C.c_int]
    assert R.lib.cvref_convertMaps(mx.ctypes.data, mx.strides[0], my.ctypes.data, my.strides[0], 50, 40,
                                   o1.ctypes.data, o1.strides[0], o2.ctypes.data, o2.strides[0], 0) == 0
    m1, m2 = convert_maps_fixed(mx, my)
    assert (m1 == o1).all() and (m2 == o2).all()


@needs_ref
@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.int16])
def test_border_transparent_vs_reference(dtype):
    """BORDER_TRANSPARENT incl. the renormalised last row / column of remapBilinear (imgwarp.cpp:726-755)"""
    R = cvref()
    rng = np.random.default_rng(9)
    sh, sw, dh, dw = 37, 53, 41, 67
    for cn in (1, 3, 4):
        src = rand_image(rng, sh, sw, cn, dtype)
        mx = rng.uniform(-3, sw + 3, (dh, dw)).astype(np.float32); my = rng.uniform(-3, sh + 3, (dh, dw)).astype(np.float32)
        mx[::3, ::2] = np.round(mx[::3, ::2]); my[::2, ::3] = np.round(my[::2, ::3]); mx[5, :] = sw - 1; my[:, 7] = sh - 1
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4):
            for M in (rotation_inverse(sw, sh, 21, 0.8), np.array([[1, 0, 0.5], [0, 1, 0.25]]), np.array([[1, 0, 2.0], [0, 1, 1.0]])):
                init = rand_image(rng, dh, dw, cn, dtype)
                a, b = init.copy(), init.copy()
                assert O.warpAffine(src, a, M, interp, 5) == 0 and R.warpAffine(src, b, M, interp, 5) == 0
                assert_same(a, b, f"transparent warpAffine cn{cn} i{interp}")
            init = rand_image(rng, dh, dw, cn, dtype)
            a, b = init.copy(), init.copy()
            assert O.remap32f(src, a, mx, my, interp, 5) == 0 and R.remap32f(src, b, mx, my, interp, 5) == 0
            assert_same(a, b, f"transparent remap cn{cn} i{interp}")


@needs_ref
def test_blob_from_images_vs_reference():
    """orc_blobFromImages against cv::dnn::blobFromImagesWithParams (dnn_utils.cpp compiled into oracle/_ref)"""
    R = C.CDLL(cvref_path())
    O.lib.orc_blobFromImages.argtypes = BLOB_SIG
    R.cvref_blobFromImages.argtypes = BLOB_SIG
    rng = np.random.default_rng(2)
    for cn in (1, 3, 4):
        for (sh, sw), (bh, bw) in BLOB_CASES:
            for swap in (0, 1):
                src = rng.integers(0, 256, (3, sh, sw, cn), dtype=np.uint8)
                sc, mn = blob_params(cn, swap)
                a = np.zeros((3, cn, bh, bw), np.float32); b = np.zeros_like(a)
                args = ((cn - 1) << 3, src.ctypes.data, sh * sw * cn, sw * cn, sw, sh)
                assert O.lib.orc_blobFromImages(*args, a.ctypes.data, bw, bh, 3, sc, mn, swap) == 0
                assert R.cvref_blobFromImages(*args, b.ctypes.data, bw, bh, 3, sc, mn, swap) == 0
                assert_same(a, b, f"blobFromImages cn{cn} {sh}x{sw}->{bh}x{bw} swap{swap}")


@needs_ref
def test_blob_from_images_padding_modes_vs_reference():
    """orc_blobFromImagesWithParams (DNN_PMODE_CROP_CENTER, DNN_PMODE_LETTERBOX) against cv::dnn::blobFromImagesWithParams"""
    R = C.CDLL(cvref_path())
    O.lib.orc_blobFromImagesWithParams.argtypes = BLOB_PAD_SIG
    R.cvref_blobFromImagesWithParams.argtypes = BLOB_PAD_SIG
    rng = np.random.default_rng(4)
    bv = (C.c_double * 4)(114.0, 3.7, 250.2, 300.0)
    for cn in (1, 3, 4):
        for (sh, sw), (bh, bw) in BLOB_PAD_CASES:
            for mode in (1, 2):
                for swap in (0, 1):
                    src = rng.integers(0, 256, (2, sh, sw, cn), dtype=np.uint8)
                    sc, mn = blob_params(cn, swap)
                    a = np.zeros((2, cn, bh, bw), np.float32); b = np.zeros_like(a)
                    args = ((cn - 1) << 3, src.ctypes.data, sh * sw * cn, sw * cn, sw, sh)
                    assert R.cvref_blobFromImagesWithParams(*args, b.ctypes.data, bw, bh, 2, sc, mn, swap, mode, bv) == 0
                    assert O.lib.orc_blobFromImagesWithParams(*args, a.ctypes.data, bw, bh, 2, sc, mn, swap, mode, bv) == 0
                    assert_same(a, b, f"blobFromImagesWithParams cn{cn} {sh}x{sw}->{bh}x{bw} mode{mode} swap{swap}")


@needs_ref
def test_convert_maps_every_pair_vs_reference():
    """orc_convertMaps against cv::convertMaps itself (oracle/ref_shim.cpp) for every accepted pair of map types."""
    R = C.CDLL(cvref_path())
    O.lib.orc_convertMaps.argtypes = CONVERT_MAPS_SIG
    R.cvref_convertMaps_any.argtypes = CONVERT_MAPS_SIG
    cases = convert_maps_cases(np.random.default_rng(3))
    assert len(cases) >= 25
    for case in cases:
        a1, a2 = convert_maps_outputs(case)
        b1, b2 = convert_maps_outputs(case)
        assert call_convert_maps(O.lib.orc_convertMaps, case, a1, a2) == 0
        rc = call_convert_maps(R.cvref_convertMaps_any, case, b1, b2)
        assert rc >= 0, case[0:1] + case[4:]
        assert_same(a1, b1, f"convertMaps map1 {case[0]}->{case[4]} nn{case[5]}")
        if rc & 0x10000:
            assert_same(a2, b2, f"convertMaps map2 {case[0]}->{case[4]} nn{case[5]}")
        else:
            assert (a2 == 7).all(), "second plane written where the reference creates none"
    # pairs the reference rejects
    s2 = np.zeros((4, 4, 2), np.int16); d1, d2 = np.zeros((4, 4, 2), np.float32), np.zeros((4, 4), np.float32)
    assert call_convert_maps(O.lib.orc_convertMaps, (CV_16SC2, s2, 0, None, CV_32FC2, 0), d1, d2) == 1
    assert call_convert_maps(R.cvref_convertMaps_any, (CV_16SC2, s2, 0, None, CV_32FC2, 0), d1, d2) == -1


RESIZE_SIZES = [((97, 131), (113, 150)), ((97, 131), (48, 65)), ((96, 132), (48, 66)), ((96, 132), (32, 44)),
                ((96, 132), (24, 66)), ((96, 132), (48, 33)), ((97, 131), (40, 50)), ((97, 131), (200, 300)),
                ((97, 131), (97, 60)), ((97, 131), (150, 70)), ((97, 131), (50, 200)), ((64, 64), (16, 16)),
                ((1, 50), (1, 20)), ((50, 1), (20, 1)), ((5, 5), (11, 13)), ((100, 100), (7, 9))]


@needs_ref
@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.int16])
@pytest.mark.parametrize("cn", [1, 2, 3, 4])
def test_oracle_vs_reference_resize(dtype, cn):
    R = cvref()
    rng = np.random.default_rng(200 + cn)
    for (sh, sw), (dh, dw) in RESIZE_SIZES:
        src = rand_image(rng, sh, sw, cn, dtype)
        dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        exact = (INTER_NEAREST_EXACT,) + ((INTER_LINEAR_EXACT,) if dtype in (np.uint8, np.uint16, np.int16) else ())
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA, INTER_CUBIC, INTER_LANCZOS4) + exact:
            a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
            assert O.resize(src, a, interp) == 0 and R.resize(src, b, interp) == 0
            assert_same(a, b, f"resize {sh}x{sw}->{dh}x{dw} i{interp}")
    # ragged right/bottom edges: explicit fx,fy on odd sizes (resize.cpp:3028-3050, :4015-4021)
    for (sh, sw), f in [((27, 35), 0.5), ((35, 27), 0.5), ((31, 37), 1 / 3.), ((29, 35), 0.25)]:
        src = rand_image(rng, sh, sw, cn, dtype)
        dw, dh = int(np.rint(sw * f)), int(np.rint(sh * f))
        dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA):
            a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
            assert O.resize(src, a, interp, (f, f)) == 0 and R.resize(src, b, interp, (f, f)) == 0
            assert_same(a, b, f"ragged resize {sh}x{sw} f={f:.3f} i{interp}")


@needs_ref
def test_oracle_vs_reference_relative_maps():
    # WARP_RELATIVE_MAP (imgwarp.cpp:293-299, :647-652; test_imgwarp.cpp:1319-1382): the map holds offsets from the dst pixel
    R = cvref()
    rng = np.random.default_rng(77)
    for dtype, cn in ((np.uint8, 3), (np.uint16, 1), (np.float32, 4)):
        src = rand_image(rng, 50, 70, cn, dtype)
        dshape = (50, 70) if cn == 1 else (50, 70, cn)
        mx = rng.uniform(-9, 9, (50, 70)).astype(np.float32)
        my = rng.uniform(-9, 9, (50, 70)).astype(np.float32)
        mxy = np.ascontiguousarray(np.stack([mx, my], -1))
        m1, m2 = convert_maps_fixed(mx, my)
        for interp in (INTER_NEAREST, INTER_LINEAR):
            for border in (BORDER_CONSTANT, BORDER_REFLECT_101):
                for name, f in (("32f", lambda L, d: L.remap32f(src, d, mx, my, interp | WARP_RELATIVE_MAP, border, (1, 2, 3, 4))),
                                ("32fc2", lambda L, d: L.remap32fc2(src, d, mxy, interp | WARP_RELATIVE_MAP, border, (1, 2, 3, 4))),
                                ("16s", lambda L, d: L.remap16s(src, d, m1, m2, interp | WARP_RELATIVE_MAP, border, (1, 2, 3, 4)))):
                    a, b = np.zeros(dshape, dtype), np.zeros(dshape, dtype)
                    assert f(O, a) == 0 and f(R, b) == 0
                    assert_same(a, b, f"relative remap {name} {dtype.__name__} i{interp} b{border}")


@needs_ref
@pytest.mark.parametrize("m1type", [11, 5])
def test_fisheye_map_restatement_vs_reference(m1type):
    # orc_fisheye_initUndistortRectifyMap against the reference's cv::fisheye::initUndistortRectifyMap (calib3d/src/fisheye.cpp:530-631
    # compiled into oracle/_ref), fed the reference's own (P*R)^-1: same libm, same operation order -> bit for bit
    for case in fisheye_cases():
        r1, r2, iR = ref_fisheye_map(case, m1type)
        o1, o2 = oracle_fisheye_map(case, m1type, iR)
        if m1type == 5:
            assert (r1.view(np.uint32) == o1.view(np.uint32)).all() and (r2.view(np.uint32) == o2.view(np.uint32)).all(), case[0]
        else:
            assert_same(o1, r1, case[0] + " map1"); assert_same(o2, r2, case[0] + " map2")
    # the edge cases are really present
    m1, _m2, _ = ref_fisheye_map(fisheye_cases()[2], 5)
    assert np.isinf(m1).any(), "the behind-the-camera case must produce infinities"


@needs_ref
def test_64f_restatement_vs_reference():
    # CV_64F INTER_LINEAR (imgwarp.cpp:1546 remapBilinear<Cast<double,double>, RemapNoVec, float>; resize.cpp linear_tab 64F row)
    R = cvref()
    rng = np.random.default_rng(6)
    for cn in (1, 3, 4):
        src = (rng.random((97, 131, cn)) * 2000 - 1000).squeeze()
        shape = (110, 140) if cn == 1 else (110, 140, cn)
        M = rotation_inverse(131, 97, 20, 0.9)
        Hm = rand_homography_inverse(rng, 131, 97)
        mx = rng.uniform(-10, 140, (110, 140)).astype(np.float32); my = rng.uniform(-10, 105, (110, 140)).astype(np.float32)
        for border in (BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101):
            for interp in (INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4):      # :1546, :1564, :1588: Cast<double,double>, float weights
                for call in (lambda L, d: L.warpAffine(src, d, M, interp, border, (1.5, 2.5, 3.5, 4.5)),
                             lambda L, d: L.warpPerspective(src, d, Hm, interp, border, (1.5, 2.5, 3.5, 4.5)),
                             lambda L, d: L.remap32f(src, d, mx, my, interp, border, (1.5, 2.5, 3.5, 4.5))):
                    a, b = np.zeros(shape), np.zeros(shape)
                    assert call(O, a) == 0 and call(R, b) == 0
                    assert (a == b).all(), f"64F interp {interp} border {border} cn {cn}"
        for (dh, dw) in ((150, 200), (50, 77), (97, 262), (48, 65), (32, 43), (24, 33)):
            for interp in (INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4, INTER_AREA):      # the 64F rows of resize's function tables
                a = np.zeros((dh, dw) if cn == 1 else (dh, dw, cn)); b = np.zeros_like(a)
                assert O.resize(src, a, interp) == 0 and R.resize(src, b, interp) == 0
                assert (a == b).all(), f"64F resize interp {interp} cn {cn} -> {dw}x{dh}"


@needs_ref
@pytest.mark.parametrize("m1type", [11, 5, 13])
def test_undistort_map_restatement_vs_reference(m1type):
    # orc_initUndistortRectifyMap against the reference's cv::initUndistortRectifyMap as this host dispatches it (AVX2 variant: fused
    # vector body + scalar tail, calib3d/src/undistort.simd.hpp:128-277), fed the reference's own iR and tilt matrix
    lib = C.CDLL(cvref_path())
    if not lib.cvref_have_avx2():
        pytest.skip("host without AVX2: the reference dispatches its baseline variant, which rounds differently")
    for case in undistort_cases():
        r1, r2, iR, tilt = ref_undistort_map(case, m1type)
        o1, o2 = oracle_undistort_map(case, m1type, iR, tilt)
        assert (r1.view(np.uint8) == o1.view(np.uint8)).all(), case[0] + " map1"
        if r2 is not None:
            assert (r2.view(np.uint8) == o2.view(np.uint8)).all(), case[0] + " map2"


@needs_ref
def test_undistort_matrices_vs_reference():
    # orc_undistort_params AND the product's host helper b200_undistort_params (no device work: runs here) against the reference's own
    # (newK*R)^-1 and tilt matrix (undistort.dispatch.cpp:123, :155-157) -- bit for bit, a rectifying R included
    from opencv_b200 import capi
    for case in undistort_cases():
        _n, K, dist, R, P, (w, h) = case
        _r1, _r2, iR, tilt = ref_undistort_map(case, 13)
        newK = P
        if newK is None:                                # :108 getDefaultNewCameraMatrix(A, size, true)
            newK = K.copy(); newK[0, 2] = (w - 1) * 0.5; newK[1, 2] = (h - 1) * 0.5
        tx, ty = (dist[12], dist[13]) if len(dist) >= 14 else (0.0, 0.0)
        oi, ot = oracle_undistort_params(newK, R, tx, ty)
        assert (oi == iR).all() and (ot == tilt).all(), case[0]
        gi, gt = (C.c_double * 9)(), (C.c_double * 9)()
        Rc = None if R is None else (C.c_double * 9)(*R.ravel())
        assert capi.lib.b200_undistort_params((C.c_double * 9)(*newK.ravel()), Rc, tx, ty, gi, gt) == 0
        assert (np.array(gi) == iR).all() and (np.array(gt) == tilt).all(), case[0] + " (b200_undistort_params)"


@needs_ref
def test_undistort_image_restatement_vs_reference():
    # orc_undistort (stripes of max(1, 4096 / cols) rows, each with its own inverted matrix) against cv::undistort
    lib = C.CDLL(cvref_path())
    if not lib.cvref_have_avx2():
        pytest.skip("host without AVX2")
    rng = np.random.default_rng(5)
    for case in undistort_image_cases():
        _n, K, dist, newK, (w, h) = case
        src = rand_image(rng, h, w, 3, np.uint8)
        want = np.zeros_like(src)
        Kc, dc = np.ascontiguousarray(K), np.ascontiguousarray(dist)
        Pc = None if newK is None else np.ascontiguousarray(newK)
        assert lib.cvref_undistort(16, src.ctypes.data_as(C.c_void_p), C.c_size_t(w * 3), w, h, want.ctypes.data_as(C.c_void_p), C.c_size_t(w * 3),
                                   Kc.ctypes.data_as(C.c_void_p), dc.ctypes.data_as(C.c_void_p), len(dc),
                                   Pc.ctypes.data_as(C.c_void_p) if Pc is not None else None) == 0
        got = oracle_undistort(case, src)[0]
        assert (got == want).all(), case[0]


@needs_ref
def test_pyrdown_restatement_vs_reference():
    # orc_pyrDown against cv::pyrDown (imgproc/src/pyramids.cpp): integer depths exact by construction; CV_32F needs the reference's
    # two evaluation orders (vector body / scalar tail) placed where the reference places them, and the right-hand border table quirk
    lib = C.CDLL(cvref_path())
    rng = np.random.default_rng(17)
    n = 0
    for (dt, cn, (w, h), border, (dw, dh)) in pyrdown_cases():
        src = pyr_image(rng, dt, h, w, cn)
        want = np.zeros((dh, dw, cn), dt)
        assert lib.cvref_pyrDown(cv_type(src), src.ctypes.data_as(C.c_void_p), C.c_size_t(src.strides[0]), w, h, want.ctypes.data_as(C.c_void_p),
                                 C.c_size_t(want.strides[0]), dw, dh, border, None) == 0
        got = oracle_pyrDown(src, dw, dh, border)
        assert (got.view(np.uint8) == want.view(np.uint8)).all(), f"{dt.__name__}C{cn} {w}x{h} -> {dw}x{dh} border {border}"
        n += 1
    assert n > 500


@needs_ref
def test_pyrup_restatement_vs_reference():
    # orc_pyrUp against cv::pyrUp (imgproc/src/pyramids.cpp:1374-1416): all five depths, byte for byte
    lib = C.CDLL(cvref_path())
    rng = np.random.default_rng(23)
    n = 0
    for (dt, cn, (w, h), (dw, dh)) in pyrup_cases():
        src = pyrup_image(rng, dt, h, w, cn)
        want = np.zeros((dh, dw, cn), dt)
        assert lib.cvref_pyrUp(cv_type(src), src.ctypes.data_as(C.c_void_p), C.c_size_t(src.strides[0]), w, h, want.ctypes.data_as(C.c_void_p),
                               C.c_size_t(want.strides[0]), dw, dh, BORDER_REFLECT_101) == 0
        got = oracle_pyrUp(src, dw, dh)
        assert (got.view(np.uint8) == want.view(np.uint8)).all(), f"{dt.__name__}C{cn} {w}x{h} -> {dw}x{dh}"
        n += 1
    assert n > 400


@needs_ref
@pytest.mark.parametrize("kind", [0, 1, 2])
def test_stitching_warper_maps_restatement_vs_reference(kind):
    # orc_stitch_buildMaps against cv::detail::PlaneWarper / SphericalWarper / CylindricalWarper::buildMaps (stitching/src/warpers.cpp
    # compiled into oracle/_ref), fed the projector's own k_rinv / t and the ROI the warper detected: bit for bit on this host
    for case in stitch_cases():
        xm, ym, tl, kr, tt = ref_stitch_maps(kind, case)
        ox, oy = oracle_stitch_maps(kind, kr, tt, case[3], tl, xm.shape[1], xm.shape[0])
        assert (ox.view(np.uint32) == xm.view(np.uint32)).all() and (oy.view(np.uint32) == ym.view(np.uint32)).all()


@needs_ref
def test_linear_exact_signed_restatement_vs_reference():
    # INTER_LINEAR_EXACT on CV_8S / CV_16S (both in fixedpoint32) and CV_32S (fixedpoint64); extremes included: products and sums
    # saturate in the reference's types
    lib = C.CDLL(cvref_path())
    rng = np.random.default_rng(9)
    n = 0
    for dt in (np.int8, np.int16, np.int32):
        for cn in (1, 2, 3, 4):
            for (w, h) in ((64, 48), (37, 21), (5, 6), (130, 9), (1, 7)):
                for (dw, dh) in ((33, 47), (128, 96), (9, 130), (200, 3), (18, 10)):
                    info = np.iinfo(dt)
                    src = rng.integers(info.min, info.max + 1, (h, w, cn)).astype(dt)
                    if n % 5 == 0:
                        src[:] = info.min
                    if n % 7 == 0:
                        src[:] = info.max
                    want = np.zeros((dh, dw, cn), dt)
                    rc = lib.cvref_cv_resize(cv_type(src), src.ctypes.data_as(C.c_void_p), C.c_size_t(src.strides[0]), w, h, want.ctypes.data_as(C.c_void_p),
                                             C.c_size_t(want.strides[0]), dw, dh, C.c_double(0), C.c_double(0), INTER_LINEAR_EXACT, None)
                    if rc != 0:
                        continue                   # exact 2x on 8S / 32S: the reference has no area function for these depths and throws
                    got = np.zeros_like(want)
                    assert O.resize(src, got, INTER_LINEAR_EXACT) == 0
                    assert (got == want).all(), f"{dt.__name__}C{cn} {w}x{h}->{dw}x{dh}"
                    n += 1
    assert n > 250


@needs_ref
def test_reference_with_more_than_four_channels_equals_groups_of_four():
    # what hal_warp relies on for cn > 4: the reference's warpAffine / warpPerspective / remap on a 6- or 9-channel image give, channel
    # group by channel group, what they give on the 4-channel groups alone with the same border value (cval[k] = borderValue[k & 3])
    R = cvref()
    rng = np.random.default_rng(77)
    sh, sw, dh, dw = 61, 83, 70, 90
    M = rotation_inverse(sw, sh, 20, 0.9)
    Hm = rand_homography_inverse(rng, sw, sh)
    mx = rng.uniform(-10, sw + 10, (dh, dw)).astype(np.float32); my = rng.uniform(-10, sh + 10, (dh, dw)).astype(np.float32)
    bv = (1.5, 22, 3, 47)
    for dtype in (np.uint8, np.float32):
        for cn in (6, 9):
            src = np.stack([rand_image(rng, sh, sw, 1, dtype) for _ in range(cn)], -1)
            assert R.warpAffine(src, np.zeros((dh, dw, cn), dtype), M, INTER_CUBIC, BORDER_CONSTANT, bv) != 0      # cn <= 4 asserted for CUBIC / LANCZOS4
            for interp, border in ((INTER_NEAREST, BORDER_CONSTANT), (INTER_LINEAR, BORDER_CONSTANT), (INTER_LINEAR, BORDER_REFLECT_101), (INTER_AREA, BORDER_WRAP)):
                for call in (lambda L, s_, d_: L.warpAffine(s_, d_, M, interp, border, bv),
                             lambda L, s_, d_: L.warpPerspective(s_, d_, Hm, interp, border, bv),
                             lambda L, s_, d_: L.remap32f(s_, d_, mx, my, interp, border, bv)):
                    full = np.zeros((dh, dw, cn), dtype)
                    assert call(R, src, full) == 0
                    for c0 in range(0, cn, 4):
                        k = min(4, cn - c0)
                        sg = np.ascontiguousarray(src[..., c0:c0 + k]); sg = sg.reshape(sh, sw) if k == 1 else sg
                        want = np.zeros((dh, dw) if k == 1 else (dh, dw, k), dtype)
                        assert call(O, sg, want) == 0
                        assert_same(full[..., c0:c0 + k].reshape(want.shape), want, f"cn{cn} group at {c0} interp{interp} border{border}")


@needs_ref
def test_reference_resize_with_more_than_four_channels_equals_groups_of_four():
    # what b200_hal_resize relies on for cn > 4: where cv::resize treats channels alike whatever their number, the reference on 5 / 6 / 9
    # channels equals the oracle on the groups of 4.  It does NOT in three places, which the hook leaves to the CPU (resize_cn_gt4_declined):
    # the 2x2 decimation (1 / 3 / 4 channels: vector path, round half up; more: generic path, round half to even), fractional INTER_AREA
    # (CV_Assert(cn <= 4), resize.cpp:3851) and floating-point CUBIC / LANCZOS4 (vector body / scalar tail split moves with width * cn)
    R = cvref()
    rng = np.random.default_rng(78)
    n = differs = 0
    for dtype in (np.uint8, np.uint16, np.int16, np.float32, np.float64):
        for cn in (5, 6, 9):
            for (sh, sw, dh, dw) in ((61, 83, 70, 97), (64, 96, 32, 48), (90, 120, 30, 40), (90, 120, 31, 47), (60, 80, 120, 160)):
                src = (rng.random((sh, sw, cn)) * 1000 - 300) if dtype == np.float64 else np.stack([rand_image(rng, sh, sw, 1, dtype) for _ in range(cn)], -1)
                for interp in (INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4, INTER_AREA, INTER_LINEAR_EXACT, INTER_NEAREST_EXACT):
                    if interp == INTER_LINEAR_EXACT and dtype in (np.float32, np.float64):
                        continue
                    full = np.zeros((dh, dw, cn), dtype)
                    rc = R.resize(src, full, interp)
                    declined = resize_cn_gt4_declined(dtype, interp, sw, sh, dw, dh)
                    if rc != 0:
                        assert declined and interp == INTER_AREA             # the fractional area path
                        continue
                    same = True
                    for c0 in range(0, cn, 4):
                        k = min(4, cn - c0)
                        sg = np.ascontiguousarray(src[..., c0:c0 + k]); sg = sg.reshape(sh, sw) if k == 1 else sg
                        want = np.zeros((dh, dw) if k == 1 else (dh, dw, k), dtype)
                        assert O.resize(sg, want, interp) == 0
                        same = same and (np.ascontiguousarray(full[..., c0:c0 + k]).reshape(want.shape).view(np.uint8) == want.view(np.uint8)).all()
                    assert same or declined, f"{dtype.__name__} cn{cn} interp{interp} {sw}x{sh}->{dw}x{dh}"
                    differs += not same
                    n += 1
    assert n > 400 and differs > 20


@needs_ref
def test_warppolar_forward_restatement_vs_reference():
    # orc_warpPolar_maps + the oracle's remap against cv::warpPolar (imgwarp.cpp:3163-3224), linear and semilog, outliers filled (BORDER_CONSTANT)
    # or kept (BORDER_TRANSPARENT): the image is identical, so the maps are (cos / sin / exp / log come from the same host libm)
    lib = C.CDLL(cvref_path())
    lib.cvref_warpPolar.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_float, C.c_double, C.c_int]
    rng = np.random.default_rng(31)
    n = 0
    for (dt, cn, (sw, sh), (dw, dh), (cx, cy), rad, flags) in warppolar_cases():
        src = rand_image(rng, sh, sw, cn, dt)
        want = np.full((dh, dw) if cn == 1 else (dh, dw, cn), 7, dt); got = want.copy()
        assert lib.cvref_warpPolar(cv_type(src), src.ctypes.data, src.strides[0], sw, sh, want.ctypes.data, want.strides[0], dw, dh, cx, cy, rad, flags) == 0
        assert oracle_warpPolar(src, got, cx, cy, rad, flags) == 0
        assert_same(got, want, f"warpPolar {dt.__name__}C{cn} {sw}x{sh}->{dw}x{dh} flags {flags}")
        n += 1
    assert n >= 75
