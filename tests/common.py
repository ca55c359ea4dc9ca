"""Shared test plumbing: ctypes bindings for the three implementations of the hook ABI.

  * ``oracle()``   -- oracle/libgeom_oracle.so  (plain-C restatement, prefix ``orc_``)
  * ``cvref()``    -- oracle/_ref/libcvref.so   (the reference's own sources, prefix ``cvref_``)
  * the CUDA C-ABI is bound by ``opencv_b200.capi`` (prefix ``b200_``)

All three share the argument lists of the HAL replacement hooks
(reference: modules/imgproc/src/hal_replacement.hpp:369,387,428,483-487,506-509,530-534),
so one set of numpy-level callers drives every implementation.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CV_8U, CV_16U, CV_16S, CV_32F = 0, 2, 3, 5
INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_AREA = 0, 1, 2, 3
INTER_LANCZOS4 = 4
INTER_LINEAR_EXACT, INTER_NEAREST_EXACT = 5, 6
BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101 = 0, 1, 2, 3, 4
WARP_INVERSE_MAP = 16
WARP_RELATIVE_MAP = 32

_DEPTH_OF = {np.dtype(np.uint8): CV_8U, np.dtype(np.uint16): CV_16U, np.dtype(np.int16): CV_16S,
             np.dtype(np.float32): CV_32F, np.dtype(np.float64): 6, np.dtype(np.int8): 1, np.dtype(np.int32): 4}


def cv_type(arr):
    cn = 1 if arr.ndim == 2 else arr.shape[2]
    return _DEPTH_OF[arr.dtype] + ((cn - 1) << 3)


_vp, _sz, _i, _d = C.c_void_p, C.c_size_t, C.c_int, C.c_double
_IMG = [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i]
SIGS = {
    "warpAffine": _IMG + [_vp, _i, _i, _vp],
    "warpPerspective": _IMG + [_vp, _i, _i, _vp],
    "remap32f": _IMG + [_vp, _sz, _vp, _sz, _i, _i, _vp],
    "remap32fc2": _IMG + [_vp, _sz, _i, _i, _vp],
    "remap16s": _IMG + [_vp, _sz, _vp, _sz, _i, _i, _vp],
    "resize": _IMG + [_d, _d, _i],
}


class HookLib:
    """numpy-level callers over one implementation of the hook ABI."""

    def __init__(self, lib, names):
        self.lib = lib
        self.fn = {}
        for op, sym in names.items():
            f = getattr(lib, sym)
            f.argtypes = SIGS[op]
            f.restype = _i
            self.fn[op] = f

    @staticmethod
    def _img(a):
        assert a.shape[-1] == 1 or a.strides[-1] == a.itemsize
        h, w = a.shape[:2]
        return a.ctypes.data, a.strides[0], w, h

    def _call(self, op, src, dst, *rest):
        sp, ss, sw, sh = self._img(src)
        dp, ds, dw, dh = self._img(dst)
        return self.fn[op](cv_type(src), sp, ss, sw, sh, dp, ds, dw, dh, *rest)

    @staticmethod
    def _bv(bv):
        bv = list(bv) + [0.0] * (4 - len(bv))
        return (C.c_double * 4)(*bv)

    def warpAffine(self, src, dst, M, interp, border, bv=(0, 0, 0, 0)):
        Mc = (C.c_double * 6)(*np.asarray(M, np.float64).ravel())
        return self._call("warpAffine", src, dst, Mc, interp, border, self._bv(bv))

    def warpPerspective(self, src, dst, M, interp, border, bv=(0, 0, 0, 0)):
        Mc = (C.c_double * 9)(*np.asarray(M, np.float64).ravel())
        return self._call("warpPerspective", src, dst, Mc, interp, border, self._bv(bv))

    def remap32f(self, src, dst, mapx, mapy, interp, border, bv=(0, 0, 0, 0)):
        return self._call("remap32f", src, dst, mapx.ctypes.data, mapx.strides[0], mapy.ctypes.data,
                          mapy.strides[0], interp, border, self._bv(bv))

    def remap32fc2(self, src, dst, mp, interp, border, bv=(0, 0, 0, 0)):
        return self._call("remap32fc2", src, dst, mp.ctypes.data, mp.strides[0], interp, border, self._bv(bv))

    def remap16s(self, src, dst, m1, m2, interp, border, bv=(0, 0, 0, 0)):
        p2, s2 = (m2.ctypes.data, m2.strides[0]) if m2 is not None else (None, 0)
        return self._call("remap16s", src, dst, m1.ctypes.data, m1.strides[0], p2, s2, interp, border, self._bv(bv))

    def resize(self, src, dst, interp, inv_scale=None):
        if inv_scale is None:
            inv_scale = (dst.shape[1] / src.shape[1], dst.shape[0] / src.shape[0])
        return self._call("resize", src, dst, float(inv_scale[0]), float(inv_scale[1]), interp)


_cache = {}


def oracle():
    if "orc" not in _cache:
        path = os.path.join(ROOT, "oracle", "libgeom_oracle.so")
        src = os.path.join(ROOT, "oracle", "geom_oracle.c")
        if not os.path.exists(path) or os.path.getmtime(path) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
        lib = C.CDLL(path)
        _cache["orc"] = HookLib(lib, {k: "orc_" + k for k in SIGS})
    return _cache["orc"]


def cvref_path():
    return os.path.join(ROOT, "oracle", "_ref", "libcvref.so")


def cvref():
    """The reference's own code (None when oracle/_ref has not been built)."""
    if "ref" not in _cache:
        p = cvref_path()
        if not os.path.exists(p):
            _cache["ref"] = None
        else:
            lib = C.CDLL(p)
            names = {"warpAffine": "cvref_hal_warpAffine", "warpPerspective": "cvref_hal_warpPerspective",
                     "resize": "cvref_hal_resize", "remap32f": "cvref_remap32f",
                     "remap32fc2": "cvref_remap32fc2", "remap16s": "cvref_remap16s"}
            _cache["ref"] = HookLib(lib, names)
    return _cache["ref"]


# ---------------------------------------------------------------- input generators ----
def rand_image(rng, h, w, cn, dtype):
    shape = (h, w) if cn == 1 else (h, w, cn)
    if dtype == np.uint8:
        return rng.integers(0, 256, shape, dtype=np.uint8)
    if dtype == np.uint16:
        return rng.integers(0, 65536, shape, dtype=np.uint16)
    if dtype == np.int16:
        return rng.integers(-32768, 32768, shape, dtype=np.int16)
    return (rng.random(shape, dtype=np.float32) * 2000 - 1000).astype(np.float32)


def rotation_inverse(w, h, angle_deg, scale, cx=None, cy=None):
    """inverse (dst->src) 2x3 matrix of getRotationMatrix2D, inverted the way cv::warpAffine does
    (reference imgwarp.cpp:2401-2411)."""
    cx = (w - 1) * 0.5 if cx is None else cx
    cy = (h - 1) * 0.5 if cy is None else cy
    a = np.deg2rad(angle_deg)
    al, be = scale * np.cos(a), scale * np.sin(a)
    M = np.array([al, be, (1 - al) * cx - be * cy, -be, al, be * cx + (1 - al) * cy], np.float64)
    out = (C.c_double * 6)()
    oracle().lib.orc_invert_affine((C.c_double * 6)(*M), out)
    return np.array(out[:], np.float64).reshape(2, 3)


def rand_homography_inverse(rng, w, h, jitter=0.15):
    """dst->src 3x3 homography from randomly jittered corners (numpy solve, float64)."""
    src = np.array([[0, 0], [w - 1, 0], [w - 1, h - 1], [0, h - 1]], np.float64)
    dst = src + rng.uniform(-jitter, jitter, (4, 2)) * np.array([w, h])
    A, b = [], []
    for (x, y), (u, v) in zip(dst, src):      # maps dst -> src directly
        A.append([x, y, 1, 0, 0, 0, -u * x, -u * y]); b.append(u)
        A.append([0, 0, 0, x, y, 1, -v * x, -v * y]); b.append(v)
    hh = np.linalg.solve(np.array(A), np.array(b))
    return np.append(hh, 1.0).reshape(3, 3)


def fisheye_maps(w, h, f=None):
    """SURVEY 8(d) C4 fisheye-undistort map (fp64 -> fp32)."""
    f = 600.0 * w / 1280.0 if f is None else f
    cx, cy = (w - 1) * 0.5, (h - 1) * 0.5
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    X, Y = (xs - cx) / f, (ys - cy) / f
    r = np.sqrt(X * X + Y * Y)
    th = np.arctan(r)
    thd = th * (1 - 0.05 * th ** 2 + 0.01 * th ** 4)
    s = np.where(r < 1e-8, 1.0, thd / np.maximum(r, 1e-300))
    return (f * X * s + cx).astype(np.float32), (f * Y * s + cy).astype(np.float32)


def convert_maps_fixed(mapx, mapy):
    """cv::convertMaps(..., CV_16SC2, nninterpolation=false) restated
    (reference imgwarp.cpp:1760-1800: ix=saturate_cast<int>(x*32); xy=sat16(ix>>5); a=(iy&31)*32+(ix&31))."""
    def rne_i32(v):
        r = np.rint(v.astype(np.float32) * np.float32(32.0)).astype(np.float64)
        bad = ~((r >= -2147483648.0) & (r <= 2147483647.0))
        r = np.where(bad, -2147483648.0, r)
        return r.astype(np.int64).astype(np.int32)
    ix, iy = rne_i32(mapx), rne_i32(mapy)
    m1 = np.stack([np.clip(ix >> 5, -32768, 32767), np.clip(iy >> 5, -32768, 32767)], -1).astype(np.int16)
    m2 = ((iy & 31) * 32 + (ix & 31)).astype(np.uint16)
    return np.ascontiguousarray(m1), np.ascontiguousarray(m2)


def assert_same(a, b, what=""):
    """bit-exact for integer types; value-exact (so +0 == -0, NaN == NaN) for float."""
    assert a.shape == b.shape and a.dtype == b.dtype, what
    if a.dtype == np.float32:
        ok = (a == b) | (np.isnan(a) & np.isnan(b))
        n = int((~ok).sum())
        assert n == 0, f"{what}: {n} float mismatches, max abs diff {np.nanmax(np.abs(a.astype(np.float64) - b)):g}"
    else:
        n = int((a != b).sum())
        assert n == 0, f"{what}: {n} of {a.size} values differ (max |d|={np.abs(a.astype(np.int64) - b).max()})"


# ---------------------------------------------------------------- cv::convertMaps cases ----
CV_16UC1, CV_16SC1, CV_32FC1, CV_16SC2, CV_32FC2 = 2, 3, 5, 11, 13
CONVERT_MAPS_SIG = [_i, _vp, _sz, _i, _vp, _sz, _i, _vp, _sz, _vp, _sz, _i, _i, _i]


def convert_maps_cases(rng, h=37, w=61):
    """(map1_type, map1, map2_type, map2, dstmap1_type, nninterpolation) over every pair cv::convertMaps accepts
    (imgwarp.cpp:1693-1696), with NaN / inf / out-of-int-range entries in the float maps.
    Left out: CV_16SC2 -> CV_32FC1 with nninterpolation, where the reference writes through a null pointer."""
    mx = rng.uniform(-50, 700, (h, w)).astype(np.float32)
    my = rng.uniform(-50, 500, (h, w)).astype(np.float32)
    mx[0, :6] = [np.nan, np.inf, -np.inf, 3e9, -3e9, 40000.7]
    my[1, :4] = [0.015625, 0.984375, 32767.49, -32768.6]
    mxy = np.ascontiguousarray(np.stack([mx, my], -1))
    s2 = rng.integers(-300, 900, (h, w, 2)).astype(np.int16)
    u1 = rng.integers(0, 65536, (h, w)).astype(np.uint16)
    out = []
    for nn in (0, 1):
        for dt in (CV_16SC2, CV_32FC1, CV_32FC2, 0):
            out += [(CV_32FC1, mx, CV_32FC1, my, dt, nn), (CV_32FC2, mxy, 0, None, dt, nn), (CV_16SC2, s2, CV_16UC1, u1, dt, nn)]
            if nn and dt != CV_32FC1:
                out.append((CV_16SC2, s2, 0, None, dt, nn))
    return [c for c in out if not (c[0] == CV_16SC2 and c[4] == CV_32FC1 and c[5])]


def convert_maps_outputs(case):
    t1, a1, _t2, _a2, dt, _nn = case
    h, w = a1.shape[:2]
    eff = dt if dt > 0 else (CV_32FC2 if t1 == CV_16SC2 else CV_16SC2)
    d1 = {CV_16SC2: np.full((h, w, 2), -7, np.int16), CV_32FC1: np.full((h, w), -7, np.float32),
          CV_32FC2: np.full((h, w, 2), -7, np.float32)}[eff]
    d2 = np.full((h, w), 7, np.uint16) if eff == CV_16SC2 else np.full((h, w), 7, np.float32)
    return d1, d2


def call_convert_maps(fn, case, d1, d2, *extra):
    t1, a1, t2, a2, dt, nn = case
    h, w = a1.shape[:2]
    return fn(t1, a1.ctypes.data, a1.strides[0], t2, a2.ctypes.data if a2 is not None else None,
              a2.strides[0] if a2 is not None else 0, dt, d1.ctypes.data, d1.strides[0], d2.ctypes.data, d2.strides[0],
              w, h, nn, *extra)


# ---------------------------------------------------------------- dnn::blobFromImages cases ----
BLOB_SIG = [_i, _vp, _sz, _sz, _i, _i, _vp, _i, _i, _i, _vp, _vp, _i]
BLOB_CASES = [((120, 160), (64, 64)), ((120, 160), (60, 80)), ((97, 131), (224, 224)), ((50, 70), (50, 70)),
              ((121, 161), (60, 80)), ((96, 128), (32, 32))]     # generic linear, exact 2x box, upscale, no resize, odd /2, integer 3x


BLOB_PAD_SIG = BLOB_SIG + [_i, _vp]
# (source (h, w), blob (h, w)) for the padding modes: wide -> square, tall -> square, upscale, exact 2x in one axis, crop of one pixel
BLOB_PAD_CASES = [((480, 640), (224, 224)), ((500, 300), (320, 240)), ((97, 33), (70, 50)), ((64, 128), (32, 32)), ((64, 64), (128, 128)),
                  ((720, 1280), (640, 640)), ((100, 101), (50, 50))]


def blob_params(cn, swap):
    if cn == 1 and not swap:
        return (C.c_double * 4)(1, 1, 1, 1), (C.c_double * 4)(0, 0, 0, 0)
    return (C.c_double * 4)(1 / 127.5, 0.5, 2.0, 1.0), (C.c_double * 4)(104.0, 117.5, 123.25, 0.0)


# ---------------------------------------------------------------- cv::fisheye::initUndistortRectifyMap cases ----
def fisheye_cases():
    """(name, K, D, R, P, (w, h)): a wide-angle camera as in BASELINE config 4, one with rectification + new camera matrix,
    one whose rotation puts part of the image behind the camera (_w <= 0: the reference writes +-inf there), and one whose
    principal point is a pixel centre with P = K (r == 0 at that pixel)."""
    K = np.array([[600., 0, 639.5], [0, 600., 359.5], [0, 0, 1]])
    D = np.array([-0.05, 0.01, 0.002, -0.0005])
    Rm = np.array([[0.9998, -0.01, 0.015], [0.0102, 0.9999, -0.008], [-0.0149, 0.0082, 0.9998]])
    P = np.array([[520., 0, 640], [0, 520., 360], [0, 0, 1.]])
    c, s = np.cos(1.35), np.sin(1.35)
    Rbehind = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
    Kc = np.array([[300., 0, 160.], [0, 300., 120.], [0, 0, 1]])
    return [("wide 720p", K, D, None, None, (1280, 720)), ("rectified 720p", K, D, Rm, P, (1280, 720)),
            ("behind the camera", Kc, D, Rbehind, Kc, (321, 243)), ("r == 0 pixel", Kc, np.zeros(4), None, Kc, (320, 240)),
            ("odd width", K, D * 3, Rm, P, (1001, 37))]


def fisheye_map_buffers(w, h, m1type):
    if m1type == 11:
        return np.zeros((h, w, 2), np.int16), np.zeros((h, w), np.uint16)
    return np.zeros((h, w), np.float32), np.zeros((h, w), np.float32)


def _dp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def ref_fisheye_map(case, m1type):
    """the reference's own cv::fisheye::initUndistortRectifyMap (oracle/_ref) -> (map1, map2, iR the reference used)"""
    _name, K, D, R, P, (w, h) = case
    lib = C.CDLL(cvref_path())
    m1, m2 = fisheye_map_buffers(w, h, m1type)
    iR = np.zeros(9)
    Kc, Dc = np.ascontiguousarray(K), np.ascontiguousarray(D)
    Rc = np.ascontiguousarray(R) if R is not None else None
    Pc = np.ascontiguousarray(P) if P is not None else None
    rc = lib.cvref_fisheye_initUndistortRectifyMap(_dp(Kc), _dp(Dc), _dp(Rc), _dp(Pc), w, h, m1type, _dp(m1), C.c_size_t(m1.strides[0]),
                                                   _dp(m2), C.c_size_t(m2.strides[0]), _dp(iR))
    assert rc == 0
    return m1, m2, iR


def oracle_fisheye_map(case, m1type, iR):
    _name, K, D, _R, _P, (w, h) = case
    m1, m2 = fisheye_map_buffers(w, h, m1type)
    Kc, Dc, iRc = np.ascontiguousarray(K), np.ascontiguousarray(D), np.ascontiguousarray(iR)
    rc = oracle().lib.orc_fisheye_initUndistortRectifyMap(_dp(Kc), _dp(Dc), _dp(iRc), w, h, m1type, _dp(m1), C.c_size_t(m1.strides[0]),
                                                          _dp(m2), C.c_size_t(m2.strides[0]))
    assert rc == 0
    return m1, m2


# ---------------------------------------------------------------- cv::initUndistortRectifyMap cases ----
def undistort_cases():
    """(name, K, dist, R, newK, (w, h)): 5-, 8-, 14-coefficient (tilted sensor) and zero models; widths that end in a scalar tail
    (w % 8 != 0), a row shorter than one vector chunk, rectification + new camera matrix"""
    K = np.array([[1000., 0, 640.3], [0, 1010., 359.8], [0, 0, 1]])
    Rm = np.array([[0.9998, -0.01, 0.015], [0.0102, 0.9999, -0.008], [-0.0149, 0.0082, 0.9998]])
    P = np.array([[900., 0, 630], [0, 905., 365], [0, 0, 1.]])
    d5 = np.array([-0.28, 0.09, 0.001, -0.0005, -0.012])
    d8 = np.array([-0.28, 0.09, 0.001, -0.0005, -0.012, 0.01, -0.02, 0.003])
    d14 = np.array([-0.28, 0.09, 0.001, -0.0005, -0.012, 0.01, -0.02, 0.003, 0.001, -0.002, 0.0015, 0.0007, 0.01, -0.015])
    return [("5 coeffs 720p", K, d5, None, None, (1280, 720)), ("8 coeffs rectified", K, d8, Rm, P, (1283, 37)),
            ("14 coeffs tilted", K, d14, Rm, P, (1001, 53)), ("no distortion", K, np.zeros(5), None, P, (644, 11)),
            ("narrower than a chunk", K, d5, Rm, P, (7, 5)), ("wide", K, d14, None, None, (4099, 3))]


def undistort_buffers(w, h, m1type):
    if m1type == 11:
        return np.zeros((h, w, 2), np.int16), np.zeros((h, w), np.uint16)
    if m1type == 5:
        return np.zeros((h, w), np.float32), np.zeros((h, w), np.float32)
    return np.zeros((h, w, 2), np.float32), None


def ref_undistort_map(case, m1type):
    """the reference's cv::initUndistortRectifyMap (oracle/_ref) -> (map1, map2, iR, tilt) with the reference's own iR / tilt"""
    _n, K, dist, R, P, (w, h) = case
    lib = C.CDLL(cvref_path())
    m1, m2 = undistort_buffers(w, h, m1type)
    iR, tilt = np.zeros(9), np.zeros(9)
    Kc, Dc = np.ascontiguousarray(K), np.ascontiguousarray(dist)
    Rc = np.ascontiguousarray(R) if R is not None else None
    Pc = np.ascontiguousarray(P) if P is not None else None
    rc = lib.cvref_initUndistortRectifyMap(_dp(Kc), _dp(Dc), len(Dc), _dp(Rc), _dp(Pc), w, h, m1type, _dp(m1), C.c_size_t(m1.strides[0]),
                                           _dp(m2), C.c_size_t(m2.strides[0] if m2 is not None else 0), _dp(iR), _dp(tilt))
    assert rc == 0
    return m1, m2, iR, tilt


def dist12_of(dist):
    d = np.zeros(12)
    d[:min(12, len(dist))] = np.asarray(dist)[:12]
    return d


def oracle_undistort_map(case, m1type, iR, tilt):
    _n, K, dist, _R, _P, (w, h) = case
    m1, m2 = undistort_buffers(w, h, m1type)
    Kc, d12, iRc, tc = np.ascontiguousarray(K), dist12_of(dist), np.ascontiguousarray(iR), np.ascontiguousarray(tilt)
    rc = oracle().lib.orc_initUndistortRectifyMap(_dp(Kc), _dp(d12), _dp(iRc), _dp(tc), w, h, m1type, _dp(m1), C.c_size_t(m1.strides[0]),
                                                  _dp(m2), C.c_size_t(m2.strides[0] if m2 is not None else 0))
    assert rc == 0
    return m1, m2


def oracle_undistort_params(newK, R, tau_x, tau_y):
    iR, tilt = np.zeros(9), np.zeros(9)
    Rc = None if R is None else np.ascontiguousarray(R, np.float64)
    assert oracle().lib.orc_undistort_params(_dp(np.ascontiguousarray(newK, np.float64)), _dp(Rc), C.c_double(tau_x), C.c_double(tau_y), _dp(iR), _dp(tilt)) == 0
    return iR, tilt


def undistort_image_cases():
    """(name, K, dist, newK, (w, h)) for cv::undistort: stripes of 3 rows (1280 wide), of 6 with a short last stripe, of one row
    (wider than 4096), the image as one stripe; 14 coefficients with tilt"""
    K = np.array([[1000., 0, 640.3], [0, 1010., 359.8], [0, 0, 1]])
    P = np.array([[900., 0.3, 630], [0, 905., 365], [0, 0, 1.]])
    d14 = np.array([-0.28, 0.09, 0.001, -0.0005, -0.012, 0.01, -0.02, 0.003, 0.001, -0.002, 0.0015, 0.0007, 0.01, -0.015])
    return [("720p, 5 coeffs", K, d14[:5], None, (1280, 720)), ("643x37, 14 coeffs, new matrix", K, d14, P, (643, 37)),
            ("4099x9, 8 coeffs", K, d14[:8], None, (4099, 9)), ("one stripe", K, d14[:12], P, (100, 30))]


def oracle_undistort(case, src):
    _n, K, dist, newK, (w, h) = case
    tilt = oracle_undistort_params(K, None, dist[12], dist[13])[1] if len(dist) >= 14 else np.eye(3).ravel().copy()
    dst = np.zeros_like(src)
    m1, m2 = np.zeros((h, w, 2), np.int16), np.zeros((h, w), np.uint16)
    Kc, d12 = np.ascontiguousarray(K), dist12_of(dist)
    Pc = None if newK is None else np.ascontiguousarray(newK)
    rc = oracle().lib.orc_undistort(cv_type(src), _dp(src), C.c_size_t(src.strides[0]), w, h, _dp(dst), C.c_size_t(dst.strides[0]), _dp(Kc), _dp(d12),
                                    _dp(Pc), _dp(tilt), _dp(m1), C.c_size_t(w * 4), _dp(m2), C.c_size_t(w * 2))
    assert rc == 0
    return dst, m1, m2, tilt


def tilt_matrix(tau_x, tau_y):
    return oracle_undistort_params(np.eye(3), None, tau_x, tau_y)[1].reshape(3, 3)


# ---------------------------------------------------------------- cv::pyrDown ----
def pyrdown_cases():
    """(dtype, cn, (w, h), border, (dw, dh)): every depth x channel count the kernel takes, odd and even sizes, images smaller than
    the filter, rows shorter than one vector step, every border, and the non-default sizes pyrDown_ accepts (|2*dw - w| <= 2)"""
    out = []
    for dt in (np.uint8, np.uint16, np.int16, np.float32):
        for cn in (1, 2, 3, 4):
            for (w, h) in ((64, 48), (37, 21), (5, 6), (1, 1), (2, 3), (131, 9), (7, 7)):
                for border in (BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101):
                    for (a, b) in ((0, 0), (1, 1), (-1, -1)):
                        dw, dh = (w + 1) // 2 + a, (h + 1) // 2 + b
                        if dw >= 1 and dh >= 1 and abs(dw * 2 - w) <= 2 and abs(dh * 2 - h) <= 2 and ((a, b) == (0, 0) or border == BORDER_REFLECT_101):
                            out.append((dt, cn, (w, h), border, (dw, dh)))
    return out


def pyr_image(rng, dt, h, w, cn):
    if dt == np.float32:
        return (rng.random((h, w, cn)) * 1000 - 300).astype(np.float32)
    info = np.iinfo(dt)
    return rng.integers(info.min, info.max + 1, (h, w, cn)).astype(dt)


def oracle_pyrDown(src, dw, dh, border):
    h, w, cn = src.shape
    dst = np.zeros((dh, dw, cn), src.dtype)
    rc = oracle().lib.orc_pyrDown(cv_type(src), _dp(src), C.c_size_t(src.strides[0]), w, h, _dp(dst), C.c_size_t(dst.strides[0]), dw, dh, border)
    assert rc == 0
    return dst


def resize_cn_gt4_declined(dtype, interp, sw, sh, dw, dh):
    """the cells b200_hal_resize leaves to the CPU for more than 4 channels (capi.cu, b200_hal_resize)"""
    two = sw == 2 * dw and sh == 2 * dh
    down = dw <= sw and dh <= sh
    if two and interp in (INTER_LINEAR, INTER_AREA, INTER_LINEAR_EXACT):
        return True
    if interp == INTER_AREA and down and (sw % dw or sh % dh):
        return True
    return interp in (INTER_CUBIC, INTER_LANCZOS4) and np.dtype(dtype).kind == "f"


# ---------------------------------------------------------------- cv::pyrUp ----
def pyrup_cases():
    """(dtype, cn, (w, h), (dw, dh)): every depth x channel count cv::pyrUp takes, sizes down to one pixel, rows shorter than a
    vector step, and the three destination sizes per axis pyrUp_ accepts (2n, 2n - 1, 2n + 1)"""
    out = []
    for dt in (np.uint8, np.uint16, np.int16, np.float32, np.float64):
        for cn in (1, 2, 3, 4):
            for (w, h) in ((32, 24), (19, 11), (3, 4), (1, 1), (2, 1), (1, 3), (67, 5)):
                for (a, b) in ((0, 0), (-1, -1), (1, 1), (1, -1)):
                    dw, dh = 2 * w + a, 2 * h + b
                    if not (w == 1 and dw == 3):                # the reference reads an unwritten cell there
                        out.append((dt, cn, (w, h), (dw, dh)))
    return out


def pyrup_image(rng, dt, h, w, cn):
    if dt == np.float64:
        return rng.random((h, w, cn)) * 1000 - 300
    return pyr_image(rng, dt, h, w, cn)


def oracle_pyrUp(src, dw, dh):
    h, w, cn = src.shape
    dst = np.zeros((dh, dw, cn), src.dtype)
    rc = oracle().lib.orc_pyrUp(cv_type(src), _dp(src), C.c_size_t(src.strides[0]), w, h, _dp(dst), C.c_size_t(dst.strides[0]), dw, dh, BORDER_REFLECT_101)
    assert rc == 0
    return dst


# ---------------------------------------------------------------- cv::warpPolar (forward) ----
WARP_FILL_OUTLIERS, WARP_INVERSE_MAP, WARP_POLAR_LOG, BORDER_TRANSPARENT = 8, 16, 256, 5


def warppolar_cases():
    """(dtype, cn, (sw, sh), (dw, dh), (cx, cy), maxRadius, flags): linear and semilog, outliers filled or kept, centres inside and
    off-centre, radii reaching past the image"""
    out = []
    for dt, cn in ((np.uint8, 1), (np.uint8, 3), (np.uint16, 4), (np.float32, 1), (np.float32, 3)):
        for (sw, sh), (dw, dh), (cx, cy), rad in (((160, 120), (100, 200), (80.0, 60.0), 75.0), ((131, 97), (64, 181), (40.25, 70.5), 120.0),
                                                  ((64, 64), (33, 47), (31.5, 31.5), 31.0)):
            for flags in (INTER_LINEAR | WARP_FILL_OUTLIERS, INTER_LINEAR, INTER_NEAREST | WARP_FILL_OUTLIERS | WARP_POLAR_LOG,
                          INTER_CUBIC | WARP_FILL_OUTLIERS, INTER_LINEAR | WARP_POLAR_LOG):
                out.append((dt, cn, (sw, sh), (dw, dh), (cx, cy), rad, flags))
    return out


def oracle_warpPolar_maps(dw, dh, cx, cy, rad, semilog):
    mx, my = np.zeros((dh, dw), np.float32), np.zeros((dh, dw), np.float32)
    f = oracle().lib.orc_warpPolar_maps
    f.argtypes = [C.c_int, C.c_int, C.c_float, C.c_float, C.c_double, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    f.restype = C.c_int
    assert f(dw, dh, cx, cy, rad, int(bool(semilog)), _dp(mx), mx.strides[0], _dp(my), my.strides[0]) == 0
    return mx, my


def oracle_warpPolar(src, dst, cx, cy, rad, flags):
    """forward cv::warpPolar = the maps + cv::remap (imgwarp.cpp:3223); dst comes pre-filled"""
    dh, dw = dst.shape[:2]
    mx, my = oracle_warpPolar_maps(dw, dh, cx, cy, rad, flags & WARP_POLAR_LOG)
    return oracle().remap32f(src, dst, mx, my, flags & 7, BORDER_CONSTANT if flags & WARP_FILL_OUTLIERS else BORDER_TRANSPARENT)


# ---------------------------------------------------------------- stitching warpers ----
def stitch_cases():
    """(K, R, T, scale, (w, h)) float32 cameras: small and larger rotations, a translation (plane warper only)"""
    def rot(ax, ay, az):
        cx, sx, cy, sy, cz, sz = np.cos(ax), np.sin(ax), np.cos(ay), np.sin(ay), np.cos(az), np.sin(az)
        Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]); Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
        Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
        return np.ascontiguousarray((Rz @ Ry @ Rx).astype(np.float32))
    K = np.array([[800, 0, 320.5], [0, 805, 239.5], [0, 0, 1]], np.float32)
    return [(K, rot(0.02, -0.1, 0.01), np.zeros(3, np.float32), 800.0, (640, 480)),
            (K, rot(-0.05, 0.3, 0.03), np.array([0.1, -0.05, 0.02], np.float32), 650.5, (640, 480)),
            (K, rot(0.0, 0.0, 0.0), np.zeros(3, np.float32), 1000.0, (333, 201))]


def ref_stitch_maps(kind, case):
    """the reference warper's buildMaps -> (xmap, ymap, (tl_x, tl_y), k_rinv, t)"""
    K, R, T, scale, (w, h) = case
    lib = C.CDLL(cvref_path())
    roi = (C.c_int * 4)(); kr = np.zeros(9, np.float32); tt = np.zeros(3, np.float32)
    a = (kind, w, h, _dp(K), _dp(R), _dp(T), C.c_float(scale))
    assert lib.cvref_stitch_buildMaps(*a, None, C.c_size_t(0), None, C.c_size_t(0), roi, _dp(kr), _dp(tt)) == 0
    xm = np.zeros((roi[3], roi[2]), np.float32); ym = np.zeros_like(xm)
    assert lib.cvref_stitch_buildMaps(*a, _dp(xm), C.c_size_t(xm.strides[0]), _dp(ym), C.c_size_t(ym.strides[0]), roi, _dp(kr), _dp(tt)) == 0
    return xm, ym, (roi[0], roi[1]), kr, tt


def oracle_stitch_maps(kind, kr, tt, scale, tl, w, h):
    xm = np.zeros((h, w), np.float32); ym = np.zeros_like(xm)
    assert oracle().lib.orc_stitch_buildMaps(kind, _dp(kr), _dp(tt), C.c_float(scale), tl[0], tl[1], w, h, _dp(xm), C.c_size_t(w * 4), _dp(ym), C.c_size_t(w * 4)) == 0
    return xm, ym
