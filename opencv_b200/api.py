"""Host-side mirror of cv::warpAffine / cv::warpPerspective / cv::remap / cv::resize over the C ABI."""
import ctypes as C

import numpy as np

from . import capi

INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_AREA, INTER_LANCZOS4 = 0, 1, 2, 3, 4
INTER_LINEAR_EXACT, INTER_NEAREST_EXACT = 5, 6
INTER_MAX = 7
WARP_INVERSE_MAP = 16
WARP_RELATIVE_MAP = 32
BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT, BORDER_WRAP, BORDER_REFLECT_101 = 0, 1, 2, 3, 4

_NP_DEPTH = {"uint8": 0, "int8": 1, "uint16": 2, "int16": 3, "int32": 4, "float32": 5, "float64": 6}


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class _Img:
    """pointer / step / size / type of a HWC image held in numpy (host) or torch (device)."""

    def __init__(self, a):
        self.obj = a
        self.device = _is_torch(a)
        if self.device:
            if not a.is_cuda:
                raise TypeError("torch tensors must live on a CUDA device (pass numpy arrays for the host path)")
            shape, dt = tuple(a.shape), str(a.dtype).replace("torch.", "")
            strides = tuple(s * a.element_size() for s in a.stride())
            self.ptr, item = a.data_ptr(), a.element_size()
        else:
            shape, dt, strides = a.shape, a.dtype.name, a.strides
            self.ptr, item = a.ctypes.data, a.itemsize
        if len(shape) == 2:
            shape, strides = shape + (1,), strides + (item,)
        if len(shape) != 3:
            raise ValueError("expected an HxW or HxWxC image")
        self.h, self.w, self.cn = shape
        if strides[2] != item or strides[1] != item * self.cn:
            raise ValueError("pixels must be densely packed within a row (row stride is free)")
        self.step = strides[0]
        self.type = _NP_DEPTH[dt] + ((self.cn - 1) << 3)


def _alloc_like(src, h, w):
    cn = src.cn
    shape = (h, w) if (src.obj.ndim == 2) else (h, w, cn)
    if src.device:
        import torch
        return torch.empty(shape, dtype=src.obj.dtype, device=src.obj.device)
    return np.empty(shape, src.obj.dtype)


def _stream(img):
    import torch
    return C.c_void_p(torch.cuda.current_stream(img.obj.device).cuda_stream)


def _on_device(img):
    """Make the tensor's own device current for the duration of a *_dev call: the library keys its per-device state
    (interpolation tables, plan caches) on cudaGetDevice(), and the stream handed over belongs to that device."""
    import torch
    return torch.cuda.device(img.obj.device)


def _bv(v):
    if np.isscalar(v):
        v = (v, 0.0, 0.0, 0.0)      # cv::Scalar(v) sets only the first component
    v = list(v) + [0.0] * (4 - len(v))
    return (C.c_double * 4)(*[float(x) for x in v[:4]])


def invert_affine(M):
    """cv::warpAffine's own closed-form inversion, in its exact operation order
    (reference imgwarp.cpp:2401-2411); the hooks receive the result."""
    M = np.array(M, np.float64).reshape(2, 3).copy()
    m = M.ravel()
    D = m[0] * m[4] - m[1] * m[3]
    D = 1. / D if D != 0 else 0.
    A11, A22 = m[4] * D, m[0] * D
    m[0] = A11; m[1] *= -D
    m[3] *= -D; m[4] = A22
    b1 = -m[0] * m[2] - m[1] * m[5]
    b2 = -m[3] * m[2] - m[4] * m[5]
    m[2] = b1; m[5] = b2
    return m.reshape(2, 3)


def _warp(kind, src, M, dsize, flags, borderMode, borderValue, dst):
    s = _Img(src)
    interp = flags & INTER_MAX
    if interp == INTER_AREA:
        interp = INTER_LINEAR                         # imgwarp.cpp:2395, :2855
    n = 6 if kind == "warpAffine" else 9
    M = np.asarray(M)
    if M.dtype not in (np.float32, np.float64) or M.size != n:
        raise ValueError("M must be a float32/float64 %s matrix" % ("2x3" if n == 6 else "3x3"))
    M = M.astype(np.float64)
    if not (flags & WARP_INVERSE_MAP):
        # cv::warpPerspective uses cv::invert (LU); numpy's LAPACK LU may differ in the last ulp, so for
        # bit-exact reproduction of a reference run pass the inverse with WARP_INVERSE_MAP.
        M = invert_affine(M) if n == 6 else np.linalg.inv(M.reshape(3, 3))
    w, h = (s.w, s.h) if not dsize or dsize[0] <= 0 else dsize
    d = _Img(dst) if dst is not None else None
    if d is None:
        dst = _alloc_like(s, h, w)
        d = _Img(dst)
    Mc = (C.c_double * n)(*M.ravel())
    args = [s.type, s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, Mc, interp, borderMode, _bv(borderValue)]
    if s.device:
        fn = getattr(capi.lib, f"b200_{kind}_dev")
        with _on_device(s):
            capi.check(fn(*args, _stream(s)), kind)
    else:
        fn = getattr(capi.lib, f"b200_hal_{kind}")
        capi.check(fn(*args), kind)
    return dst


def warpAffine(src, M, dsize=None, flags=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0, dst=None):
    """cv::warpAffine (reference imgwarp.cpp:2367-2415)."""
    return _warp("warpAffine", src, M, dsize, flags, borderMode, borderValue, dst)


def warpPerspective(src, M, dsize=None, flags=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0, dst=None):
    """cv::warpPerspective (reference imgwarp.cpp:2829-2866)."""
    return _warp("warpPerspective", src, M, dsize, flags, borderMode, borderValue, dst)


def remap(src, map1, map2=None, interpolation=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0, dst=None):
    """cv::remap (reference imgwarp.cpp:1519-1679): map1/map2 = 2xCV_32FC1, CV_32FC2, or CV_16SC2[+CV_16UC1]."""
    s = _Img(src)
    m1 = _Img(map1)
    m2 = _Img(map2) if map2 is not None else None
    if m1.device != s.device or (m2 is not None and m2.device != s.device):
        raise ValueError("src and maps must all be host arrays or all be device tensors")
    d = _Img(dst) if dst is not None else None
    if d is None:
        dst = _alloc_like(s, m1.h, m1.w)
        d = _Img(dst)
    bv = _bv(borderValue)
    img = [s.type, s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h]
    F32C1, F32C2, S16C2, U16C1 = 5, 5 + 8, 3 + 8, 2
    if m1.type == F32C1 and m2 is not None and m2.type == F32C1:
        name, rest = "remap32f", [m1.ptr, m1.step, m2.ptr, m2.step]
    elif m1.type == F32C2 and m2 is None:
        name, rest = "remap32fc2", [m1.ptr, m1.step]
    elif m1.type == S16C2 and (m2 is None or m2.type in (U16C1, 3)):      # CV_16UC1 or CV_16SC1 (imgwarp.cpp:1555: both accepted)
        name, rest = "remap16s", [m1.ptr, m1.step, m2.ptr if m2 else None, m2.step if m2 else 0]
    else:
        raise NotImplementedError("map types not covered by the HAL hooks (imgwarp.cpp:1608-1622)")
    if s.device:
        with _on_device(s):
            capi.check(getattr(capi.lib, f"b200_{name}_dev")(*img, *rest, interpolation, borderMode, bv, _stream(s)), name)
    else:
        capi.check(getattr(capi.lib, f"b200_hal_{name}")(*img, *rest, interpolation, borderMode, bv), name)
    return dst


def _rne(v):
    r = np.rint(v)
    return int(r) if -2147483648.0 <= r <= 2147483647.0 else -2147483648


def resize(src, dsize=None, fx=0., fy=0., interpolation=INTER_LINEAR, dst=None):
    """cv::resize (reference resize.cpp:4007-4052): dsize from (fx, fy) or scales from dsize, same-size copy
    shortcut, then the hook with inv_scale_x/y."""
    s = _Img(src)
    if not dsize or dsize[0] <= 0 or dsize[1] <= 0:
        if not (fx > 0 and fy > 0):
            raise ValueError("either dsize or both fx and fy must be positive")
        dsize = (_rne(s.w * fx), _rne(s.h * fy))                     # resize.cpp:4015-4021
        inv_x, inv_y = float(fx), float(fy)
    else:
        inv_x, inv_y = dsize[0] / s.w, dsize[1] / s.h               # :4023-4027
    w, h = dsize
    if w <= 0 or h <= 0:
        raise ValueError("empty destination size")
    if interpolation == INTER_LINEAR_EXACT and (s.type & 7) in (5, 6):
        interpolation = INTER_LINEAR                                  # resize.cpp:4029
    d = _Img(dst) if dst is not None else None
    if d is None:
        dst = _alloc_like(s, h, w)
        d = _Img(dst)
    if (w, h) == (s.w, s.h):                                         # :4044-4049 never reaches the HAL
        if s.device:
            dst.copy_(src)
        else:
            np.copyto(dst, src)
        return dst
    args = [s.type, s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, inv_x, inv_y, interpolation]
    if s.device:
        with _on_device(s):
            capi.check(capi.lib.b200_resize_dev(*args, _stream(s)), "resize")
    else:
        capi.check(capi.lib.b200_hal_resize(*args), "resize")
    return dst


def pyrDown(src, dstsize=None, borderType=BORDER_REFLECT_101, dst=None):
    """cv::pyrDown (reference imgproc/src/pyramids.cpp:1266-1310): 5x5 binomial filter + decimation by 2; default size
    ((w + 1) / 2, (h + 1) / 2).  numpy arrays go through the cv_hal_pyrdown replacement, CUDA tensors through the device entry."""
    s = _Img(src)
    if borderType == BORDER_CONSTANT:
        raise ValueError("cv::pyrDown does not accept BORDER_CONSTANT")            # pyramids.cpp:1270
    w, h = dstsize if dstsize and dstsize[0] > 0 and dstsize[1] > 0 else ((s.w + 1) // 2, (s.h + 1) // 2)
    if abs(w * 2 - s.w) > 2 or abs(h * 2 - s.h) > 2:
        raise ValueError("dstsize must be within 2 of half the source size")      # :892-894
    d = _Img(dst) if dst is not None else None
    if d is None:
        dst = _alloc_like(s, h, w)
        d = _Img(dst)
    if s.device:
        with _on_device(s):
            capi.check(capi.lib.b200_pyrDown_dev(s.type, s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, borderType, _stream(s)), "pyrDown")
    else:
        capi.check(capi.lib.b200_hal_pyrdown(s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, s.type & 7, s.cn, borderType), "pyrDown")
    return dst


def pyrUp(src, dstsize=None, borderType=BORDER_REFLECT_101, dst=None):
    """cv::pyrUp (reference imgproc/src/pyramids.cpp:1374-1416): upsample by 2 with the binomial kernel; default size (2w, 2h).
    numpy arrays go through the cv_hal_pyrup replacement, CUDA tensors through the device entry."""
    s = _Img(src)
    if borderType != BORDER_REFLECT_101:
        raise ValueError("cv::pyrUp accepts BORDER_DEFAULT only")                  # pyramids.cpp:1378
    w, h = dstsize if dstsize and dstsize[0] > 0 and dstsize[1] > 0 else (s.w * 2, s.h * 2)
    if abs(w - s.w * 2) != w % 2 or abs(h - s.h * 2) != h % 2:
        raise ValueError("dstsize must be 2n, 2n - 1 or 2n + 1 per axis")          # :1055-1056
    d = _Img(dst) if dst is not None else None
    if d is None:
        dst = _alloc_like(s, h, w)
        d = _Img(dst)
    if s.device:
        with _on_device(s):
            capi.check(capi.lib.b200_pyrUp_dev(s.type, s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, borderType, _stream(s)), "pyrUp")
    else:
        capi.check(capi.lib.b200_hal_pyrup(s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, s.type & 7, s.cn, borderType), "pyrUp")
    return dst


def buildPyramid(src, maxlevel, borderType=BORDER_REFLECT_101):
    """cv::buildPyramid (pyramids.cpp:1533-1560): [src, pyrDown(src), pyrDown(pyrDown(src)), ...], maxlevel + 1 images"""
    out = [src]
    for _ in range(maxlevel):
        out.append(pyrDown(out[-1], None, borderType))
    return out


def resize_warpAffine_frames(frames, mid_size, M, dsize=None, resize_interpolation=INTER_LINEAR,
                             flags=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0, dst=None, chunk_frames=0):
    """A stream of host frames through cv::resize(frame, mid_size) then cv::warpAffine(.., M, dsize): `frames` is a
    numpy array (N, H, W[, C]), the result (N, dsize[1], dsize[0][, C]).  Upload, kernels and download of
    consecutive chunks overlap (b200_hal_resize_warpAffine_batch); results equal the per-frame calls."""
    a = np.asarray(frames)
    if a.ndim not in (3, 4) or a.strides[-1] != a.itemsize and a.shape[-1] != 1:
        raise ValueError("frames must be (N, H, W) or (N, H, W, C) with contiguous pixels")
    n, h, w = a.shape[:3]
    cn = 1 if a.ndim == 3 else a.shape[3]
    if a.ndim == 4 and cn > 1 and a.strides[2] != cn * a.itemsize:
        raise ValueError("pixels of a row must be contiguous")
    s_type = _Img(a[0]).type if n else 0
    dw, dh = mid_size if not dsize or dsize[0] <= 0 else dsize
    if dst is None:
        dst = np.empty((n, dh, dw) + (() if a.ndim == 3 else (cn,)), a.dtype)
    if n == 0:
        return dst
    interp = flags & INTER_MAX
    if interp == INTER_AREA:
        interp = INTER_LINEAR
    M = np.asarray(M, np.float64).reshape(2, 3)
    if not (flags & WARP_INVERSE_MAP):
        M = invert_affine(M)
    Mc = (C.c_double * 6)(*M.ravel())
    capi.check(capi.lib.b200_hal_resize_warpAffine_batch(
        s_type, a.ctypes.data, a.strides[0], a.strides[1], w, h, mid_size[0], mid_size[1],
        dst.ctypes.data, dst.strides[0], dst.strides[1], dw, dh, n, resize_interpolation, Mc, interp, borderMode,
        _bv(borderValue), chunk_frames), "resize_warpAffine_frames")
    return dst


def _frames(frames):
    a = np.asarray(frames)
    if a.ndim not in (3, 4):
        raise ValueError("frames must be (N, H, W) or (N, H, W, C)")
    cn = 1 if a.ndim == 3 else a.shape[3]
    if a.strides[-1] != a.itemsize or (a.ndim == 4 and a.strides[2] != cn * a.itemsize):
        raise ValueError("pixels of a row must be contiguous")
    return a, cn


def warpPerspective_frames(frames, M, dsize, flags=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0, chunk_frames=0):
    """A stream of host frames (N, H, W[, C]) through cv::warpPerspective with one matrix: upload, kernels and download of consecutive
    chunks overlap (b200_hal_warpPerspective_batch); results equal the per-frame calls."""
    a, cn = _frames(frames)
    n, h, w = a.shape[:3]
    dw, dh = dsize
    dst = np.empty((n, dh, dw) + (() if a.ndim == 3 else (cn,)), a.dtype)
    if n == 0:
        return dst
    interp = flags & INTER_MAX
    if interp == INTER_AREA:
        interp = INTER_LINEAR
    M = np.asarray(M, np.float64).reshape(3, 3)
    if not (flags & WARP_INVERSE_MAP):
        M = np.linalg.inv(M)
    capi.check(capi.lib.b200_hal_warpPerspective_batch(_Img(a[0]).type, a.ctypes.data, a.strides[0], a.strides[1], w, h, dst.ctypes.data, dst.strides[0],
                                                       dst.strides[1], dw, dh, n, (C.c_double * 9)(*M.ravel()), interp, borderMode, _bv(borderValue), chunk_frames),
               "warpPerspective_frames")
    return dst


def remap_frames(frames, map1, map2, interpolation=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0, chunk_frames=0):
    """A stream of host frames through cv::remap with ONE map (float planes, CV_32FC2, or CV_16SC2 [+ CV_16UC1]): the map is uploaded once,
    the frames stream in chunks (b200_hal_remap_batch)."""
    a, cn = _frames(frames)
    n, h, w = a.shape[:3]
    m1 = np.ascontiguousarray(map1)
    m2 = None if map2 is None else np.ascontiguousarray(map2)
    dh, dw = m1.shape[:2]
    if m1.dtype == np.float32 and m1.ndim == 2:
        mt = 5
    elif m1.dtype == np.float32:
        mt = 13
    elif m1.dtype == np.int16:
        mt = 11
    else:
        raise TypeError("map1 must be CV_32FC1, CV_32FC2 or CV_16SC2")
    dst = np.empty((n, dh, dw) + (() if a.ndim == 3 else (cn,)), a.dtype)
    if n == 0:
        return dst
    capi.check(capi.lib.b200_hal_remap_batch(_Img(a[0]).type, a.ctypes.data, a.strides[0], a.strides[1], w, h, dst.ctypes.data, dst.strides[0], dst.strides[1],
                                             dw, dh, n, mt, m1.ctypes.data, m1.strides[0], m2.ctypes.data if m2 is not None else None,
                                             m2.strides[0] if m2 is not None else 0, interpolation, borderMode, _bv(borderValue), chunk_frames),
               "remap_frames")
    return dst


def fisheye_initUndistortRectifyMap(K, D, R, P, size, m1type=11, device=None):
    """cv::fisheye::initUndistortRectifyMap (reference calib3d/src/fisheye.cpp:530-631) on the device: returns (map1, map2) as
    CUDA tensors, CV_16SC2 + CV_16UC1 (m1type 11, default) or two CV_32FC1 planes (m1type 5), ready for remap().
    K: 3x3 camera matrix, D: 4 coefficients (or None), R: 3x3 rectification (or None), P: 3x3 / 3x4 new camera matrix (or None).
    (P*R)^-1 is taken with numpy here; the reference uses its own SVD-based inverse (:575) -- pass the result of that instead
    of (R, P) as `R=None, P=inv` if the last ulp matters."""
    import torch
    w, h = size
    K = np.asarray(K, np.float64).reshape(3, 3)
    RR = np.eye(3) if R is None else np.asarray(R, np.float64).reshape(3, 3)
    PP = np.eye(3) if P is None else np.asarray(P, np.float64)[:, :3]
    iR = np.linalg.inv(PP @ RR)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if m1type <= 0:
        m1type = 11
    if m1type == 11:
        map1 = torch.empty((h, w, 2), dtype=torch.int16, device=dev)
        map2 = torch.empty((h, w), dtype=torch.int16, device=dev)        # CV_16UC1 bits
    elif m1type == 5:
        map1 = torch.empty((h, w), dtype=torch.float32, device=dev)
        map2 = torch.empty((h, w), dtype=torch.float32, device=dev)
    else:
        raise NotImplementedError("m1type must be CV_16SC2 (11) or CV_32FC1 (5)")
    Dc = (C.c_double * 4)(*[float(x) for x in np.asarray(D, np.float64).ravel()[:4]]) if D is not None else None
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        capi.check(capi.lib.b200_fisheye_initUndistortRectifyMap_dev((C.c_double * 9)(*K.ravel()), Dc, (C.c_double * 9)(*iR.ravel()), w, h, m1type,
                                                                     map1.data_ptr(), map1.stride(0) * map1.element_size(), map2.data_ptr(),
                                                                     map2.stride(0) * map2.element_size(), st), "fisheye_initUndistortRectifyMap")
    return map1, map2


def _undistort_params(newK, R, distCoeffs):
    """(iR, tilt, dist12) as ctypes arrays: the matrices of undistort.dispatch.cpp:123, :155-157 through b200_undistort_params"""
    d = np.zeros(14)
    if distCoeffs is not None:
        v = np.asarray(distCoeffs, np.float64).ravel()
        if v.size not in (4, 5, 8, 12, 14):
            raise ValueError("distCoeffs must have 4, 5, 8, 12 or 14 entries")
        d[:v.size] = v
    iR, tilt = (C.c_double * 9)(), (C.c_double * 9)()
    Rc = None if R is None else (C.c_double * 9)(*np.asarray(R, np.float64).reshape(9))
    capi.check(capi.lib.b200_undistort_params((C.c_double * 9)(*np.asarray(newK, np.float64)[:, :3].reshape(9)), Rc, float(d[12]), float(d[13]), iR, tilt),
               "undistort_params")
    return iR, tilt, (C.c_double * 12)(*d[:12])


def initUndistortRectifyMap(cameraMatrix, distCoeffs, R, newCameraMatrix, size, m1type=11, device=None):
    """cv::initUndistortRectifyMap (reference calib3d/src/undistort.dispatch.cpp:82-161) on the device: (map1, map2) as CUDA tensors,
    CV_16SC2 + CV_16UC1 (m1type 11), two CV_32FC1 planes (5) or one CV_32FC2 (13, map2 = None).  newCameraMatrix=None: the camera
    matrix with its principal point moved to the image centre (:108, getDefaultNewCameraMatrix(A, size, true))."""
    import torch
    w, h = size
    K = np.asarray(cameraMatrix, np.float64).reshape(3, 3)
    if newCameraMatrix is None:
        Ar = K.copy()
        Ar[0, 2], Ar[1, 2] = (w - 1) * 0.5, (h - 1) * 0.5
    else:
        Ar = np.asarray(newCameraMatrix, np.float64)
    iR, tilt, d12 = _undistort_params(Ar, R, distCoeffs)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if m1type <= 0:
        m1type = 11
    if m1type == 11:
        map1, map2 = torch.empty((h, w, 2), dtype=torch.int16, device=dev), torch.empty((h, w), dtype=torch.int16, device=dev)
    elif m1type == 5:
        map1, map2 = torch.empty((h, w), dtype=torch.float32, device=dev), torch.empty((h, w), dtype=torch.float32, device=dev)
    elif m1type == 13:
        map1, map2 = torch.empty((h, w, 2), dtype=torch.float32, device=dev), None
    else:
        raise NotImplementedError("m1type must be CV_16SC2 (11), CV_32FC1 (5) or CV_32FC2 (13)")
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        capi.check(capi.lib.b200_initUndistortRectifyMap_dev((C.c_double * 9)(*K.ravel()), d12, iR, tilt, w, h, m1type,
                                                             map1.data_ptr(), map1.stride(0) * map1.element_size(),
                                                             map2.data_ptr() if map2 is not None else None,
                                                             map2.stride(0) * map2.element_size() if map2 is not None else 0, st),
                   "initUndistortRectifyMap")
    return map1, map2


def undistort(src, cameraMatrix, distCoeffs, newCameraMatrix=None):
    """cv::undistort (reference calib3d/src/undistort.dispatch.cpp:284-328) for a CUDA tensor, stripe-wise maps as the reference
    builds them.  Returns (dst, (map1, map2)); the maps serve the following frames through remap()."""
    import torch
    s = _Img(src)
    if not s.device:
        raise TypeError("undistort works on CUDA tensors")
    K = np.asarray(cameraMatrix, np.float64).reshape(3, 3)
    _iR, tilt, d12 = _undistort_params(K, None, distCoeffs)
    newK = None if newCameraMatrix is None else (C.c_double * 9)(*np.asarray(newCameraMatrix, np.float64)[:, :3].reshape(9))
    dst = torch.empty_like(src)
    map1 = torch.empty((s.h, s.w, 2), dtype=torch.int16, device=src.device)
    map2 = torch.empty((s.h, s.w), dtype=torch.int16, device=src.device)
    with torch.cuda.device(src.device):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        capi.check(capi.lib.b200_undistort_dev(s.type, s.ptr, s.step, s.w, s.h, dst.data_ptr(), dst.stride(0) * dst.element_size(),
                                               (C.c_double * 9)(*K.ravel()), d12, newK, tilt, map1.data_ptr(), s.w * 4, map2.data_ptr(), s.w * 2, st),
                   "undistort")
    return dst, (map1, map2)


WARP_FILL_OUTLIERS, WARP_INVERSE_MAP, WARP_POLAR_LOG = 8, 16, 256


def warpPolar(src, dsize, center, maxRadius, flags, dst=None):
    """cv::warpPolar without WARP_INVERSE_MAP (reference imgproc/src/imgwarp.cpp:3163-3224) for a CUDA tensor: polar maps on the device,
    then remap with `flags & 7` and BORDER_CONSTANT (WARP_FILL_OUTLIERS) or BORDER_TRANSPARENT (dst, zeros unless given, keeps its
    contents where the map leaves the source).  dsize follows the reference's defaults (:3168-3176).  Returns (dst, (mapx, mapy))."""
    import torch
    s = _Img(src)
    if not s.device:
        raise TypeError("warpPolar works on CUDA tensors (host arrays: cv::warpPolar reaches the remap hook by itself)")
    if flags & WARP_INVERSE_MAP:
        raise NotImplementedError("WARP_INVERSE_MAP is left to the reference")
    w, h = dsize if dsize else (0, 0)
    if w <= 0 and h <= 0:
        w, h = _cv_round(maxRadius), _cv_round(maxRadius * np.pi)
    elif h <= 0:
        h = _cv_round(w * np.pi)
    if dst is None:
        dst = torch.zeros((h, w) + tuple(src.shape[2:]), dtype=src.dtype, device=src.device)
    d = _Img(dst)
    mapx = torch.empty((h, w), dtype=torch.float32, device=src.device)
    mapy = torch.empty((h, w), dtype=torch.float32, device=src.device)
    with _on_device(s):
        capi.check(capi.lib.b200_warpPolar_dev(s.type, s.ptr, s.step, s.w, s.h, d.ptr, d.step, d.w, d.h, float(center[0]), float(center[1]),
                                               float(maxRadius), int(flags), mapx.data_ptr(), w * 4, mapy.data_ptr(), w * 4, _stream(s)), "warpPolar")
    return dst, (mapx, mapy)


def _cv_round(v):
    return int(np.rint(v))                                  # cvRound: round half to even


def fisheye_undistortImage(distorted, K, D, Knew=None, new_size=None, maps=None):
    """cv::fisheye::undistortImage (reference calib3d/src/fisheye.cpp:636-646) for a CUDA tensor: builds the CV_16SC2 + CV_16UC1 maps on
    the device (or re-uses `maps` from an earlier call) and remaps with INTER_LINEAR / BORDER_CONSTANT.  Returns (undistorted, maps)."""
    s = _Img(distorted)
    if not s.device:
        raise TypeError("fisheye_undistortImage works on CUDA tensors (use remap() with host arrays)")
    w, h = (s.w, s.h) if not new_size else new_size
    if maps is None:
        maps = fisheye_initUndistortRectifyMap(K, D, None, np.eye(3) if Knew is None else Knew, (w, h), 11, device=distorted.device)
    return remap(distorted, maps[0], maps[1], interpolation=INTER_LINEAR, borderMode=BORDER_CONSTANT), maps


DNN_PMODE_NULL, DNN_PMODE_CROP_CENTER, DNN_PMODE_LETTERBOX = 0, 1, 2


def blobFromImages(images, scalefactor=1.0, size=None, mean=(0, 0, 0, 0), swapRB=False, crop=False, paddingmode=None, borderValue=(0, 0, 0, 0)):
    """cv::dnn::blobFromImages / blobFromImagesWithParams for a CUDA uint8 tensor (N, H, W[, C]) -> CUDA float32 tensor
    (N, C, size[1], size[0]): cv::resize INTER_LINEAR to `size` (skipped when equal), (v - mean[c]) * scalefactor[c], optional swapRB
    (reference dnn/src/dnn_utils.cpp:130-241, :275-353).  crop=True (blobFromImages, :96) or paddingmode=DNN_PMODE_CROP_CENTER resize
    by the larger ratio and crop the centre; DNN_PMODE_LETTERBOX resizes by the smaller one and pads with borderValue.  One fused kernel."""
    import torch
    if not _is_torch(images) or not images.is_cuda or images.dtype != torch.uint8 or not images.is_contiguous():
        raise ValueError("images must be a contiguous CUDA uint8 tensor (N, H, W[, C])")
    if images.dim() == 3:
        images = images.unsqueeze(-1)
    n, h, w, cn = images.shape
    bw, bh = (w, h) if not size or size[0] <= 0 else size
    out = torch.empty((n, cn, bh, bw), dtype=torch.float32, device=images.device)
    sf = (float(scalefactor),) * 4 if np.isscalar(scalefactor) else tuple(scalefactor) + (1.0,) * (4 - len(scalefactor))
    mn = tuple(mean) + (0.0,) * (4 - len(mean))
    with torch.cuda.device(images.device):
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        mode = paddingmode if paddingmode is not None else (DNN_PMODE_CROP_CENTER if crop else DNN_PMODE_NULL)
        bvv = tuple(borderValue) + (0.0,) * (4 - len(borderValue)) if not np.isscalar(borderValue) else (float(borderValue), 0.0, 0.0, 0.0)
        capi.check(capi.lib.b200_blobFromImagesWithParams_dev((cn - 1) << 3, images.data_ptr(), h * w * cn, w * cn, w, h, out.data_ptr(), bw, bh, n,
                                                              (C.c_double * 4)(*sf[:4]), (C.c_double * 4)(*[float(x) for x in mn[:4]]), int(bool(swapRB)),
                                                              int(mode), (C.c_double * 4)(*[float(x) for x in bvv[:4]]), st),
                   "blobFromImages")
    return out
