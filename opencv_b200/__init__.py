"""opencv_b200 -- B200-native (sm_100a) backend for OpenCV's geometric-transform hot path.

Python mirror of the reference's public API for this path, so tests read like the reference's own:

    warpAffine(src, M, dsize, flags=INTER_LINEAR, borderMode=BORDER_CONSTANT, borderValue=0)
    warpPerspective(src, M, dsize, ...)           remap(src, map1, map2, interpolation, ...)
    resize(src, dsize, fx=0, fy=0, interpolation=INTER_LINEAR)

`src` may be a numpy array (host path: goes through the b200_hal_* hooks, i.e. exactly what OpenCV's
cv_hal_* replacement would call, H2D/D2H inside) or a CUDA torch tensor in HWC layout (device-resident
path, asynchronous on torch's current stream).  All arithmetic runs in libb200geom.so; there is no
fallback.  Reference: modules/imgproc/src/imgwarp.cpp:1519,2367,2829; resize.cpp:4007.
"""
from .api import (BORDER_CONSTANT, BORDER_REFLECT, BORDER_REFLECT_101, BORDER_REPLICATE, BORDER_WRAP,  # noqa: F401
                  INTER_AREA, INTER_CUBIC, INTER_LANCZOS4, INTER_LINEAR, INTER_LINEAR_EXACT, INTER_NEAREST, INTER_NEAREST_EXACT, WARP_INVERSE_MAP, WARP_RELATIVE_MAP,
                  blobFromImages, fisheye_initUndistortRectifyMap, fisheye_undistortImage, initUndistortRectifyMap, invert_affine, buildPyramid, pyrDown, pyrUp, warpPolar, remap, remap_frames, resize, resize_warpAffine_frames, warpPerspective_frames, undistort, warpAffine, warpPerspective)
from . import capi  # noqa: F401

__all__ = ["warpAffine", "warpPerspective", "remap", "resize", "invert_affine", "blobFromImages", "resize_warpAffine_frames", "warpPerspective_frames", "remap_frames", "fisheye_initUndistortRectifyMap", "fisheye_undistortImage", "initUndistortRectifyMap", "undistort", "pyrDown", "pyrUp", "warpPolar", "buildPyramid", "capi"]
