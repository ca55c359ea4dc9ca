"""ctypes binding of libb200geom.so (include/b200_geom.h).  The library is REQUIRED: importing this
module raises if the CUDA extension has not been built -- there is no CPU or PyTorch fallback."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200geom.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: build it with `make` (or __graft_entry__.build()); "
                      "opencv_b200 has no fallback path")

lib = C.CDLL(LIB_PATH)

_vp, _sz, _i, _d, _ll = C.c_void_p, C.c_size_t, C.c_int, C.c_double, C.c_longlong
_IMG = [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i]

# every symbol include/b200_geom.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "b200_hal_resize": (_i, _IMG + [_d, _d, _i]),
    "b200_hal_warpAffine": (_i, _IMG + [_vp, _i, _i, _vp]),
    "b200_hal_warpPerspective": (_i, _IMG + [_vp, _i, _i, _vp]),
    "b200_hal_remap32f": (_i, _IMG + [_vp, _sz, _vp, _sz, _i, _i, _vp]),
    "b200_hal_remap32fc2": (_i, _IMG + [_vp, _sz, _i, _i, _vp]),
    "b200_hal_remap16s": (_i, _IMG + [_vp, _sz, _vp, _sz, _i, _i, _vp]),
    "b200_hal_set_min_pixels": (_ll, [_ll]),
    "b200_hal_set_policy": (_i, [_i]),
    "b200_resize_dev": (_i, _IMG + [_d, _d, _i, _vp]),
    "b200_warpAffine_dev": (_i, _IMG + [_vp, _i, _i, _vp, _vp]),
    "b200_warpPerspective_dev": (_i, _IMG + [_vp, _i, _i, _vp, _vp]),
    "b200_remap32f_dev": (_i, _IMG + [_vp, _sz, _vp, _sz, _i, _i, _vp, _vp]),
    "b200_remap32fc2_dev": (_i, _IMG + [_vp, _sz, _i, _i, _vp, _vp]),
    "b200_remap16s_dev": (_i, _IMG + [_vp, _sz, _vp, _sz, _i, _i, _vp, _vp]),
    "b200_resize_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _d, _d, _i, _vp]),
    "b200_warpAffine_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "b200_convertMaps_dev": (_i, [_i, _vp, _sz, _i, _vp, _sz, _i, _vp, _sz, _vp, _sz, _i, _i, _i, _vp]),
    "b200_warpPerspective_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "b200_remap_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _sz, _vp, _sz, _i, _i, _vp, _vp]),
    "b200_blobFromImages_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _i, _i, _i, _vp, _vp, _i, _vp]),
    "b200_blobFromImagesWithParams_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp]),
    "b200_sched_run_pipeline_host": (_i, [_vp, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "b200_hal_resize_batch": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, C.c_double, C.c_double, _i, _i]),
    "b200_hal_warpAffine_batch": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _vp, _i, _i, _vp, _i]),
    "b200_hal_warpPerspective_batch": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _vp, _i, _i, _vp, _i]),
    "b200_hal_remap_batch": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _sz, _vp, _sz, _i, _i, _vp, _i]),
    "b200_hal_resize_warpAffine_batch": (_i, [_i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _sz, _sz, _i, _i,
                                              _i, _i, _vp, _i, _i, _vp, _i]),
    "b200_resize_warpAffine_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i,
                                              _vp, _sz, _sz, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "b200_resize_warpAffine_noalloc_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "b200_fisheye_initUndistortRectifyMap_dev": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp, _sz, _vp]),
    "b200_stitch_buildMaps_dev": (_i, [_i, _vp, _vp, C.c_float, _i, _i, _i, _i, _vp, _sz, _vp, _sz, _vp]),
    "b200_stitch_warp_dev": (_i, [_i, _vp, _vp, C.c_float, _i, _i, _i, _vp, _sz, _i, _i, _vp, _sz, _i, _i, _vp, _sz, _vp, _sz, _i, _i, _vp]),
    "b200_warpPolar_buildMaps_dev": (_i, [_i, _i, C.c_float, C.c_float, _d, _i, _vp, _sz, _vp, _sz, _vp]),
    "b200_warpPolar_dev": (_i, [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i, C.c_float, C.c_float, _d, _i, _vp, _sz, _vp, _sz, _vp]),
    "b200_hal_pyrdown": (_i, [_vp, _sz, _i, _i, _vp, _sz, _i, _i, _i, _i, _i]),
    "b200_pyrDown_dev": (_i, [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i, _i, _vp]),
    "b200_pyrDown_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp]),
    "b200_hal_pyrup": (_i, [_vp, _sz, _i, _i, _vp, _sz, _i, _i, _i, _i, _i]),
    "b200_pyrUp_dev": (_i, [_i, _vp, _sz, _i, _i, _vp, _sz, _i, _i, _i, _vp]),
    "b200_pyrUp_batch_dev": (_i, [_i, _vp, _sz, _sz, _i, _i, _vp, _sz, _sz, _i, _i, _i, _i, _vp]),
    "b200_initUndistortRectifyMap_dev": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _sz, _vp, _sz, _vp]),
    "b200_undistort_params": (_i, [_vp, _vp, C.c_double, C.c_double, _vp, _vp]),
    "b200_undistort_dev": (_i, [_i, _vp, _sz, _i, _i, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _sz, _vp, _sz, _vp]),
    "b200_fisheye_undistortImage_dev": (_i, _IMG + [_vp, _vp, _vp, _vp, _sz, _vp, _sz, _vp]),
    "b200_synth_frames_dev": (_i, [_vp, _sz, _sz, _i, _i, _i, _i, _i, C.c_uint, _vp]),
    "b200_sched_run_pipeline": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp, C.c_uint, _vp, _vp]),
    "b200_sched_partition": (None, [_i, _i, _i, _vp, _vp]),
    "b200_last_error": (C.c_char_p, []),
    "b200_device_count": (_i, []),
    "b200_version": (C.c_char_p, []),
    "b200_launch_count": (C.c_ulonglong, []),
    "b200_set_force_generic": (_i, [_i]),
    "b200_profile_begin": (None, []),
    "b200_profile_end": (_i, [_vp, _vp, _i]),
}

for _name, (_res, _args) in SYMBOLS.items():
    _f = getattr(lib, _name)          # AttributeError here == header/library mismatch: fail loudly
    _f.restype = _res
    _f.argtypes = _args

# The Python mirror has no CPU path to fall back to, so the hooks it calls take every cell they support (the C default,
# "profitable only", is meant for an OpenCV that has its own CPU code behind the hook); OPENCV_B200_POLICY still wins.
if "OPENCV_B200_POLICY" not in os.environ:
    lib.b200_hal_set_policy(1)


def last_error():
    return lib.b200_last_error().decode()


def check(rc, what):
    if rc == 0:
        return
    if rc == 1:
        raise NotImplementedError(f"{what}: combination not implemented by the B200 backend "
                                  "(OpenCV would fall back to its CPU path)")
    raise RuntimeError(f"{what} failed: {last_error()}")
