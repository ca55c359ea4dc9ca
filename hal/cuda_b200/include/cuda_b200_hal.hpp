// hal/cuda_b200 -- OpenCV HAL backend for the geometric-transform hot path on NVIDIA B200 (sm_100a).
//
// Registered like hal/ipp or hal/carotene: this header is pulled into modules/imgproc/src/hal_replacement.hpp
// through the generated custom_hal.hpp (hal_replacement.hpp:1596) and swaps the no-op hal_ni_* stubs
// (hal_replacement.hpp:369,387,428,483-487,506-509,530-534,1461,1508) for the C-ABI entry points of libb200geom.so.
// Every hook returns CV_HAL_ERROR_OK (handled on the GPU), CV_HAL_ERROR_NOT_IMPLEMENTED (OpenCV silently runs
// its CPU path: unsupported type / interpolation / border, images below the pixel gate, no CUDA device) or
// CV_HAL_ERROR_UNKNOWN (CUDA failure).  The header is self-contained: it needs only
// opencv2/core/hal/interface.h types, as the core and features2d TUs that also include custom_hal.hpp require.
#ifndef OPENCV_HAL_CUDA_B200_HPP
#define OPENCV_HAL_CUDA_B200_HPP

#include "opencv2/core/hal/interface.h"
#include "b200_geom.h"

#undef cv_hal_resize
#define cv_hal_resize b200_hal_resize
#undef cv_hal_warpAffine
#define cv_hal_warpAffine b200_hal_warpAffine
#undef cv_hal_warpPerspective
#define cv_hal_warpPerspective b200_hal_warpPerspective
#undef cv_hal_remap32f
#define cv_hal_remap32f b200_hal_remap32f
#undef cv_hal_remap32fc2
#define cv_hal_remap32fc2 b200_hal_remap32fc2
#undef cv_hal_remap16s
#define cv_hal_remap16s b200_hal_remap16s
#undef cv_hal_pyrdown
#define cv_hal_pyrdown b200_hal_pyrdown
#undef cv_hal_pyrup
#define cv_hal_pyrup b200_hal_pyrup
// cv_hal_warpAffineBlockline[NN] / cv_hal_warpPerspectiveBlockline[NN] are left to the CPU on purpose: they
// only run inside OpenCV's CPU fallback, per <=4096-pixel tile.

#endif
