// Hot-cell kernels (CV_8U): TMA-staged warpAffine and the vectorised 2x2 box.  See fast_kernels.cu.
#pragma once
#include "warp_kernels.cuh"
#include "resize_kernels.cuh"
#include <cmath>
#include <cstring>

namespace b200 {

// 0 = launched, -1 = not applicable (use the generic kernel), >0 = cudaError_t
int try_launch_warp_affine_tma_u8(const WarpArgs& a, int cn, int n_frames, cudaStream_t stream);

// warpPerspective 8U INTER_LINEAR, TMA-staged; same return convention
int try_launch_warp_persp_tma_u8(const WarpArgs& a, int cn, int n_frames, cudaStream_t stream);

// resize (exact 2x2 box) -> warpAffine over a batch in ONE persistent launch; mid_bytes = size of the scratch behind ra.dst
int try_launch_pipe_box_warp_u8(const ResizeArgs& ra, const WarpArgs& wa, int cn, int n_frames, size_t mid_bytes, cudaStream_t stream);

bool box2x2_u8_applicable(const ResizeArgs& a, int cn);
int launch_box2x2_u8(const ResizeArgs& a, int cn, int n_frames, cudaStream_t stream);

// integer-scale INTER_AREA (kx in 2..4, any ky, not 2x2) for 8U with word loads
bool area_int_u8_applicable(const ResizeArgs& a, int cn);
int launch_area_int_u8(const ResizeArgs& a, int cn, int n_frames, cudaStream_t stream);

} // namespace b200
