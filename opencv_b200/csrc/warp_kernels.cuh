// Argument block + launcher declarations for the warpAffine / warpPerspective / remap kernel family.
#pragma once
#include "geom_common.cuh"

namespace b200 {

enum CoordKind { COORD_AFFINE = 0, COORD_PERSP = 1, COORD_MAP32F = 2, COORD_MAP32FC2 = 3, COORD_MAP16S = 4 };

struct WarpArgs {
    Image src;
    ImageOut dst;
    long long src_frame_stride, dst_frame_stride;   // bytes between frames of a batch (0 for one frame)
    int n_frames, frames_per_thread;                // generic kernel: a thread derives its coordinate once and samples this many frames
    double M[9];                                    // inverse (dst->src) matrix; affine uses M[0..5]
    int border;
    int linear;                                     // 0 = INTER_NEAREST, 1 = any interpolating mode (5 fractional bits)
    int interp;                                     // INTER_NEAREST 0, LINEAR 1, CUBIC 2, LANCZOS4 4
    const float* tab1d;                             // CUBIC / LANCZOS4: 32 x ksize 1-D coefficients (device)
    const short* tab2d_i;                           // CUBIC / LANCZOS4, 8U: 1024 x ksize^2 int16 weights (device)
    int relative;                                   // WARP_RELATIVE_MAP (remap only)
    int force_generic;                              // testing/diagnostics: skip the fast paths
    int keep_in_l2;                    // 1: plain stores (a consumer re-reads dst from L2 soon); 0: evict-first (.cs) stores
    int vec4;                                       // 4-byte pixels with 4-byte aligned src/dst/cval: word loads and stores
    int bw0_shift;                                  // log2(bw0) when bw0 is a power of two, else -1
    int bw0;                                        // warpPerspective coordinate block width (imgwarp.cpp:2695-2697)
    const uint8_t* map1; int map1_step;
    const uint8_t* map2; int map2_step;
    alignas(16) uint8_t cval[32];                   // border pixel, saturate_cast<T>(borderValue[k&3]) for k < cn (4 x 64F at most)
};

// returns cudaError_t as int; `depth`/`cn` select the instantiation, coord the coordinate generator
int launch_warp(const WarpArgs& a, int depth, int cn, int coord, int n_frames, cudaStream_t stream);

} // namespace b200
