// Argument block + launcher declarations for the resize kernel family.
#pragma once
#include "geom_common.cuh"

namespace b200 {

enum ResizeMode { RS_NEAREST = 0, RS_LINEAR = 1, RS_AREA_FAST = 2, RS_AREA = 3, RS_NEAREST_EXACT = 4, RS_LINEAR_EXACT = 5, RS_CUBIC = 6, RS_LANCZOS4 = 7 };

struct ResizeArgs {
    Image src;
    ImageOut dst;
    long long src_frame_stride, dst_frame_stride;
    double inv_scale_x, inv_scale_y;   // as received by the hook (resize.cpp:3642-3646)
    double scale_x, scale_y;           // 1./inv_scale  (host fp64 division, resize.cpp:3764)
    double ifxd, ifyd;                 // nearest: 1./fx, 1./fy (resize.cpp:1128) -- same values as scale_*
    int cn;                            // channels (mode selection needs it: LINEAR_EXACT reroutes 2x unless cn == 2)
    int iscale_x, iscale_y;            // integer scales of the area-fast path (resize.cpp:3766-3767)
    int ifx, ifx0, ify, ify0;          // INTER_NEAREST_EXACT 16.16 steps and offsets (resize.cpp:1270-1273)
    int area_mode;                     // RS_LINEAR reached through INTER_AREA (resize.cpp:3877,3905-3916)
    int lin_separable;                 // RS_LINEAR: vertical scale < 2, horizontal results are shared between destination rows
    int force_generic;                 // testing/diagnostics: skip the fast paths
    int keep_in_l2;                    // 1: plain stores (a consumer re-reads dst from L2 soon); 0: evict-first (.cs) stores
    int kmax_x, kmax_y;                // RS_AREA: upper bound of table entries per dst column / row (shared-memory sizing)
    int nmax_x;                        // RS_AREA: the largest entry count any dst column really has (kernel tap bound)
    const float* ltab_x; const float* ltab_y;   // RS_LANCZOS4: 8 coefficients per dst column / row, built on the host (device memory)
};

// picks the reference's mode for (interpolation, scales) exactly as resize.cpp:3771-3873 does;
// returns the ResizeMode or -1 for unsupported interpolation.
int resize_select_mode(ResizeArgs& a, int interpolation);

int launch_resize(const ResizeArgs& a, int mode, int depth, int cn, int n_frames, cudaStream_t stream);

// dnn::blobFromImages (dnn/src/dnn_utils.cpp:130-241, :275-353) for CV_8U frames: cv::resize INTER_LINEAR to the blob size
// (or no resize when the sizes agree), then planar fp32 NCHW with (v - mean[ch]) * scale[ch], optional swapRB.
struct BlobArgs {
    float* blob;                       // [n][cn][dh][dw]
    float mean[4], scale[4];           // per BLOB channel (after the swap), as the reference applies them
    int swap_rb;
    // padding modes (dnn_utils.cpp:309-331): the blob (bw x bh) is a window of the resized image (ResizeArgs::dst.w x dst.h) placed
    // at (ox, oy) in it -- inside it for the centre crop, around it for the letterbox, whose outside is `pad` (per image channel)
    int bw, bh, ox, oy;
    uint8_t pad[4];
};
// mode: RS_LINEAR, RS_AREA_FAST (exact 2x2 only) or -2 for equal sizes; returns cudaError_t as int
int launch_blob_u8(const ResizeArgs& a, const BlobArgs& b, int mode, int cn, int n_frames, cudaStream_t stream);

} // namespace b200
