// b200_resize_warpAffine_noalloc_dev: the resize -> warpAffine frame pipeline for callers that must not
// allocate.  Think of a real-time video loop that promised its integrator "no allocator traffic after
// start-up", a plugin running inside somebody else's memory budget, or code executing where cudaMalloc would
// serialise against other work on the device.  Those callers can hand us inputs and outputs but have nowhere
// to put the intermediate (resized) frame, so it lives in the fixed device array below.
//
// The array is part of the module image: every GPU context that loads libb200geom.so carries these
// B200_NOALLOC_SCRATCH_BYTES until the process exits, whether or not this entry point is ever called.
#include "../../include/b200_geom.h"
#include <cuda_runtime.h>
#include <cstdio>

// Intermediate frame storage.  In host code this name is only a handle for the runtime (not a device
// address); the address kernels may use is obtained per device with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

namespace {

// Asks the runtime where the current device's copy of the array lives.  Done on every call: the lookup allocates nothing,
// costs well under a microsecond, and a cached address would go stale if the application reset the device
// (cudaDeviceReset) between two calls.  Returns nullptr (and reports why) if that fails.
unsigned char* scratch_for_current_device()
{
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
        int dev = -1;
        cudaGetDevice(&dev);
        std::fprintf(stderr, "b200_resize_warpAffine_noalloc_dev: cudaGetSymbolAddress failed on device %d: %s\n",
                     dev, cudaGetErrorString(e));
        return nullptr;
    }
    return p;
}

int elem_bytes(int type)
{
    const int depth = type & 7, cn = ((type >> 3) & 511) + 1;
    const int e1 = depth <= 1 ? 1 : depth <= 3 ? 2 : depth <= 5 ? 4 : 8;
    return e1 * cn;
}

} // namespace

extern "C" __attribute__((visibility("default")))
int b200_resize_warpAffine_noalloc_dev(int src_type,
                                       const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                       int mid_width, int mid_height,
                                       void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                                       int n_frames, int resize_interpolation,
                                       const double M[6], int warp_interpolation, int borderType,
                                       const double borderValue[4], void* stream)
{
    if (n_frames <= 0) return B200_OK;
    if (mid_width <= 0 || mid_height <= 0) return B200_NOT_IMPLEMENTED;
    // rows of the intermediate frame are padded to 256 bytes so every fast path of the two stages applies
    const size_t mid_step = (((size_t)mid_width * elem_bytes(src_type)) + 255) & ~(size_t)255;
    const size_t need = mid_step * (size_t)mid_height;
    if (need > (size_t)B200_NOALLOC_SCRATCH_BYTES) {
        std::fprintf(stderr, "b200_resize_warpAffine_noalloc_dev: intermediate frame %dx%d needs %zu bytes, the built-in "
                             "scratch holds %zu; use b200_resize_warpAffine_batch_dev with your own buffer\n",
                     mid_width, mid_height, need, (size_t)B200_NOALLOC_SCRATCH_BYTES);
        return B200_NOALLOC_TOO_LARGE;
    }
    unsigned char* mid = scratch_for_current_device();
    if (!mid) return B200_ERROR;
    // Frames go through one at a time: resize frame i into the scratch, warp the scratch into dst frame i.
    // Stream order keeps a frame's warp ahead of the next frame's resize.  Nothing in the scratch has to start
    // at zero: the resize stage writes every pixel the warp stage can read (row padding is never read).
    for (int f = 0; f < n_frames; f++) {
        const unsigned char* s = (const unsigned char*)src + (size_t)f * src_frame_stride;
        unsigned char* d = (unsigned char*)dst + (size_t)f * dst_frame_stride;
        int rc = b200_resize_dev(src_type, s, src_step, src_width, src_height, mid, mid_step, mid_width, mid_height,
                                 0.0, 0.0, resize_interpolation, stream);
        if (rc != B200_OK) return rc;
        rc = b200_warpAffine_dev(src_type, mid, mid_step, mid_width, mid_height, d, dst_step, dst_width, dst_height,
                                 M, warp_interpolation, borderType, borderValue, stream);
        if (rc != B200_OK) return rc;
    }
    return B200_OK;
}
