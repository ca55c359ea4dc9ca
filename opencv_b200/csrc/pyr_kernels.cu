// cv::pyrDown (reference imgproc/src/pyramids.cpp:880-1036, 1266-1310): 5x5 binomial filter (1 4 6 4 1 per axis) + decimation by 2,
// CV_8U / CV_16U / CV_16S (int accumulators, (v + 128) >> 8) and CV_32F, 1-4 channels, every border the reference accepts.
//
// One CTA = 128 destination elements x 16 destination rows: the 35 source rows it needs are filtered horizontally (and decimated) into
// shared memory once, then filtered vertically.  A source element is read from global memory 2.5 times on average, all within the
// CTA's rows (L1), so DRAM sees the source once plus a 3-row halo per 32 rows.
//
// CV_32F reproduces the reference's TWO evaluation orders per pass: its vector bodies (4 float lanes in the default x86 build) and its
// scalar borders / tails associate the five terms differently (pyramids.cpp:322-404, :494-514 against :955-1034), and which element
// gets which depends on its position in the row.  The rule is restated in oracle/geom_oracle.c (orc_pyrDown), pinned to the reference.
#include "../../include/b200_geom.h"
#include "geom_common.cuh"
#include "trace.cuh"
#include <type_traits>
#include <cstdlib>

namespace b200 {

void count_launch();
int report_cuda_error(const char* what, cudaError_t e);

struct PyrArgs {
    const uint8_t* src; int src_step; long long src_frame_stride; int sw, sh;
    uint8_t* dst; int dst_step; long long dst_frame_stride; int dw, dh;
    int border, width0, simd_h, simd_v;
};

namespace {

constexpr int PY_TWE = 128, PY_TH = 16, PY_NT = 256, PY_ROWS = 2 * PY_TH + 3;

template <typename T> struct PyrWork { typedef int type; };
template <> struct PyrWork<float> { typedef float type; };

template <typename T, int CN>
__global__ void __launch_bounds__(PY_NT) pyrdown_kernel(const PyrArgs a)
{
    typedef typename PyrWork<T>::type WT;
    constexpr bool FLT = std::is_same<WT, float>::value;
    __shared__ WT h[PY_ROWS][PY_TWE];
    __shared__ int rowoff[PY_ROWS];                            // border-interpolated source row of every filtered row
    const int n = a.dw * CN, x0 = blockIdx.x * PY_TWE, y0 = blockIdx.y * PY_TH, tid = threadIdx.x;
    const uint8_t* sf = a.src + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* df = a.dst + (long long)blockIdx.z * a.dst_frame_stride;
    const int rows = min(PY_TH, a.dh - y0), nrows = 2 * rows + 3;
    if (tid < nrows) rowoff[tid] = border_idx(2 * y0 - 2 + tid, a.sh, a.border);
    // a thread owns ONE element column of the tile (its tap offsets and evaluation order are fixed) and every second row
    const int e = tid % PY_TWE, r0 = tid / PY_TWE, x = x0 + e;
    constexpr int RSTEP = PY_NT / PY_TWE;
    const bool live = x < n;
    const int px = x / CN, c = x - px * CN;
    int t0, t1, t2, t3, t4;
    if (px >= 1 && px < a.width0) { t0 = (2 * px - 2) * CN + c; t1 = t0 + CN; t2 = t1 + CN; t3 = t2 + CN; t4 = t3 + CN; }
    else {
        // tabL (pyramids.cpp:899) / tabR (:900, :1015-1021: indexed by (pixel - width0) + tap -- a second right-hand pixel, which
        // only a non-default dsize has, takes its taps ONE source pixel to the right of the first one's)
        const int b = px < 1 ? 2 * px - 2 : 2 * a.width0 - 2 + (px - a.width0);
        t0 = border_idx(b, a.sw, a.border) * CN + c; t1 = border_idx(b + 1, a.sw, a.border) * CN + c;
        t2 = border_idx(b + 2, a.sw, a.border) * CN + c; t3 = border_idx(b + 3, a.sw, a.border) * CN + c;
        t4 = border_idx(b + 4, a.sw, a.border) * CN + c;
    }
    const bool vec_h = px >= 1 && px < a.width0 && (x - CN) < a.simd_h, vec_v = x < a.simd_v;
    __syncthreads();
    // ---- horizontal pass: source rows 2*y0 - 2 .. 2*(y0 + rows), decimated columns
    if (live)
#pragma unroll 3
        for (int r = r0; r < nrows; r += RSTEP) {
            const T* s = reinterpret_cast<const T*>(sf + (long long)rowoff[r] * a.src_step);
            if constexpr (FLT) {
                const float s0 = __ldg(s + t0), s1 = __ldg(s + t1), s2 = __ldg(s + t2), s3 = __ldg(s + t3), s4 = __ldg(s + t4);
                h[r][e] = vec_h ? __fadd_rn(__fmul_rn(s2, 6.f), __fadd_rn(__fmul_rn(__fadd_rn(s1, s3), 4.f), __fadd_rn(s0, s4)))
                                : __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(s2, 6.f), __fmul_rn(__fadd_rn(s1, s3), 4.f)), s0), s4);
            } else {
                h[r][e] = (int)__ldg(s + t2) * 6 + ((int)__ldg(s + t1) + (int)__ldg(s + t3)) * 4 + (int)__ldg(s + t0) + (int)__ldg(s + t4);
            }
        }
    __syncthreads();
    // ---- vertical pass
    if (live)
#pragma unroll 2
        for (int ry = r0; ry < rows; ry += RSTEP) {
            const WT q0 = h[2 * ry][e], q1 = h[2 * ry + 1][e], q2 = h[2 * ry + 2][e], q3 = h[2 * ry + 3][e], q4 = h[2 * ry + 4][e];
            T* d = reinterpret_cast<T*>(df + (long long)(y0 + ry) * a.dst_step);
            if constexpr (FLT) {
                d[x] = vec_v
                     ? __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fadd_rn(q1, q3), q2), 4.f), __fadd_rn(__fadd_rn(q0, q4), __fadd_rn(q2, q2))), 1.f / 256)
                     : __fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q2, 6.f), __fmul_rn(__fadd_rn(q1, q3), 4.f)), q0), q4), 1.f / 256);
            } else {
                d[x] = (T)((q2 * 6 + (q1 + q3) * 4 + q0 + q4 + 128) >> 8);
            }
        }
}

template <typename T>
int launch_t(const PyrArgs& a, int cn, int n_frames, cudaStream_t s)
{
    dim3 grid((a.dw * cn + PY_TWE - 1) / PY_TWE, (a.dh + PY_TH - 1) / PY_TH, n_frames);
    switch (cn) {
    case 1: pyrdown_kernel<T, 1><<<grid, PY_NT, 0, s>>>(a); break;
    case 2: pyrdown_kernel<T, 2><<<grid, PY_NT, 0, s>>>(a); break;
    case 3: pyrdown_kernel<T, 3><<<grid, PY_NT, 0, s>>>(a); break;
    case 4: pyrdown_kernel<T, 4><<<grid, PY_NT, 0, s>>>(a); break;
    default: return B200_NOT_IMPLEMENTED;
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return report_cuda_error("pyrDown kernel launch", e);
    count_launch();
    return B200_OK;
}

} // namespace

// argument rules of cv::pyrDown / pyrDown_ (pyramids.cpp:1270, :892-894); B200_NOT_IMPLEMENTED = the caller's CPU path decides
int pyrdown_check(int type, int sw, int sh, int dw, int dh, int border)
{
    const int depth = B200_DEPTH(type), cn = B200_CN(type);
    if (!(depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F) || cn < 1 || cn > 4) return B200_NOT_IMPLEMENTED;
    border &= ~16;                                         // BORDER_ISOLATED
    if (!(border == B_REPLICATE || border == B_REFLECT || border == B_WRAP || border == B_REFLECT_101)) return B200_NOT_IMPLEMENTED;
    if (sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0 || sw > 32767 * 2 || sh > 32767 * 2) return B200_NOT_IMPLEMENTED;
    if (abs(dw * 2 - sw) > 2 || abs(dh * 2 - sh) > 2) return B200_NOT_IMPLEMENTED;
    return B200_OK;
}

int launch_pyrdown(int type, const void* src, size_t src_step, long long sfs, int sw, int sh,
                   void* dst, size_t dst_step, long long dfs, int dw, int dh, int n_frames, int border, cudaStream_t s)
{
    int rc = pyrdown_check(type, sw, sh, dw, dh, border);
    if (rc != B200_OK) return rc;
    if (n_frames <= 0) return B200_OK;
    if (!src || !dst || src_step > 0x7fffffffu || dst_step > 0x7fffffffu || n_frames > 65535) return B200_NOT_IMPLEMENTED;
    const int depth = B200_DEPTH(type), cn = B200_CN(type), e1 = depth == D_8U ? 1 : depth == D_32F ? 4 : 2;
    if (((uintptr_t)src % e1) || (src_step % e1) || ((uintptr_t)dst % e1) || (dst_step % e1) || (sfs % e1) || (dfs % e1)) return B200_NOT_IMPLEMENTED;
    PyrArgs a;
    a.src = (const uint8_t*)src; a.src_step = (int)src_step; a.src_frame_stride = sfs; a.sw = sw; a.sh = sh;
    a.dst = (uint8_t*)dst; a.dst_step = (int)dst_step; a.dst_frame_stride = dfs; a.dw = dw; a.dh = dh;
    a.border = border & ~16;
    a.width0 = (sw - 3) / 2 + 1; if (a.width0 > dw) a.width0 = dw;                 // pyramids.cpp:895
    // elements handled by the vector bodies of the reference's float code (PyrDownVecH<float,float,cn>, PyrDownVecV<float,float>)
    int W = a.width0 * cn - cn; if (W < 0) W = 0;
    a.simd_h = (cn == 1 || cn == 4) ? (W / 4) * 4 : cn == 2 ? (W / 8) * 8 : (W >= 4 ? ((W - 4) / 3 + 1) * 3 : 0);
    a.simd_v = (dw * cn / 4) * 4;
    switch (depth) {
    case D_8U: return launch_t<uint8_t>(a, cn, n_frames, s);
    case D_16U: return launch_t<uint16_t>(a, cn, n_frames, s);
    case D_16S: return launch_t<int16_t>(a, cn, n_frames, s);
    default: return launch_t<float>(a, cn, n_frames, s);
    }
}

} // namespace b200

extern "C" __attribute__((visibility("default")))
int b200_pyrDown_dev(int type, const void* src, size_t src_step, int src_width, int src_height,
                     void* dst, size_t dst_step, int dst_width, int dst_height, int border, void* stream)
{
    B200_TRACE();
    return b200::launch_pyrdown(type, src, src_step, 0, src_width, src_height, dst, dst_step, 0, dst_width, dst_height, 1, border, (cudaStream_t)stream);
}

extern "C" __attribute__((visibility("default")))
int b200_pyrDown_batch_dev(int type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                           void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames, int border, void* stream)
{
    B200_TRACE();
    return b200::launch_pyrdown(type, src, src_step, (long long)src_frame_stride, src_width, src_height, dst, dst_step, (long long)dst_frame_stride,
                                dst_width, dst_height, n_frames, border, (cudaStream_t)stream);
}
