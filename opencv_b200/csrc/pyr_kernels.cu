// cv::pyrDown (reference imgproc/src/pyramids.cpp:880-1036, 1266-1310): 5x5 binomial filter (1 4 6 4 1 per axis) + decimation by 2,
// CV_8U / CV_16U / CV_16S (int accumulators, (v + 128) >> 8) and CV_32F, 1-4 channels, every border the reference accepts.
//
// One CTA = 128 destination elements x 16 destination rows: the 35 source rows it needs are filtered horizontally (and decimated) into
// shared memory once, then filtered vertically, 4 neighbouring elements per thread and store.  The 1- and 2-byte depths copy the
// source span into shared memory as 32-bit words first; CV_32F reads its taps from global memory (L1).  DRAM sees the source once
// plus a 3-row halo per 32 rows.
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

// source elements a tile's interior destination pixels read per row: 2 * (128 / CN + 1) + 3 pixels, 273 elements at most (CN = 3);
// with up to 3 bytes of word alignment that is 69 words of CV_8U, 137 of the 16-bit depths
template <typename T> struct PyrRawWords { static constexpr int value = sizeof(T) == 1 ? 72 : 144; };

template <typename T, int CN>
__global__ void __launch_bounds__(PY_NT, 8) pyrdown_kernel(const PyrArgs a)
{
    typedef typename PyrWork<T>::type WT;
    constexpr bool FLT = std::is_same<WT, float>::value;
    // 1- and 2-byte depths: the source rows are copied into shared memory as aligned 32-bit words first (a byte-granular global load per
    // tap made the load/store unit the limit); CV_32F reads its taps straight from global memory
    constexpr bool STAGE = sizeof(T) <= 2;
    // CV_8U keeps the horizontally filtered values (<= 16 * 255) as 16-bit halves: the vertical pass then works on two elements per
    // register -- no lane of q0 + q4 + 6*q2 + 4*(q1 + q3) + 128 exceeds 65408
    constexpr bool PK = sizeof(T) == 1;
    typedef typename std::conditional<PK, uint16_t, WT>::type HT;
    constexpr int Q = PK ? 8 : 4, QPR = PY_TWE / Q;            // elements per vertical task, tasks per row
    __shared__ __align__(16) HT h[PY_ROWS][PY_TWE];
    __shared__ int rowoff[PY_ROWS];                            // border-interpolated source row of every filtered row
    constexpr int RAWW = PyrRawWords<T>::value;
    __shared__ uint32_t raw[STAGE ? PY_ROWS : 1][STAGE ? RAWW : 1];
    __shared__ int rsh[STAGE ? PY_ROWS : 1];                   // elements between a staged row's first word and its first element
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
    const bool inner = px >= 1 && px < a.width0;
    int t0, t1, t2, t3, t4;
    if (inner) { t0 = (2 * px - 2) * CN + c; t1 = t0 + CN; t2 = t1 + CN; t3 = t2 + CN; t4 = t3 + CN; }
    else {
        // tabL (pyramids.cpp:899) / tabR (:900, :1015-1021: indexed by (pixel - width0) + tap -- a second right-hand pixel, which
        // only a non-default dsize has, takes its taps ONE source pixel to the right of the first one's)
        const int b = px < 1 ? 2 * px - 2 : 2 * a.width0 - 2 + (px - a.width0);
        t0 = border_idx(b, a.sw, a.border) * CN + c; t1 = border_idx(b + 1, a.sw, a.border) * CN + c;
        t2 = border_idx(b + 2, a.sw, a.border) * CN + c; t3 = border_idx(b + 3, a.sw, a.border) * CN + c;
        t4 = border_idx(b + 4, a.sw, a.border) * CN + c;
    }
    const bool vec_h = inner && (x - CN) < a.simd_h;
    // staged span: the taps of the tile's interior pixels, source pixels 2*p_lo - 2 .. 2*p_hi + 2 inside the image
    const int e_lo = max(2 * (x0 / CN) - 2, 0) * CN;
    const int e_cnt = (min(2 * ((min(x0 + PY_TWE, n) - 1) / CN) + 2, a.sw - 1) + 1) * CN - e_lo;
    __syncthreads();
    if constexpr (STAGE) {
        for (int r = tid >> 5; r < nrows; r += PY_NT / 32) {                                   // a warp per row
            const uintptr_t first = reinterpret_cast<uintptr_t>(sf + (long long)rowoff[r] * a.src_step) + (uintptr_t)e_lo * sizeof(T);
            const uint32_t* w = reinterpret_cast<const uint32_t*>(first & ~(uintptr_t)3);
            const int sh = (int)(first & 3), words = (sh + e_cnt * (int)sizeof(T) + 3) >> 2;   // an aligned word never leaves its 32-byte sector
            if ((tid & 31) == 0) rsh[r] = sh / (int)sizeof(T);
#pragma unroll
            for (int k = tid & 31; k < RAWW; k += 32)
                if (k < words) raw[r][k] = __ldg(w + k);
        }
        __syncthreads();
    }
    // ---- horizontal pass: source rows 2*y0 - 2 .. 2*(y0 + rows), decimated columns
    if (live) {
        if (STAGE && inner) {
            const int o = t0 - e_lo;
#pragma unroll 3
            for (int r = r0; r < nrows; r += RSTEP) {
                const T* s = reinterpret_cast<const T*>(raw[r]) + rsh[r] + o;
                h[r][e] = (HT)((int)s[2 * CN] * 6 + ((int)s[CN] + (int)s[3 * CN]) * 4 + (int)s[0] + (int)s[4 * CN]);
            }
        } else {
#pragma unroll 3
            for (int r = r0; r < nrows; r += RSTEP) {
                const T* s = reinterpret_cast<const T*>(sf + (long long)rowoff[r] * a.src_step);
                if constexpr (FLT) {
                    const float s0 = __ldg(s + t0), s1 = __ldg(s + t1), s2 = __ldg(s + t2), s3 = __ldg(s + t3), s4 = __ldg(s + t4);
                    h[r][e] = vec_h ? __fadd_rn(__fmul_rn(s2, 6.f), __fadd_rn(__fmul_rn(__fadd_rn(s1, s3), 4.f), __fadd_rn(s0, s4)))
                                    : __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(s2, 6.f), __fmul_rn(__fadd_rn(s1, s3), 4.f)), s0), s4);
                } else {
                    h[r][e] = (HT)((int)__ldg(s + t2) * 6 + ((int)__ldg(s + t1) + (int)__ldg(s + t3)) * 4 + (int)__ldg(s + t0) + (int)__ldg(s + t4));
                }
            }
        }
    }
    __syncthreads();
    // ---- vertical pass: a task = Q neighbouring elements of one destination row
    for (int t = tid; t < rows * QPR; t += PY_NT) {
        const int ry = t / QPR, eq = (t - ry * QPR) * Q, xq = x0 + eq;
        if (xq >= n) continue;
        __align__(16) T out[Q];
        if constexpr (PK) {
            const uint4 v0 = *reinterpret_cast<const uint4*>(&h[2 * ry][eq]), v1 = *reinterpret_cast<const uint4*>(&h[2 * ry + 1][eq]),
                        v2 = *reinterpret_cast<const uint4*>(&h[2 * ry + 2][eq]), v3 = *reinterpret_cast<const uint4*>(&h[2 * ry + 3][eq]),
                        v4 = *reinterpret_cast<const uint4*>(&h[2 * ry + 4][eq]);
            auto f = [](uint32_t q0, uint32_t q1, uint32_t q2, uint32_t q3, uint32_t q4) {
                return ((q0 + q4 + q2 * 6u + ((q1 + q3) << 2) + 0x00800080u) >> 8) & 0x00ff00ffu;            // two (v + 128) >> 8
            };
            const uint32_t r0 = f(v0.x, v1.x, v2.x, v3.x, v4.x), r1 = f(v0.y, v1.y, v2.y, v3.y, v4.y),
                           r2 = f(v0.z, v1.z, v2.z, v3.z, v4.z), r3 = f(v0.w, v1.w, v2.w, v3.w, v4.w);
            const uint2 pk = make_uint2(__byte_perm(r0, r1, 0x6420), __byte_perm(r2, r3, 0x6420));
            *reinterpret_cast<uint2*>(out) = pk;
        } else {
#pragma unroll
        for (int k = 0; k < Q; k++) {
            const WT q0 = h[2 * ry][eq + k], q1 = h[2 * ry + 1][eq + k], q2 = h[2 * ry + 2][eq + k], q3 = h[2 * ry + 3][eq + k], q4 = h[2 * ry + 4][eq + k];
            if constexpr (FLT) {
                out[k] = xq + k < a.simd_v
                       ? __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fadd_rn(q1, q3), q2), 4.f), __fadd_rn(__fadd_rn(q0, q4), __fadd_rn(q2, q2))), 1.f / 256)
                       : __fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(q2, 6.f), __fmul_rn(__fadd_rn(q1, q3), 4.f)), q0), q4), 1.f / 256);
            } else {
                out[k] = (T)((q2 * 6 + (q1 + q3) * 4 + q0 + q4 + 128) >> 8);
            }
        }
        }
        T* d = reinterpret_cast<T*>(df + (long long)(y0 + ry) * a.dst_step) + xq;
        if (xq + Q <= n && (reinterpret_cast<uintptr_t>(d) % (Q * sizeof(T))) == 0) {
            if constexpr (Q * sizeof(T) == 4) *reinterpret_cast<uint32_t*>(d) = *reinterpret_cast<const uint32_t*>(out);
            else if constexpr (Q * sizeof(T) == 8) *reinterpret_cast<uint2*>(d) = *reinterpret_cast<const uint2*>(out);
            else *reinterpret_cast<uint4*>(d) = *reinterpret_cast<const uint4*>(out);
        } else {
#pragma unroll
            for (int k = 0; k < Q; k++) if (xq + k < n) d[k] = out[k];
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

// ------------------------------------------------------------------------------------------------------------------------------
// cv::pyrUp (reference imgproc/src/pyramids.cpp:1036-1155, 1374-1416): every source pixel becomes a 2x2 block filtered with the even
// (1 6 1) and odd (4 4) phases of the binomial kernel, / 64.  CV_8U / CV_16U / CV_16S (int, (v + 32) >> 6), CV_32F, CV_64F.
//
// One CTA = 256 destination elements x 16 source rows (32 destination rows; 8 / 16 for CV_64F): the 18 source rows are copied into
// shared memory (all loads in flight at once), expanded horizontally there (one thread per destination element column: its three taps
// and its formula are fixed), then every thread filters vertically 16 bytes' worth of neighbouring elements per task and stores them
// with one instruction.  DRAM sees the source once plus a 2-row halo per 16 rows, and the destination (4x the source) once.
namespace b200 {

struct PyrUpArgs {
    const uint8_t* src; int src_step; long long src_frame_stride; int sw, sh;
    uint8_t* dst; int dst_step; long long dst_frame_stride; int dw, dh;
    int simd_v;
};

namespace {

constexpr int PU_TWE = 256, PU_NT = 256, PU_RAW = 144;
template <typename T> struct PuRows { static constexpr int value = sizeof(T) == 8 ? 8 : 16; };         // source rows per CTA (shared memory <= 48 KB)

template <typename T> struct PyrUpWork { typedef int type; };
template <> struct PyrUpWork<float> { typedef float type; };
template <> struct PyrUpWork<double> { typedef double type; };

__device__ __forceinline__ int pu_mul(int a, int b) { return a * b; }
__device__ __forceinline__ int pu_add(int a, int b) { return a + b; }
__device__ __forceinline__ float pu_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float pu_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double pu_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double pu_add(double a, double b) { return __dadd_rn(a, b); }

template <typename T, typename WT> __device__ __forceinline__ T pu_cast(WT v)
{
    if constexpr (std::is_same<WT, int>::value) return (T)((v + 32) >> 6);                     // FixPtCast<T, 6>
    else return pu_mul(v, (WT)(1. / 64));                                                      // FltCast<T, 6>
}

template <typename T, int CN>
__global__ void __launch_bounds__(PU_NT) pyrup_kernel(const PyrUpArgs a)
{
    typedef typename PyrUpWork<T>::type WT;
    // CV_8U keeps the horizontally expanded values (<= 8 * 255) as 16-bit halves: the vertical pass works on two elements per register
    constexpr bool PK = sizeof(T) == 1;
    typedef typename std::conditional<PK, uint16_t, WT>::type HT;
    constexpr int Q = sizeof(T) == 8 ? 2 : PK ? 8 : 4, QPR = PU_TWE / Q;                       // elements per vertical task, tasks per row
    constexpr int SR = PuRows<T>::value, HR = SR + 2;
    __shared__ __align__(16) HT h[HR][PU_TWE];
    __shared__ T raw[HR][PU_RAW];                                                              // the source rows, only the span this tile reads
    const int n = a.dw * CN, x0 = blockIdx.x * PU_TWE, y0 = blockIdx.y * SR, tid = threadIdx.x, W = a.sw;
    const uint8_t* sf = a.src + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* df = a.dst + (long long)blockIdx.z * a.dst_frame_stride;
    const int rows = min(SR, a.sh - y0);
    // ---- source rows y0 - 1 .. y0 + rows (row 1 above the top, H - 1 below the bottom: pyramids.cpp:1075) into shared memory: every
    // global load of the CTA is in flight at once.  Span: the tile's source pixels plus one on either side, <= 35 + (128 / CN) pixels.
    const int i_lo = max((min(x0 / CN, 2 * W - 1) >> 1) - 1, 0);
    const int i_hi = min((min((min(x0 + PU_TWE, n) - 1) / CN, 2 * W - 1) >> 1) + 1, W - 1);
    const int e_lo = i_lo * CN, cnt = (i_hi + 1) * CN - e_lo;
    for (int r = tid >> 5; r < rows + 2; r += PU_NT / 32) {                                   // a warp per row
        const int sy = y0 - 1 + r, ry = sy < 0 ? min(1, a.sh - 1) : min(sy, a.sh - 1);
        const T* srow = reinterpret_cast<const T*>(sf + (long long)ry * a.src_step) + e_lo;
#pragma unroll
        for (int k = tid & 31; k < PU_RAW; k += 32)
            if (k < cnt) raw[r][k] = __ldg(srow + k);
    }
    __syncthreads();
    // ---- horizontal pass: one thread per destination element column (its three taps and its formula are fixed)
    {
        const int x = x0 + tid;
        if (x < n) {
            int p = x / CN; const int c = x - p * CN;
            p = min(p, 2 * W - 1);                                                             // a destination 2W + 1 wide repeats column 2W - 1
            const int i = p >> 1;
            const int oa = max(i - 1, 0) * CN + c - e_lo, ob = i * CN + c - e_lo, od = min(i + 1, W - 1) * CN + c - e_lo;
            // formula of this column (pyramids.cpp): 0 = b*8 (:1083, :1096), 1 = b*6 + d*2 (:1090), 2 = a + b*7 (:1095),
            // 3 = (a + b*6) + d (:1108), 4 = (b + d)*4 (:1091, :1109) -- evaluated without branches, in the reference's order
            const int kind = W == 1 ? 0 : !(p & 1) ? (i == 0 ? 1 : i == W - 1 ? 2 : 3) : (i == W - 1 ? 0 : 4);
            const bool k4 = kind == 4, has_a = kind == 2 || kind == 3, has_d = kind == 1 || kind == 3;
            const WT cb = (WT)(kind == 0 ? 8 : kind == 2 ? 7 : kind == 4 ? 4 : 6), cd = (WT)(kind == 1 ? 2 : 1);
#pragma unroll 6
            for (int r = 0; r < rows + 2; r++) {
                const WT va = (WT)raw[r][oa], vb = (WT)raw[r][ob], vd = (WT)raw[r][od];
                const WT t = pu_mul(k4 ? pu_add(vb, vd) : vb, cb);
                const WT w = has_a ? pu_add(va, t) : t;
                h[r][tid] = (HT)(has_d ? pu_add(w, pu_mul(vd, cd)) : w);
            }
        }
    }
    __syncthreads();
    // ---- vertical pass: destination rows 2*y0 .. 2*(y0 + rows) - 1; a task = Q neighbouring elements of an even / odd row pair
    for (int t = tid; t < rows * QPR; t += PU_NT) {
        const int ry = t / QPR, e = (t - ry * QPR) * Q, x = x0 + e, Y = 2 * (y0 + ry);
        if (x >= n) continue;
        const bool pair = Y + 1 < a.dh;                                                        // a destination 2H - 1 tall has no last odd row
        __align__(16) T ev[Q], od[Q];
        if constexpr (PK) {
            const uint4 v0 = *reinterpret_cast<const uint4*>(&h[ry][e]), v1 = *reinterpret_cast<const uint4*>(&h[ry + 1][e]),
                        v2 = *reinterpret_cast<const uint4*>(&h[ry + 2][e]);
            auto fe = [](uint32_t h0, uint32_t h1, uint32_t h2) { return ((h0 + h2 + h1 * 6u + 0x00200020u) >> 6) & 0x00ff00ffu; };   // two (v + 32) >> 6
            auto fo = [](uint32_t h1, uint32_t h2) { return ((h1 + h2 + 0x00080008u) >> 4) & 0x00ff00ffu; };                       // ((h1 + h2)*4 + 32) >> 6
            *reinterpret_cast<uint2*>(ev) = make_uint2(__byte_perm(fe(v0.x, v1.x, v2.x), fe(v0.y, v1.y, v2.y), 0x6420),
                                                       __byte_perm(fe(v0.z, v1.z, v2.z), fe(v0.w, v1.w, v2.w), 0x6420));
            *reinterpret_cast<uint2*>(od) = make_uint2(__byte_perm(fo(v1.x, v2.x), fo(v1.y, v2.y), 0x6420),
                                                       __byte_perm(fo(v1.z, v2.z), fo(v1.w, v2.w), 0x6420));
        } else {
#pragma unroll
        for (int k = 0; k < Q; k++) {
            const WT h0 = h[ry][e + k], h1 = h[ry + 1][e + k], h2 = h[ry + 2][e + k];
            ev[k] = pu_cast<T, WT>(pu_add(pu_add(h0, pu_mul(h1, (WT)6)), h2));
            // PyrUpVecV<float,float> (:708-726) writes the odd row as (1/16)*(h1 + h2), the scalar tail as ((h1 + h2)*4)*(1/64)
            if (std::is_same<T, float>::value && x + k < a.simd_v) od[k] = (T)pu_mul((WT)(1. / 16), pu_add(h1, h2));
            else od[k] = pu_cast<T, WT>(pu_mul(pu_add(h1, h2), (WT)4));
        }
        }
        auto put = [&](int row, const T* v) {
            T* d = reinterpret_cast<T*>(df + (long long)row * a.dst_step) + x;
            if (x + Q <= n && (reinterpret_cast<uintptr_t>(d) % (Q * sizeof(T))) == 0) {
                if constexpr (Q * sizeof(T) == 4) *reinterpret_cast<uint32_t*>(d) = *reinterpret_cast<const uint32_t*>(v);
                else if constexpr (Q * sizeof(T) == 8) *reinterpret_cast<uint2*>(d) = *reinterpret_cast<const uint2*>(v);
                else *reinterpret_cast<uint4*>(d) = *reinterpret_cast<const uint4*>(v);
            } else {
#pragma unroll
                for (int k = 0; k < Q; k++) if (x + k < n) d[k] = v[k];
            }
        };
        put(Y, ev);
        if (pair) put(Y + 1, od);
        if (Y == 2 * a.sh - 2 && a.dh > 2 * a.sh) put(2 * a.sh, ev);                           // a destination 2H + 1 tall repeats row 2H - 2 (:1144-1153)
    }
}

template <typename T>
int launch_up_t(const PyrUpArgs& a, int cn, int n_frames, cudaStream_t s)
{
    dim3 grid((a.dw * cn + PU_TWE - 1) / PU_TWE, (a.sh + PuRows<T>::value - 1) / PuRows<T>::value, n_frames);
    switch (cn) {
    case 1: pyrup_kernel<T, 1><<<grid, PU_NT, 0, s>>>(a); break;
    case 2: pyrup_kernel<T, 2><<<grid, PU_NT, 0, s>>>(a); break;
    case 3: pyrup_kernel<T, 3><<<grid, PU_NT, 0, s>>>(a); break;
    case 4: pyrup_kernel<T, 4><<<grid, PU_NT, 0, s>>>(a); break;
    default: return B200_NOT_IMPLEMENTED;
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return report_cuda_error("pyrUp kernel launch", e);
    count_launch();
    return B200_OK;
}

} // namespace

// argument rules of cv::pyrUp / pyrUp_ (pyramids.cpp:1378, :1055-1056); B200_NOT_IMPLEMENTED = the caller's CPU path decides
int pyrup_check(int type, int sw, int sh, int dw, int dh, int border)
{
    const int depth = B200_DEPTH(type), cn = B200_CN(type);
    if (!(depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F || depth == D_64F) || cn < 1 || cn > 4) return B200_NOT_IMPLEMENTED;
    if (border != B_REFLECT_101) return B200_NOT_IMPLEMENTED;                                  // BORDER_DEFAULT only (:1378)
    if (sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0 || sw > 32767 || sh > 32767) return B200_NOT_IMPLEMENTED;
    if (abs(dw - sw * 2) != dw % 2 || abs(dh - sh * 2) != dh % 2) return B200_NOT_IMPLEMENTED;
    if (sw == 1 && dw == 3) return B200_NOT_IMPLEMENTED;                                       // the reference reads an unwritten cell here
    return B200_OK;
}

int launch_pyrup(int type, const void* src, size_t src_step, long long sfs, int sw, int sh,
                 void* dst, size_t dst_step, long long dfs, int dw, int dh, int n_frames, int border, cudaStream_t s)
{
    int rc = pyrup_check(type, sw, sh, dw, dh, border);
    if (rc != B200_OK) return rc;
    if (n_frames <= 0) return B200_OK;
    if (!src || !dst || src_step > 0x7fffffffu || dst_step > 0x7fffffffu || n_frames > 65535) return B200_NOT_IMPLEMENTED;
    const int depth = B200_DEPTH(type), cn = B200_CN(type), e1 = depth == D_8U ? 1 : depth == D_32F ? 4 : depth == D_64F ? 8 : 2;
    if (((uintptr_t)src % e1) || (src_step % e1) || ((uintptr_t)dst % e1) || (dst_step % e1) || (sfs % e1) || (dfs % e1)) return B200_NOT_IMPLEMENTED;
    PyrUpArgs a;
    a.src = (const uint8_t*)src; a.src_step = (int)src_step; a.src_frame_stride = sfs; a.sw = sw; a.sh = sh;
    a.dst = (uint8_t*)dst; a.dst_step = (int)dst_step; a.dst_frame_stride = dfs; a.dw = dw; a.dh = dh;
    a.simd_v = depth == D_32F ? (dw * cn / 4) * 4 : 0;
    switch (depth) {
    case D_8U: return launch_up_t<uint8_t>(a, cn, n_frames, s);
    case D_16U: return launch_up_t<uint16_t>(a, cn, n_frames, s);
    case D_16S: return launch_up_t<int16_t>(a, cn, n_frames, s);
    case D_32F: return launch_up_t<float>(a, cn, n_frames, s);
    default: return launch_up_t<double>(a, cn, n_frames, s);
    }
}

} // namespace b200

extern "C" __attribute__((visibility("default")))
int b200_pyrUp_dev(int type, const void* src, size_t src_step, int src_width, int src_height,
                   void* dst, size_t dst_step, int dst_width, int dst_height, int border, void* stream)
{
    B200_TRACE();
    return b200::launch_pyrup(type, src, src_step, 0, src_width, src_height, dst, dst_step, 0, dst_width, dst_height, 1, border, (cudaStream_t)stream);
}

extern "C" __attribute__((visibility("default")))
int b200_pyrUp_batch_dev(int type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                         void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames, int border, void* stream)
{
    B200_TRACE();
    return b200::launch_pyrup(type, src, src_step, (long long)src_frame_stride, src_width, src_height, dst, dst_step, (long long)dst_frame_stride,
                              dst_width, dst_height, n_frames, border, (cudaStream_t)stream);
}
