// Device-side arithmetic shared by every kernel: x86-exact rounding, saturation, border indexing.
// Reference semantics being reproduced (paths under /root/reference/modules):
//   cvRound / cvFloor / cvCeil   core/include/opencv2/core/fast_math.hpp:199-262,330-386
//   saturate_cast                core/include/opencv2/core/saturate.hpp:101-151
//   borderInterpolate            core/src/copy.cpp:905-950
// This TU family is compiled with -fmad=false: the reference's default build has no FMA, so every
// multiply and add must round separately for the 16U/32F results to be bit-identical.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <limits.h>

namespace b200 {

enum { D_8U = 0, D_8S = 1, D_16U = 2, D_16S = 3, D_32S = 4, D_32F = 5, D_64F = 6 };
enum { I_NEAREST = 0, I_LINEAR = 1, I_CUBIC = 2, I_AREA = 3, I_LANCZOS4 = 4, I_LINEAR_EXACT = 5, I_NEAREST_EXACT = 6 };
enum { B_CONSTANT = 0, B_REPLICATE = 1, B_REFLECT = 2, B_WRAP = 3, B_REFLECT_101 = 4, B_TRANSPARENT = 5 };
enum { WARP_RELATIVE_MAP = 32 };
enum { INTER_BITS = 5, INTER_TAB_SIZE = 32 };

#define B200_DEPTH(t) ((t) & 7)
#define B200_CN(t) ((((t) >> 3) & 511) + 1)

// cvRound(double) == cvtsd2si: nearest-even; 0x80000000 when the rounded value does not fit or is NaN.
// __double2int_rn saturates instead (and maps NaN to 0): only the positive side and NaN differ.
__device__ __forceinline__ int rne_d(double v) { return (v < 2147483647.5) ? __double2int_rn(v) : INT_MIN; }
// cvRound(float) == cvtss2si
__device__ __forceinline__ int rne_f(float v) { return (v < 2147483648.0f) ? __float2int_rn(v) : INT_MIN; }
// cvFloor: i = (int)v; i - (i > v)   (in-range values only reach these)
__device__ __forceinline__ int floor_d(double v) { return __double2int_rd(v); }
__device__ __forceinline__ int floor_f(float v) { return __float2int_rd(v); }
__device__ __forceinline__ int ceil_d(double v) { return __double2int_ru(v); }

__device__ __forceinline__ int sat_s16(int v) { return max(-32768, min(32767, v)); }
__device__ __forceinline__ int sat_u8(int v) { return max(0, min(255, v)); }
__device__ __forceinline__ int sat_u16(int v) { return max(0, min(65535, v)); }
__device__ __forceinline__ int clipi(int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; }

// borderInterpolate for the out-of-range case (copy.cpp:905-950); returns -1 for BORDER_CONSTANT.
__device__ __forceinline__ int border_idx(int p, int len, int borderType)
{
    if ((unsigned)p < (unsigned)len) return p;
    if (borderType == B_REPLICATE) return p < 0 ? 0 : len - 1;
    if (borderType == B_REFLECT || borderType == B_REFLECT_101) {
        const int delta = borderType == B_REFLECT_101;
        if (len == 1) return 0;
        do {
            if (p < 0) p = -p - 1 + delta;
            else p = len - 1 - (p - len) - delta;
        } while ((unsigned)p >= (unsigned)len);
        return p;
    }
    if (borderType == B_WRAP) {
        if (p < 0) p -= ((p - len + 1) / len) * len;
        if (p >= len) p %= len;
        return p;
    }
    return -1;
}

// integer -> float without the XU pipe: I2F runs at 16 lanes/clk/SM and was the limiter of the fp32 area / linear kernels
// (profiles/r1_step2: XU 59 % busy).  For 0 <= v < 2^23, float(v) == as_float(0x4B000000 | v) - 2^23 exactly (LOP + FADD).
__device__ __forceinline__ float u2f(unsigned v) { return __fsub_rn(__uint_as_float(0x4B000000u | v), 8388608.0f); }

// element <-> float conversions used by the fp32 paths
template <typename T> struct Elem;
template <> struct Elem<uint8_t>  { static constexpr int depth = D_8U;  static __device__ __forceinline__ uint8_t  from_f(float v) { return (uint8_t)sat_u8(rne_f(v)); }
                                    static __device__ __forceinline__ float to_f(uint8_t v) { return u2f(v); } };
template <> struct Elem<uint16_t> { static constexpr int depth = D_16U; static __device__ __forceinline__ uint16_t from_f(float v) { return (uint16_t)sat_u16(rne_f(v)); }
                                    static __device__ __forceinline__ float to_f(uint16_t v) { return u2f(v); } };
template <> struct Elem<int16_t>  { static constexpr int depth = D_16S; static __device__ __forceinline__ int16_t  from_f(float v) { return (int16_t)sat_s16(rne_f(v)); }
                                    static __device__ __forceinline__ float to_f(int16_t v) { return __fsub_rn(u2f((unsigned)(v + 32768)), 32768.0f); } };
template <> struct Elem<float>    { static constexpr int depth = D_32F; static __device__ __forceinline__ float    from_f(float v) { return v; }
                                    static __device__ __forceinline__ float to_f(float v) { return v; } };

// separately-rounded fp32 ops (belt and braces on top of -fmad=false)
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }

// destination store: always evict-first (.cs).  Write-back stores leave the whole output dirty in L2 and, measured on
// B200, cost a streaming kernel 40 % of its steady-state DRAM throughput.  A run-time choice
// between the two store flavours made ptxas duplicate the address arithmetic under predicates, so there is none.
template <typename T> __device__ __forceinline__ void store_out(T* p, T v, int /*keep_in_l2*/)
{
    __stcs(p, v);
}

struct Image {
    const uint8_t* p;
    int step;               // bytes; 32-bit on purpose: row*step is then ONE IMAD.WIDE instead of a 64x64 multiply
    int w, h;
};
struct ImageOut {
    uint8_t* p;
    int step;
    int w, h;
};

} // namespace b200
