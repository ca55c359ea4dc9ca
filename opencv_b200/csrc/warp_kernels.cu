// warpAffine / warpPerspective / remap: generic gather kernels (every depth/channel/border/map cell).
//
// One thread per destination pixel, 32x8 tiles.  Per tile the affine coordinate terms
// (adelta/bdelta per column, X0/Y0 per row -- imgwarp.cpp:2266-2270, :2030-2031) are evaluated once in
// fp64 into shared memory with separately rounded ops and the x86 cvRound rule, so the kernel needs no
// host-built tables and no scratch memory.  warpPerspective evaluates the reference's block-relative
// fp64 form per pixel (imgwarp.cpp:2712-2714, :2803-2821; imgwarp.sse4_1.cpp:313-458).
// remap decodes the three map formats per pixel in registers (imgwarp.cpp:1094-1265).
// Sampling follows remapNearest (imgwarp.cpp:268-369) and remapBilinear (imgwarp.cpp:614-843).
#include "warp_kernels.cuh"
#include "fast_kernels.cuh"
#include "persp_coords.cuh"
#include <type_traits>

namespace b200 {

// ---------------------------------------------------------------------------------------------
// bilinear blend of 4 taps.
// 8U : ((sum v_k*w_k) + 2^14) >> 15 with the int16 table w = 32*{(32-fy)(32-fx),(32-fy)fx,fy(32-fx),fy*fx}
//      (imgwarp.cpp:179-206, FixedPtCast<int,uchar,15> :250-257).  All integer, so it is evaluated
//      separably: s = (32-fy)*((32-fx)v0+fx*v1) + fy*((32-fx)v2+fx*v3), out = (s+512)>>10.  The table's
//      entry 0 is {32767,0,0,1} instead of {32768,0,0,0}; (v0*32767 + v3 + 16384)>>15 == v0 for all
//      bytes, which is what the closed form gives at fx=fy=0, so the quirk is reproduced exactly.
// 16U/16S/32F: fp32 ((v0*w0 + v1*w1) + v2*w2) + v3*w3 with w = (1-fy/32 | fy/32)*(1-fx/32 | fx/32)
//      (products of 6-bit values: exact in fp32, identical to BilinearTab_f), each op rounded.
// ---------------------------------------------------------------------------------------------
template <typename T> struct Blend;
template <> struct Blend<uint8_t> {
    int ax0, ax1, ay0, ay1;
    __device__ __forceinline__ Blend(int fx, int fy) : ax0(32 - fx), ax1(fx), ay0(32 - fy), ay1(fy) {}
    __device__ __forceinline__ uint8_t operator()(uint8_t v0, uint8_t v1, uint8_t v2, uint8_t v3) const {
        int s = ay0 * (ax0 * v0 + ax1 * v1) + ay1 * (ax0 * v2 + ax1 * v3);
        return (uint8_t)((s + 512) >> 10);
    }
};
template <typename T> struct BlendF {
    float w0, w1, w2, w3;
    __device__ __forceinline__ BlendF(int fx, int fy) {
        const float sc = 1.f / 32;
        float x1 = fx * sc, x0 = 1.f - x1, y1 = fy * sc, y0 = 1.f - y1;
        w0 = fmul(y0, x0); w1 = fmul(y0, x1); w2 = fmul(y1, x0); w3 = fmul(y1, x1);
    }
    __device__ __forceinline__ T operator()(T v0, T v1, T v2, T v3) const {
        float t = fadd(fadd(fadd(fmul(Elem<T>::to_f(v0), w0), fmul(Elem<T>::to_f(v1), w1)), fmul(Elem<T>::to_f(v2), w2)), fmul(Elem<T>::to_f(v3), w3));
        return Elem<T>::from_f(t);
    }
};
template <> struct Blend<uint16_t> : BlendF<uint16_t> { using BlendF<uint16_t>::BlendF; };
template <> struct Blend<int16_t>  : BlendF<int16_t>  { using BlendF<int16_t>::BlendF; };
template <> struct Blend<float>    : BlendF<float>    { using BlendF<float>::BlendF; };
// 64F: remapBilinear<Cast<double,double>, RemapNoVec, float> (imgwarp.cpp:1546): the FLOAT table weights, double products and sums
template <> struct Blend<double> {
    double w0, w1, w2, w3;
    __device__ __forceinline__ Blend(int fx, int fy) {
        const float sc = 1.f / 32;
        const float x1 = fx * sc, x0 = 1.f - x1, y1 = fy * sc, y0 = 1.f - y1;
        w0 = (double)fmul(y0, x0); w1 = (double)fmul(y0, x1); w2 = (double)fmul(y1, x0); w3 = (double)fmul(y1, x1);
    }
    __device__ __forceinline__ double operator()(double v0, double v1, double v2, double v3) const {
        return __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(v0, w0), __dmul_rn(v1, w1)), __dmul_rn(v2, w2)), __dmul_rn(v3, w3));
    }
};

// BORDER_TRANSPARENT, bilinear (imgwarp.cpp:726-755): a point on the last row / column lacks neighbours and is blended
// from the taps that exist, renormalised by t * (float)w_total / w_available (8U: the int16 table weights, float
// division, truncation, then the usual (t + 2^14) >> 15); a point outside the image leaves the destination untouched.
template <typename T, int CN>
__device__ __forceinline__ bool sample_linear_transparent(const WarpArgs& a, const uint8_t* __restrict__ src,
                                                       int sx, int sy, int fx, int fy, T* __restrict__ D)
{
    const int W = a.src.w, H = a.src.h;
    if (!(sx >= 0 && sx <= W - 1 && sy >= 0 && sy <= H - 1)) return false;
    const bool a1 = sx < W - 1, a2 = sy < H - 1;
    const T* S = reinterpret_cast<const T*>(src + (long long)sy * a.src.step) + sx * CN;
    const T* S1 = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(S) + (a2 ? a.src.step : 0));
    if constexpr (std::is_same<T, uint8_t>::value) {
        // the int16 table: 32 * products of (32 - f, f); entry 0 is {32767, 0, 0, 1} (imgwarp.cpp:179-206)
        int w0 = (32 - fy) * (32 - fx) * 32, w1 = (32 - fy) * fx * 32, w2 = fy * (32 - fx) * 32, w3 = fy * fx * 32;
        if (fx == 0 && fy == 0) { w0 = 32767; w3 = 1; }
        const int w_tot = w0 + (a1 ? w1 : 0) + (a2 ? w2 : 0) + (a1 && a2 ? w3 : 0);
        if (w_tot == 0) return false;
        const float w_ini = (float)(w0 + w1 + w2 + w3);
#pragma unroll
        for (int k = 0; k < CN; k++) {
            int t0 = S[k] * w0;
            if (a1) t0 += S[k + CN] * w1;
            if (a2) t0 += S1[k] * w2;
            if (a1 && a2) t0 += S1[k + CN] * w3;
            t0 = __float2int_rz(__fdiv_rn(fmul((float)t0, w_ini), (float)w_tot));
            D[k] = (uint8_t)sat_u8((t0 + (1 << 14)) >> 15);
        }
    } else {
        const float sc = 1.f / 32;
        const float x1 = fx * sc, x0 = 1.f - x1, y1 = fy * sc, y0 = 1.f - y1;
        const float w0 = fmul(y0, x0), w1 = fmul(y0, x1), w2 = fmul(y1, x0), w3 = fmul(y1, x1);
        float w_tot = fadd(0.f, w0);
        if (a1) w_tot = fadd(w_tot, w1);
        if (a2) w_tot = fadd(w_tot, w2);
        if (a1 && a2) w_tot = fadd(w_tot, w3);
        if (w_tot == 0.f) return false;
        const float w_ini = fadd(fadd(fadd(w0, w1), w2), w3);
#pragma unroll
        for (int k = 0; k < CN; k++) {
            float t0 = fadd(0.f, fmul(Elem<T>::to_f(S[k]), w0));
            if (a1) t0 = fadd(t0, fmul(Elem<T>::to_f(S[k + CN]), w1));
            if (a2) t0 = fadd(t0, fmul(Elem<T>::to_f(S1[k]), w2));
            if (a1 && a2) t0 = fadd(t0, fmul(Elem<T>::to_f(S1[k + CN]), w3));
            D[k] = Elem<T>::from_f(__fdiv_rn(fmul(t0, w_ini), w_tot));
        }
    }
    return true;
}

// returns false when the destination pixel must be left untouched (BORDER_TRANSPARENT)
template <typename T, int CN>
__device__ __forceinline__ bool sample_linear(const WarpArgs& a, const uint8_t* __restrict__ src,
                                              int sx, int sy, int fx, int fy, T* __restrict__ D)
{
    const int W = a.src.w, H = a.src.h;
    const long long step = a.src.step;
    const T* cv = reinterpret_cast<const T*>(a.cval);
    const T *p0, *p1, *p2, *p3;
    if ((unsigned)sx < (unsigned)max(W - 1, 0) && (unsigned)sy < (unsigned)max(H - 1, 0)) {
        if constexpr (std::is_same<T, uint8_t>::value && CN == 3) {
            // 8UC3 interior pixel: the two 6-byte tap pairs come from three aligned 32-bit words per row (funnel-shifted into
            // place) instead of 12 byte loads, and are blended with IDP.4A -- the exact integer arithmetic of Blend<uint8_t>.
            // Only when the 12-byte window stays inside the row pitch (so it can never leave the allocation).
            const int ob = sx * 3, ob4 = ob & ~3;
            if (a.vec4 && ob4 + 12 <= step) {
                const uint32_t* wt = reinterpret_cast<const uint32_t*>(src + (long long)sy * step + ob4);
                const uint32_t* wb = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(wt) + step);
                const uint32_t t0 = __ldg(wt), t1w = __ldg(wt + 1), t2w = __ldg(wt + 2);
                const uint32_t b0 = __ldg(wb), b1w = __ldg(wb + 1), b2w = __ldg(wb + 2);
                const int sh = (ob & 3) * 8;
                const uint32_t tL = __funnelshift_r(t0, t1w, sh), tH = __funnelshift_r(t1w, t2w, sh);
                const uint32_t bL = __funnelshift_r(b0, b1w, sh), bH = __funnelshift_r(b1w, b2w, sh);
                // weights as 16-bit (top | bottom << 16) pairs, two IDP.2A per channel (see warp_px in fast_kernels.cu)
                const uint32_t ayp = (uint32_t)fy * 0xFFFFu + 32u;
                const uint32_t WR = (uint32_t)fx * ayp, WL = (ayp << 5) - WR;
                const uint32_t p01 = __byte_perm(tL, bL, 0x5140), p23 = __byte_perm(tL, bL, 0x7362), p45 = __byte_perm(tH, bH, 0x5140);
                D[0] = (uint8_t)(__dp2a_hi(WR, p23, __dp2a_lo(WL, p01, 512u)) >> 10);
                D[1] = (uint8_t)(__dp2a_lo(WR, p45, __dp2a_hi(WL, p01, 512u)) >> 10);
                D[2] = (uint8_t)(__dp2a_hi(WR, p45, __dp2a_lo(WL, p23, 512u)) >> 10);
                return true;
            }
        }
        if constexpr (std::is_same<T, uint8_t>::value && CN == 4) {
            // 8UC4 interior pixel (pixels are aligned words: a.vec4): four word loads through the read-only path and the
            // exact integer blend as two IDP.2A per channel; keeps the common case off the generic-pointer border code
            if (a.vec4) {
                const uint32_t* wt = reinterpret_cast<const uint32_t*>(src + (long long)sy * step) + sx;
                const uint32_t* wb = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(wt) + step);
                const uint32_t v0 = __ldg(wt), v1 = __ldg(wt + 1), v2 = __ldg(wb), v3 = __ldg(wb + 1);
                const uint32_t ayp = (uint32_t)fy * 0xFFFFu + 32u;
                const uint32_t WR = (uint32_t)fx * ayp, WL = (ayp << 5) - WR;
                const uint32_t l01 = __byte_perm(v0, v2, 0x5140), l23 = __byte_perm(v0, v2, 0x7362);
                const uint32_t r01 = __byte_perm(v1, v3, 0x5140), r23 = __byte_perm(v1, v3, 0x7362);
                D[0] = (uint8_t)(__dp2a_lo(WR, r01, __dp2a_lo(WL, l01, 512u)) >> 10);
                D[1] = (uint8_t)(__dp2a_hi(WR, r01, __dp2a_hi(WL, l01, 512u)) >> 10);
                D[2] = (uint8_t)(__dp2a_lo(WR, r23, __dp2a_lo(WL, l23, 512u)) >> 10);
                D[3] = (uint8_t)(__dp2a_hi(WR, r23, __dp2a_hi(WL, l23, 512u)) >> 10);
                return true;
            }
        }
        p0 = reinterpret_cast<const T*>(src + (long long)sy * step) + sx * CN;
        p1 = p0 + CN;
        p2 = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(p0) + step);
        p3 = p2 + CN;
    } else {
        const int bt = a.border;
        if (bt == B_TRANSPARENT) {
            if constexpr (std::is_same<T, double>::value) return false;                   // declined on the host
            else return sample_linear_transparent<T, CN>(a, src, sx, sy, fx, fy, D);
        }
        if (bt == B_CONSTANT && (sx >= W || sx + 1 < 0 || sy >= H || sy + 1 < 0)) {     // imgwarp.cpp:807-813
#pragma unroll
            for (int k = 0; k < CN; k++) D[k] = cv[k];
            return true;
        }
        int sx0, sx1, sy0, sy1;
        if (bt == B_REPLICATE) {                                                      // :822-826
            sx0 = clipi(sx, 0, W); sx1 = clipi(sx + 1, 0, W);
            sy0 = clipi(sy, 0, H); sy1 = clipi(sy + 1, 0, H);
        } else {                                                                      // :834-837
            sx0 = border_idx(sx, W, bt); sx1 = border_idx(sx + 1, W, bt);
            sy0 = border_idx(sy, H, bt); sy1 = border_idx(sy + 1, H, bt);
        }
        const T* r0 = reinterpret_cast<const T*>(src + (long long)max(sy0, 0) * step);
        const T* r1 = reinterpret_cast<const T*>(src + (long long)max(sy1, 0) * step);
        p0 = (sx0 >= 0 && sy0 >= 0) ? r0 + sx0 * CN : cv;                             // :838-841
        p1 = (sx1 >= 0 && sy0 >= 0) ? r0 + sx1 * CN : cv;
        p2 = (sx0 >= 0 && sy1 >= 0) ? r1 + sx0 * CN : cv;
        p3 = (sx1 >= 0 && sy1 >= 0) ? r1 + sx1 * CN : cv;
    }
    if constexpr (std::is_same<T, uint8_t>::value && CN == 4) {
        // 8UC4: pixels are 4-byte aligned words (checked by the host: a.vec4) -> one load per tap and the same
        // exact integer blend as Blend<uint8_t>, two channels per PRMT, IDP.4A horizontally, IMAD vertically
        if (a.vec4) {
            const uint32_t v0 = *reinterpret_cast<const uint32_t*>(p0), v1 = *reinterpret_cast<const uint32_t*>(p1);
            const uint32_t v2 = *reinterpret_cast<const uint32_t*>(p2), v3 = *reinterpret_cast<const uint32_t*>(p3);
            const uint32_t ax0 = 32 - fx, ax1 = fx, ay0 = 32 - fy, ay1 = fy;
            const uint32_t te = __byte_perm(v0, v1, 0x5140), tf = __byte_perm(v0, v1, 0x7362);
            const uint32_t be = __byte_perm(v2, v3, 0x5140), bf = __byte_perm(v2, v3, 0x7362);
            const uint32_t wlo = ax0 | (ax1 << 8), whi = wlo << 16;
            const uint32_t s0 = ay0 * __dp4a(te, wlo, 0u) + ay1 * __dp4a(be, wlo, 0u) + 512u;
            const uint32_t s1 = ay0 * __dp4a(te, whi, 0u) + ay1 * __dp4a(be, whi, 0u) + 512u;
            const uint32_t s2 = ay0 * __dp4a(tf, wlo, 0u) + ay1 * __dp4a(bf, wlo, 0u) + 512u;
            const uint32_t s3 = ay0 * __dp4a(tf, whi, 0u) + ay1 * __dp4a(bf, whi, 0u) + 512u;
            D[0] = (uint8_t)(s0 >> 10); D[1] = (uint8_t)(s1 >> 10); D[2] = (uint8_t)(s2 >> 10); D[3] = (uint8_t)(s3 >> 10);
            return true;
        }
    }
    Blend<T> blend(fx, fy);
#pragma unroll
    for (int k = 0; k < CN; k++) D[k] = blend(p0[k], p1[k], p2[k], p3[k]);
    return true;
}

// nearest: copies one element of N units of type U (imgwarp.cpp:268-369)
template <typename U, int N>
__device__ __forceinline__ bool sample_nearest(const WarpArgs& a, const uint8_t* __restrict__ src,
                                               int sx, int sy, U* __restrict__ D)
{
    const int W = a.src.w, H = a.src.h;
    const U* S;
    if ((unsigned)sx < (unsigned)W && (unsigned)sy < (unsigned)H)
        S = reinterpret_cast<const U*>(src + (long long)sy * a.src.step) + sx * N;
    else {
        const int bt = a.border;
        if (bt == B_TRANSPARENT) return false;                                       // imgwarp.cpp:312
        if (bt == B_CONSTANT) S = reinterpret_cast<const U*>(a.cval);
        else {
            if (bt == B_REPLICATE) { sx = clipi(sx, 0, W); sy = clipi(sy, 0, H); }
            else { sx = border_idx(sx, W, bt); sy = border_idx(sy, H, bt); }
            S = reinterpret_cast<const U*>(src + (long long)sy * a.src.step) + sx * N;
        }
    }
#pragma unroll
    for (int k = 0; k < N; k++) D[k] = S[k];
    return true;
}

// ---------------------------------------------------------------------------------------------
// coordinate generators: produce the reference's fixed-point source coordinate for dst pixel (x,y).
//   LINEAR : (sx, sy) = sat16(X>>5), fractional (fx, fy) = X&31     NEAREST: (sx, sy) only
// ---------------------------------------------------------------------------------------------
constexpr int TILE_W = 32, TILE_H = 8;

struct AffineTile {        // shared-memory terms of one tile
    int adelta[TILE_W], bdelta[TILE_W], X0[TILE_H], Y0[TILE_H];
};

// adelta[x] = cvRound(M0*x*1024) (imgwarp.cpp:2268), X0 = cvRound((M1*y+M2)*1024)+round_delta (:2030)
__device__ __forceinline__ void affine_tile_setup(const WarpArgs& a, AffineTile& t, int x0, int y0, int tid)
{
    const int rd = a.linear ? 16 : 512;      // AB_SCALE/INTER_TAB_SIZE/2 | AB_SCALE/2   (:2011)
    if (tid < TILE_W) {
        double x = (double)(x0 + tid);
        t.adelta[tid] = rne_d(__dmul_rn(__dmul_rn(a.M[0], x), 1024.0));
        t.bdelta[tid] = rne_d(__dmul_rn(__dmul_rn(a.M[3], x), 1024.0));
    } else if (tid < TILE_W + TILE_H) {
        int r = tid - TILE_W;
        double y = (double)(y0 + r);
        // int adds wrap in the reference (CV_DISABLE_UBSAN, :2279): unsigned arithmetic here
        t.X0[r] = (int)((unsigned)rne_d(__dmul_rn(__dadd_rn(__dmul_rn(a.M[1], y), a.M[2]), 1024.0)) + (unsigned)rd);
        t.Y0[r] = (int)((unsigned)rne_d(__dmul_rn(__dadd_rn(__dmul_rn(a.M[4], y), a.M[5]), 1024.0)) + (unsigned)rd);
    }
}

template <int COORD, bool LINEAR>
__device__ __forceinline__ void map_coord(const WarpArgs& a, int x, int y, int& sx, int& sy, int& fx, int& fy)
{
    fx = fy = 0;
    if (COORD == COORD_MAP16S) {
        const short2 xy = *reinterpret_cast<const short2*>(a.map1 + (long long)y * a.map1_step + (long long)x * 4);
        sx = xy.x; sy = xy.y;
        if (LINEAR) {                                                   // imgwarp.cpp:1168-1185
            int al = *reinterpret_cast<const uint16_t*>(a.map2 + (long long)y * a.map2_step + (long long)x * 2) & 1023;
            fx = al & 31; fy = al >> 5;
        } else if (a.map2) {                                            // :1109-1121 (+NNDeltaTab, int16 wrap)
            int al = *reinterpret_cast<const uint16_t*>(a.map2 + (long long)y * a.map2_step + (long long)x * 2) & 1023;
            sx = (short)(sx + ((al & 31) >= 16)); sy = (short)(sy + ((al >> 5) >= 16));
        }
    } else {
        float mx, my;
        if (COORD == COORD_MAP32F) {
            mx = *reinterpret_cast<const float*>(a.map1 + (long long)y * a.map1_step + (long long)x * 4);
            my = *reinterpret_cast<const float*>(a.map2 + (long long)y * a.map2_step + (long long)x * 4);
        } else {
            const float2 m = *reinterpret_cast<const float2*>(a.map1 + (long long)y * a.map1_step + (long long)x * 8);
            mx = m.x; my = m.y;
        }
        if (LINEAR) {                                                   // :1186-1263
            int isx = rne_f(fmul(mx, 32.f)), isy = rne_f(fmul(my, 32.f));
            sx = sat_s16(isx >> INTER_BITS); sy = sat_s16(isy >> INTER_BITS);
            fx = isx & 31; fy = isy & 31;
        } else {                                                        // :1123-1156 saturate_cast<short>(float)
            sx = sat_s16(rne_f(mx)); sy = sat_s16(rne_f(my));
        }
    }
    if (a.relative) { sx += x; sy += y; }                               // :293-299, :647-652
}

// ---------------------------------------------------------------------------------------------
// T/CN: element type and channels for LINEAR; for NEAREST T is the copy unit and CN the unit count.
// ---------------------------------------------------------------------------------------------
// PXT rows per thread (rows y, y + TILE_H, ...): the per-thread set-up and the destination address arithmetic are a fifth of
// the ~170 instructions a perspective / remap pixel costs, so large images run with PXT = 4.
// MF (multi-frame): the frames of a batch share the transform / the map, so a thread derives the coordinate of its pixel
// ONCE and samples frames_per_thread frames with it -- the fp64 perspective coordinates (~110 of the ~170 instructions of a
// pixel) and the map decode are paid once per 8 pixels.  Without MF the frame loop folds away (one frame per blockIdx.z).
template <typename T, int CN, int COORD, bool LINEAR, int PXT = 1, bool MF = false>
// (4 CTAs per SM = 64 registers for the multi-frame form: 201 us for warpPerspective 4K 8UC4 x 8 frames, 240 us at 3 CTAs / 80
// registers; the double instantiations need the 80)
__global__ void __launch_bounds__(TILE_W * TILE_H, MF ? (sizeof(T) == 8 ? 3 : 4) : 1)
warp_generic_kernel(const WarpArgs a)
{
    static_assert(PXT == 1 || COORD != COORD_AFFINE, "the affine tile tables cover TILE_H rows");
    static_assert(!MF || PXT == 1, "multi-frame threads own one row");
    constexpr int TROWS = TILE_H * PXT;
    __shared__ AffineTile tile;
    __shared__ PerspRow prow[TROWS];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int x0 = blockIdx.x * TILE_W, y0 = blockIdx.y * TROWS;
    const int x = x0 + tx;
    const int f0 = MF ? blockIdx.z * a.frames_per_thread : blockIdx.z;
    const int nf = MF ? min(a.frames_per_thread, a.n_frames - f0) : 1;
    const uint8_t* src0 = a.src.p + (long long)f0 * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)f0 * a.dst_frame_stride;

    if (COORD == COORD_AFFINE) {
        affine_tile_setup(a, tile, x0, y0, ty * TILE_W + tx);
        __syncthreads();
    }
    // bw0 is a power of two >= 32 for every image of at least 64x16 (then bw0_shift = log2): the whole 32-wide tile lies
    // inside one coordinate block and block origins come from shifts, not from an emulated integer division per pixel
    const bool persp_shared = COORD == COORD_PERSP && a.bw0_shift >= 5;
    if (COORD == COORD_PERSP && persp_shared) {
        const int t = ty * TILE_W + tx;
        if (t < TROWS) prow[t] = persp_row_terms(a, (x0 >> a.bw0_shift) << a.bw0_shift, y0 + t);
        __syncthreads();
    }
    if (x >= a.dst.w) return;
    uint8_t* drow = dst + (long long)(y0 + ty) * a.dst.step + (long long)x * (CN * (int)sizeof(T));
#pragma unroll
    for (int i = 0; i < PXT; i++, drow += (long long)TILE_H * a.dst.step) {
        const int r = ty + i * TILE_H, y = y0 + r;
        if (y >= a.dst.h) break;
        int sx, sy, fx, fy;
        if (COORD == COORD_AFFINE) {
            int X = (int)((unsigned)tile.X0[ty] + (unsigned)tile.adelta[tx]);
            int Y = (int)((unsigned)tile.Y0[ty] + (unsigned)tile.bdelta[tx]);
            if constexpr (LINEAR) {                                         // imgwarp.cpp:2351-2358
                X >>= 5; Y >>= 5;
                sx = sat_s16(X >> 5); sy = sat_s16(Y >> 5); fx = X & 31; fy = Y & 31;
            } else {                                                        // :2301-2307
                sx = sat_s16(X >> 10); sy = sat_s16(Y >> 10); fx = fy = 0;
            }
        } else if (COORD == COORD_PERSP) {
            const int xb = persp_shared ? ((x >> a.bw0_shift) << a.bw0_shift) : (x / a.bw0) * a.bw0;
            const PerspRow pr = persp_shared ? prow[r] : persp_row_terms(a, xb, y);
            persp_coord<LINEAR>(a, pr, x - xb, min(a.bw0, a.dst.w - xb), sx, sy, fx, fy);
        } else {
            map_coord<COORD, LINEAR>(a, x, y, sx, sy, fx, fy);
        }

        const uint8_t* src = src0;
        uint8_t* dpx = drow;
        if constexpr (MF && LINEAR && std::is_same<T, uint8_t>::value && (CN == 3 || CN == 4)) {
            // 8UC3 / 8UC4 interior pixel of a multi-frame thread: whether the taps are interior, their offset, the funnel shift
            // and the weights are the same for every frame, so the tap words of FOUR frames are loaded before the first blend
            // (one frame at a time left a thread waiting a DRAM round trip per frame: 146 us for 32 frames of 720p)
            const int W = a.src.w, H = a.src.h;
            const long long step = a.src.step;
            const int ob = sx * CN, ob4 = CN == 3 ? (ob & ~3) : ob;
            if ((unsigned)sx < (unsigned)max(W - 1, 0) && (unsigned)sy < (unsigned)max(H - 1, 0) && a.vec4 &&
                (CN == 4 || ob4 + 12 <= step)) {
                constexpr int NWD = CN == 3 ? 3 : 2;                  // words per tap row
                const long long toff = (long long)sy * step + ob4;
                const int sh = (ob & 3) * 8;
                const uint32_t ayp = (uint32_t)fy * 0xFFFFu + 32u;
                const uint32_t WR = (uint32_t)fx * ayp, WL = (ayp << 5) - WR;
                for (int f = 0; f < nf; f += 4) {
                    uint32_t t[4][NWD], b[4][NWD];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (f + j < nf) {
                            const uint32_t* wt = reinterpret_cast<const uint32_t*>(src + (long long)j * a.src_frame_stride + toff);
                            const uint32_t* wb = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(wt) + step);
#pragma unroll
                            for (int k = 0; k < NWD; k++) { t[j][k] = __ldg(wt + k); b[j][k] = __ldg(wb + k); }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        if (f + j < nf) {
                            uint8_t* D = dpx + (long long)j * a.dst_frame_stride;
                            if constexpr (CN == 3) {
                                const uint32_t tL = __funnelshift_r(t[j][0], t[j][1], sh), tH = __funnelshift_r(t[j][1], t[j][2], sh);
                                const uint32_t bL = __funnelshift_r(b[j][0], b[j][1], sh), bH = __funnelshift_r(b[j][1], b[j][2], sh);
                                const uint32_t p01 = __byte_perm(tL, bL, 0x5140), p23 = __byte_perm(tL, bL, 0x7362), p45 = __byte_perm(tH, bH, 0x5140);
                                store_out(D, (uint8_t)(__dp2a_hi(WR, p23, __dp2a_lo(WL, p01, 512u)) >> 10), 0);
                                store_out(D + 1, (uint8_t)(__dp2a_lo(WR, p45, __dp2a_hi(WL, p01, 512u)) >> 10), 0);
                                store_out(D + 2, (uint8_t)(__dp2a_hi(WR, p45, __dp2a_lo(WL, p23, 512u)) >> 10), 0);
                            } else {
                                const uint32_t l01 = __byte_perm(t[j][0], b[j][0], 0x5140), l23 = __byte_perm(t[j][0], b[j][0], 0x7362);
                                const uint32_t r01 = __byte_perm(t[j][1], b[j][1], 0x5140), r23 = __byte_perm(t[j][1], b[j][1], 0x7362);
                                const uint32_t s0 = __dp2a_lo(WR, r01, __dp2a_lo(WL, l01, 512u)), s1 = __dp2a_hi(WR, r01, __dp2a_hi(WL, l01, 512u));
                                const uint32_t s2 = __dp2a_lo(WR, r23, __dp2a_lo(WL, l23, 512u)), s3 = __dp2a_hi(WR, r23, __dp2a_hi(WL, l23, 512u));
                                store_out(reinterpret_cast<uint32_t*>(D), (s0 >> 10) | ((s1 >> 10) << 8) | ((s2 >> 10) << 16) | ((s3 >> 10) << 24), 0);
                            }
                        }
                    }
                    src += 4 * a.src_frame_stride; dpx += 4 * a.dst_frame_stride;
                }
                continue;
            }
        }
#pragma unroll 2
        for (int f = 0; f < nf; f++, src += a.src_frame_stride, dpx += a.dst_frame_stride) {
            T out[CN];
            bool written;
            if constexpr (LINEAR) written = sample_linear<T, CN>(a, src, sx, sy, fx, fy, out);
            else written = sample_nearest<T, CN>(a, src, sx, sy, out);
            if (!written) continue;
            T* D = reinterpret_cast<T*>(dpx);
            if constexpr (sizeof(T) == 1 && CN == 4) {
                if (a.vec4) {
                    const uint32_t w = (uint32_t)out[0] | ((uint32_t)out[1] << 8) | ((uint32_t)out[2] << 16) | ((uint32_t)out[3] << 24);
                    store_out(reinterpret_cast<uint32_t*>(D), w, a.keep_in_l2);
                    continue;
                }
            }
#pragma unroll
            for (int k = 0; k < CN; k++) store_out(&D[k], out[k], a.keep_in_l2);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// INTER_CUBIC (4x4 taps from sx-1, sy-1) and INTER_LANCZOS4 (8x8 taps from sx-3, sy-3):
// remapBicubic imgwarp.cpp:846-949, remapLanczos4 :952-1059.  The fractional index and the coordinate
// derivation are those of INTER_LINEAR.  8U: int sums over the int16 product table (host-built with the
// reference's sum fix-up, :154-206) and (sum + 2^14) >> 15; 16U/16S/32F: fp32 with w[r][c] = vy[r]*vx[c]
// (the float table's own definition, :183-184), products added left to right one row at a time; border
// pixels as cv*ONE + sum (S - cv)*w in tap order.  Channel count and coordinate kind are run-time
// arguments here (coverage kernels, 8 instantiations instead of 160).
// ---------------------------------------------------------------------------------------------
template <typename T, int KS>
__device__ __forceinline__ bool sample_hq(const WarpArgs& a, const uint8_t* __restrict__ src, int cn,
                                          int sx, int sy, int fx, int fy, const float* __restrict__ tab1, T* __restrict__ D)
{
    constexpr bool U8 = std::is_same<T, uint8_t>::value;
    const int W = a.src.w, H = a.src.h;
    const long long step = a.src.step;
    const T* cvp = reinterpret_cast<const T*>(a.cval);
    sx -= KS / 2 - 1; sy -= KS / 2 - 1;
    const short* wi = a.tab2d_i + (fy * 32 + fx) * (KS * KS);
    const float* vx = tab1 + fx * KS;
    const float* vy = tab1 + fy * KS;
    if ((unsigned)sx < (unsigned)max(W - (KS - 1), 0) && (unsigned)sy < (unsigned)max(H - (KS - 1), 0)) {
        const T* S0 = reinterpret_cast<const T*>(src + (long long)sy * step) + sx * cn;
        for (int k = 0; k < cn; k++) {
            if constexpr (U8) {
                int sum = 0;
#pragma unroll
                for (int r = 0; r < KS; r++) {
                    const uint8_t* S = S0 + (long long)r * step + k;
#pragma unroll
                    for (int c = 0; c < KS; c++) sum += (int)S[c * cn] * (int)__ldg(wi + r * KS + c);
                }
                D[k] = (uint8_t)sat_u8((sum + (1 << 14)) >> 15);
            } else if constexpr (std::is_same<T, double>::value) {     // Cast<double,double> with float weights: double products and sums
                double sum = 0.0;
#pragma unroll
                for (int r = 0; r < KS; r++) {
                    const double* S = reinterpret_cast<const double*>(reinterpret_cast<const uint8_t*>(S0) + (long long)r * step) + k;
                    const float wy = vy[r];
                    double row = __dmul_rn(S[0], (double)fmul(wy, vx[0]));
#pragma unroll
                    for (int c = 1; c < KS; c++) row = __dadd_rn(row, __dmul_rn(S[c * cn], (double)fmul(wy, vx[c])));
                    sum = (r == 0 && KS == 4) ? row : __dadd_rn(sum, row);
                }
                D[k] = sum;
            } else {
                float sum = 0.f;
#pragma unroll
                for (int r = 0; r < KS; r++) {
                    const T* S = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(S0) + (long long)r * step) + k;
                    const float wy = vy[r];
                    float row = fmul(Elem<T>::to_f(S[0]), fmul(wy, vx[0]));
#pragma unroll
                    for (int c = 1; c < KS; c++) row = fadd(row, fmul(Elem<T>::to_f(S[c * cn]), fmul(wy, vx[c])));
                    sum = (r == 0 && KS == 4) ? row : fadd(sum, row);       // bicubic: first row initialises (:885)
                }
                D[k] = Elem<T>::from_f(sum);
            }
        }
        return true;
    }
    int bt = a.border;
    if (bt == B_TRANSPARENT) {                                                          // :862, :897-900
        if ((unsigned)(sx + KS / 2 - 1) >= (unsigned)W || (unsigned)(sy + KS / 2 - 1) >= (unsigned)H) return false;
        bt = B_REFLECT_101;
    }
    if (bt == B_CONSTANT && (sx >= W || sx + KS <= 0 || sy >= H || sy + KS <= 0)) {     // :905-912, :1011-1018
        for (int k = 0; k < cn; k++) D[k] = cvp[k];
        return true;
    }
    int xs[KS], ys[KS];
#pragma unroll
    for (int i = 0; i < KS; i++) {
        xs[i] = border_idx(sx + i, W, bt);
        ys[i] = border_idx(sy + i, H, bt);
    }
    for (int k = 0; k < cn; k++) {
        if constexpr (U8) {
            const int cv = cvp[k];
            int sum = cv << 15;
#pragma unroll
            for (int i = 0; i < KS; i++) {
                if (ys[i] < 0) continue;
                const uint8_t* S = src + (long long)ys[i] * step + k;
#pragma unroll
                for (int c = 0; c < KS; c++)
                    if (xs[c] >= 0) sum += ((int)S[xs[c] * cn] - cv) * (int)__ldg(wi + i * KS + c);
            }
            D[k] = (uint8_t)sat_u8((sum + (1 << 14)) >> 15);
        } else if constexpr (std::is_same<T, double>::value) {
            const double cv = cvp[k];
            double sum = cv;
#pragma unroll
            for (int i = 0; i < KS; i++) {
                if (ys[i] < 0) continue;
                const double* S = reinterpret_cast<const double*>(src + (long long)ys[i] * step) + k;
                const float wy = vy[i];
#pragma unroll
                for (int c = 0; c < KS; c++)
                    if (xs[c] >= 0) sum = __dadd_rn(sum, __dmul_rn(__dsub_rn(S[xs[c] * cn], cv), (double)fmul(wy, vx[c])));
            }
            D[k] = sum;
        } else {
            const float cv = Elem<T>::to_f(cvp[k]);
            float sum = cv;
#pragma unroll
            for (int i = 0; i < KS; i++) {
                if (ys[i] < 0) continue;
                const T* S = reinterpret_cast<const T*>(src + (long long)ys[i] * step) + k;
                const float wy = vy[i];
#pragma unroll
                for (int c = 0; c < KS; c++)
                    if (xs[c] >= 0) sum = fadd(sum, fmul(fsub(Elem<T>::to_f(S[xs[c] * cn]), cv), fmul(wy, vx[c])));
            }
            D[k] = Elem<T>::from_f(sum);
        }
    }
    return true;
}

template <typename T, int KS>
__global__ void __launch_bounds__(TILE_W * TILE_H)
warp_hq_kernel(const WarpArgs a, int coord, int cn)
{
    __shared__ AffineTile tile;
    __shared__ PerspRow prow[TILE_H];
    __shared__ float tab1[32 * KS];
    const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * TILE_W + tx;
    const int x0 = blockIdx.x * TILE_W, y0 = blockIdx.y * TILE_H;
    const int x = x0 + tx, y = y0 + ty;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride;
    if (tid < 32 * KS) tab1[tid] = a.tab1d[tid];
    const bool persp_shared = coord == COORD_PERSP && a.bw0_shift >= 5;
    if (coord == COORD_AFFINE) affine_tile_setup(a, tile, x0, y0, tid);
    else if (persp_shared && ty == 0 && tx < TILE_H) prow[tx] = persp_row_terms(a, (x0 >> a.bw0_shift) << a.bw0_shift, y0 + tx);
    __syncthreads();
    if (x >= a.dst.w || y >= a.dst.h) return;

    int sx, sy, fx, fy;
    if (coord == COORD_AFFINE) {                                        // imgwarp.cpp:2351-2358
        int X = (int)((unsigned)tile.X0[ty] + (unsigned)tile.adelta[tx]);
        int Y = (int)((unsigned)tile.Y0[ty] + (unsigned)tile.bdelta[tx]);
        X >>= 5; Y >>= 5;
        sx = sat_s16(X >> 5); sy = sat_s16(Y >> 5); fx = X & 31; fy = Y & 31;
    } else if (coord == COORD_PERSP) {
        const int xb = persp_shared ? ((x >> a.bw0_shift) << a.bw0_shift) : (x / a.bw0) * a.bw0;
        const PerspRow pr = persp_shared ? prow[ty] : persp_row_terms(a, xb, y);
        persp_coord<true>(a, pr, x - xb, min(a.bw0, a.dst.w - xb), sx, sy, fx, fy);
    } else if (coord == COORD_MAP32F) map_coord<COORD_MAP32F, true>(a, x, y, sx, sy, fx, fy);
    else if (coord == COORD_MAP32FC2) map_coord<COORD_MAP32FC2, true>(a, x, y, sx, sy, fx, fy);
    else map_coord<COORD_MAP16S, true>(a, x, y, sx, sy, fx, fy);

    T out[4];
    if (!sample_hq<T, KS>(a, src, cn, sx, sy, fx, fy, tab1, out)) return;
    T* D = reinterpret_cast<T*>(dst + (long long)y * a.dst.step) + x * cn;
    for (int k = 0; k < cn; k++) store_out(&D[k], out[k], a.keep_in_l2);
}

template <typename T>
static cudaError_t launch_hq(const WarpArgs& a, int cn, int coord, dim3 grid, dim3 block, cudaStream_t s)
{
    if (cn < 1 || cn > 4 || !a.tab1d || !a.tab2d_i) return cudaErrorInvalidValue;
    if (a.interp == 2) warp_hq_kernel<T, 4><<<grid, block, 0, s>>>(a, coord, cn);
    else warp_hq_kernel<T, 8><<<grid, block, 0, s>>>(a, coord, cn);
    return cudaGetLastError();
}

template <typename T, int CN, bool LINEAR>
static cudaError_t launch_coord(const WarpArgs& a, int coord, dim3 grid, dim3 block, cudaStream_t s)
{
    if constexpr (LINEAR) {
        if (a.frames_per_thread > 1 && coord != COORD_AFFINE) {
            switch (coord) {
            case COORD_PERSP:    warp_generic_kernel<T, CN, COORD_PERSP, true, 1, true><<<grid, block, 0, s>>>(a); break;
            case COORD_MAP32F:   warp_generic_kernel<T, CN, COORD_MAP32F, true, 1, true><<<grid, block, 0, s>>>(a); break;
            case COORD_MAP32FC2: warp_generic_kernel<T, CN, COORD_MAP32FC2, true, 1, true><<<grid, block, 0, s>>>(a); break;
            case COORD_MAP16S:   warp_generic_kernel<T, CN, COORD_MAP16S, true, 1, true><<<grid, block, 0, s>>>(a); break;
            default: return cudaErrorInvalidValue;
            }
            return cudaGetLastError();
        }
        // 4 rows per thread once the image keeps every SM busy for several waves (measured at 4K and on 32-frame batches)
        if (coord != COORD_AFFINE && (long long)a.dst.w * a.dst.h * grid.z >= 4000000LL) {
            dim3 g4(grid.x, (a.dst.h + TILE_H * 4 - 1) / (TILE_H * 4), grid.z);
            switch (coord) {
            case COORD_PERSP:    warp_generic_kernel<T, CN, COORD_PERSP, true, 4><<<g4, block, 0, s>>>(a); break;
            case COORD_MAP32F:   warp_generic_kernel<T, CN, COORD_MAP32F, true, 4><<<g4, block, 0, s>>>(a); break;
            case COORD_MAP32FC2: warp_generic_kernel<T, CN, COORD_MAP32FC2, true, 4><<<g4, block, 0, s>>>(a); break;
            case COORD_MAP16S:   warp_generic_kernel<T, CN, COORD_MAP16S, true, 4><<<g4, block, 0, s>>>(a); break;
            default: return cudaErrorInvalidValue;
            }
            return cudaGetLastError();
        }
    }
    switch (coord) {
    case COORD_AFFINE:   warp_generic_kernel<T, CN, COORD_AFFINE, LINEAR><<<grid, block, 0, s>>>(a); break;
    case COORD_PERSP:    warp_generic_kernel<T, CN, COORD_PERSP, LINEAR><<<grid, block, 0, s>>>(a); break;
    case COORD_MAP32F:   warp_generic_kernel<T, CN, COORD_MAP32F, LINEAR><<<grid, block, 0, s>>>(a); break;
    case COORD_MAP32FC2: warp_generic_kernel<T, CN, COORD_MAP32FC2, LINEAR><<<grid, block, 0, s>>>(a); break;
    case COORD_MAP16S:   warp_generic_kernel<T, CN, COORD_MAP16S, LINEAR><<<grid, block, 0, s>>>(a); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

template <typename T, bool LINEAR>
static cudaError_t launch_cn(const WarpArgs& a, int cn, int coord, dim3 grid, dim3 block, cudaStream_t s)
{
    switch (cn) {
    case 1: return launch_coord<T, 1, LINEAR>(a, coord, grid, block, s);
    case 2: return launch_coord<T, 2, LINEAR>(a, coord, grid, block, s);
    case 3: return launch_coord<T, 3, LINEAR>(a, coord, grid, block, s);
    case 4: return launch_coord<T, 4, LINEAR>(a, coord, grid, block, s);
    default: return cudaErrorInvalidValue;
    }
}

static int launch_generic(const WarpArgs& a, int depth, int cn, int coord, dim3 grid, dim3 block, cudaStream_t stream);

int launch_warp(const WarpArgs& a, int depth, int cn, int coord, int n_frames, cudaStream_t stream)
{
    if (coord == COORD_AFFINE && depth == D_8U && a.interp == 1 && !a.force_generic) {
        int rc = try_launch_warp_affine_tma_u8(a, cn, n_frames, stream);
        if (rc >= 0) return rc;                       // launched (0) or a CUDA error
    }
    if (coord == COORD_PERSP && depth == D_8U && a.interp == 1 && !a.force_generic) {
        int rc = try_launch_warp_persp_tma_u8(a, cn, n_frames, stream);
        if (rc >= 0) return rc;
    }
    dim3 block(TILE_W, TILE_H);
    dim3 grid((a.dst.w + TILE_W - 1) / TILE_W, (a.dst.h + TILE_H - 1) / TILE_H, n_frames);
    if (a.interp == 2 || a.interp == 4) {
        switch (depth) {
        case D_8U:  return (int)launch_hq<uint8_t>(a, cn, coord, grid, block, stream);
        case D_16U: return (int)launch_hq<uint16_t>(a, cn, coord, grid, block, stream);
        case D_16S: return (int)launch_hq<int16_t>(a, cn, coord, grid, block, stream);
        case D_32F: return (int)launch_hq<float>(a, cn, coord, grid, block, stream);
        case D_64F: return (int)launch_hq<double>(a, cn, coord, grid, block, stream);
        default: return (int)cudaErrorInvalidValue;
        }
    }
    // generic gather kernel: up to 8 frames of a batch per thread, as long as the grid still covers the machine a few times
    WarpArgs b = a;
    b.n_frames = n_frames; b.frames_per_thread = 1;
    const long long px = (long long)a.dst.w * a.dst.h;
    if (a.linear && coord != COORD_AFFINE)
        for (int fpt = 8; fpt > 1; fpt >>= 1)
            if (n_frames >= fpt && px * (n_frames / fpt) >= 600000LL) { b.frames_per_thread = fpt; break; }
    grid.z = (n_frames + b.frames_per_thread - 1) / b.frames_per_thread;
    return launch_generic(b, depth, cn, coord, grid, block, stream);
}

static int launch_generic(const WarpArgs& a, int depth, int cn, int coord, dim3 grid, dim3 block, cudaStream_t stream)
{
    if (a.linear) {
        switch (depth) {
        case D_8U:  return (int)launch_cn<uint8_t, true>(a, cn, coord, grid, block, stream);
        case D_16U: return (int)launch_cn<uint16_t, true>(a, cn, coord, grid, block, stream);
        case D_16S: return (int)launch_cn<int16_t, true>(a, cn, coord, grid, block, stream);
        case D_32F: return (int)launch_cn<float, true>(a, cn, coord, grid, block, stream);
        case D_64F: return (int)launch_cn<double, true>(a, cn, coord, grid, block, stream);
        default: return (int)cudaErrorInvalidValue;
        }
    }
    // NEAREST: pure element copy -> instantiate on the channel unit size only
    const int esz1 = depth <= D_8S ? 1 : depth <= D_16S ? 2 : depth <= D_32F ? 4 : 8;
    switch (esz1) {
    case 1: return (int)launch_cn<uint8_t, false>(a, cn, coord, grid, block, stream);
    case 2: return (int)launch_cn<uint16_t, false>(a, cn, coord, grid, block, stream);
    case 4: return (int)launch_cn<uint32_t, false>(a, cn, coord, grid, block, stream);
    default: return (int)launch_cn<unsigned long long, false>(a, cn, coord, grid, block, stream);
    }
}

} // namespace b200
