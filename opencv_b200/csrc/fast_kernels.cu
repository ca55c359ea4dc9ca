// Hot-cell kernels for CV_8U (the generic family in warp_kernels.cu / resize_kernels.cu stays the
// fallback and produces identical bytes):
//
//  * warp_affine_tma_u8<CN,TW,TH>: cv::warpAffine INTER_LINEAR for 8UC1/C3/C4.  The transform bounds the
//    source footprint of a TWxTH output tile by an axis-aligned box (both coordinate terms are monotone in
//    their index, so the corners are the exact extremes); that box is staged into shared memory by ONE
//    TMA tile load (cp.async.bulk.tensor.3d, byte tensor of width cols*cn, zero OOB fill), border pixels
//    are patched in shared memory for tiles that cross the image edge, taps are gathered from shared
//    memory with aligned 32-bit loads + funnel shifts, blended with IDP.4A + IMAD in the reference's exact
//    integer arithmetic, written to a shared output tile and stored with ONE TMA tile store (edge clipping
//    by the tensor map).  Coordinates: reference AB_BITS=10 fixed point, evaluated per tile in fp64
//    (imgwarp.cpp:2266-2270, :2030-2031, :2351-2358), blend = FixedPtCast<int,uchar,15> over
//    BilinearTab_i (imgwarp.cpp:179-206,:250-257,:614-843) in its separable closed form (warp_kernels.cu).
//
//  * box2x2_u8<CN>: cv::resize's exact-2x path for 8U ((a+b+c+d+2)>>2, resize.cpp:2937-2987, reached by
//    INTER_AREA and by INTER_LINEAR at exact 2x, :3817): rows handled as byte streams, 16-byte coalesced
//    vector loads, 8/16-byte vector stores.
#include "fast_kernels.cuh"
#include "persp_coords.cuh"
#include "tma_utils.cuh"
#include <cstdlib>
#include <cstdio>

namespace b200 {

// =========================================================================== 2x2 box, 8U =========
template <int CN> struct BoxCfg;
template <> struct BoxCfg<1> { static constexpr int G = 16, NV = 2; };   // 16 dst px <- 32 src bytes / row
template <> struct BoxCfg<3> { static constexpr int G = 8,  NV = 3; };   //  8 dst px <- 48 src bytes / row
template <> struct BoxCfg<4> { static constexpr int G = 4,  NV = 2; };   //  4 dst px <- 32 src bytes / row

__device__ __forceinline__ uint4 ldg_stream(const uint4* p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ void stg_cs(uint2* p, uint2 v) { asm volatile("st.global.cs.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory"); }
__device__ __forceinline__ void stg_cs(uint4* p, uint4 v) { asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }

// the arithmetic of one group: t / b = the group's bytes of the two source rows as words, out = G destination pixels
template <int CN, int G = BoxCfg<CN>::G>
__device__ __forceinline__ void box_words(const uint32_t* t, const uint32_t* b, uint32_t* out)
{
    constexpr int NW = G * 2 * CN / 4, OUTB = G * CN;
    // vertical sums as packed 16-bit lanes: ve[i] = bytes (4i, 4i+2), vo[i] = bytes (4i+1, 4i+3)
    uint32_t ve[NW], vo[NW];
#pragma unroll
    for (int i = 0; i < NW; i++) {
        ve[i] = __byte_perm(t[i], 0, 0x4240) + __byte_perm(b[i], 0, 0x4240);
        vo[i] = __byte_perm(t[i], 0, 0x4341) + __byte_perm(b[i], 0, 0x4341);
    }
    // lanes(p, q) = the vertical sums of source bytes p and q as one pair of 16-bit lanes (indices are compile-time after
    // unrolling: an adjacent pair of one word costs nothing, anything else one PRMT)
    auto lanes = [&](int p, int q) -> uint32_t {
        const uint32_t wp = (p & 1) ? vo[p >> 2] : ve[p >> 2], wq = (q & 1) ? vo[q >> 2] : ve[q >> 2];
        const int lp = (p >> 1) & 1, lq = (q >> 1) & 1;
        if ((p & 1) == (q & 1) && (p >> 2) == (q >> 2) && lp == 0 && lq == 1) return wp;
        return __byte_perm(wp, wq, (2 * lp) | ((2 * lp + 1) << 4) | ((4 + 2 * lq) << 8) | ((5 + 2 * lq) << 12));
    };
    auto kof = [](int ob) { return 2 * (ob / CN) * CN + ob % CN; };       // output byte -> its left source byte
#pragma unroll
    for (int wi = 0; wi < OUTB / 4; wi++) {
        // output bytes (0, 2) and (1, 3) of this word as lane pairs: left + right + 2, >> 2 on both lanes at once (a lane's
        // sum is < 1024, so the bits the shift drags in from the upper lane land above the byte that is kept)
        const int k0 = kof(4 * wi), k1 = kof(4 * wi + 1), k2 = kof(4 * wi + 2), k3 = kof(4 * wi + 3);
        const uint32_t x = (lanes(k0, k2) + lanes(k0 + CN, k2 + CN) + 0x00020002u) >> 2;
        const uint32_t y = (lanes(k1, k3) + lanes(k1 + CN, k3 + CN) + 0x00020002u) >> 2;
        out[wi] = __byte_perm(x, y, 0x6240);
    }
}

template <int CN, bool STREAM>
__device__ __forceinline__ void box_store(uint8_t* d, const uint32_t* out)
{
    if constexpr (CN == 3) {
        uint2* d2 = reinterpret_cast<uint2*>(d);
        if (STREAM) { stg_cs(d2, make_uint2(out[0], out[1])); stg_cs(d2 + 1, make_uint2(out[2], out[3])); stg_cs(d2 + 2, make_uint2(out[4], out[5])); }
        else { d2[0] = make_uint2(out[0], out[1]); d2[1] = make_uint2(out[2], out[3]); d2[2] = make_uint2(out[4], out[5]); }
    } else {
        if (STREAM) stg_cs(reinterpret_cast<uint4*>(d), make_uint4(out[0], out[1], out[2], out[3]));
        else *reinterpret_cast<uint4*>(d) = make_uint4(out[0], out[1], out[2], out[3]);
    }
}

// destination pixels [x_from, dst_w) of one row, one at a time (ragged right edge)
template <int CN>
__device__ __forceinline__ void box_tail(const uint8_t* __restrict__ s0row, const uint8_t* __restrict__ s1row, uint8_t* __restrict__ drow,
                                         int x_from, int dst_w)
{
    for (int x = x_from; x < dst_w; x++)
#pragma unroll
        for (int c = 0; c < CN; c++) {
            const int k = 2 * x * CN + c;
            drow[x * CN + c] = (uint8_t)((s0row[k] + s0row[k + CN] + s1row[k] + s1row[k + CN] + 2) >> 2);
        }
}

// one group of G destination pixels of one row: r0/r1 = the two source rows, drow = the destination row
template <int CN, bool STREAM>
__device__ __forceinline__ void box_group(const uint8_t* __restrict__ s0row, const uint8_t* __restrict__ s1row,
                                          uint8_t* __restrict__ drow, int g, int dst_w)
{
    constexpr int G = BoxCfg<CN>::G, NV = BoxCfg<CN>::NV, NW = NV * 4, OUTB = G * CN;
    const int px0 = g * G;
    if (px0 + G <= dst_w) {
        const uint4* r0 = reinterpret_cast<const uint4*>(s0row + (long long)px0 * 2 * CN);
        const uint4* r1 = reinterpret_cast<const uint4*>(s1row + (long long)px0 * 2 * CN);
        uint32_t t[NW], b[NW];
#pragma unroll
        for (int i = 0; i < NV; i++) {
            uint4 u = ldg_stream(r0 + i), v = ldg_stream(r1 + i);
            t[4 * i] = u.x; t[4 * i + 1] = u.y; t[4 * i + 2] = u.z; t[4 * i + 3] = u.w;
            b[4 * i] = v.x; b[4 * i + 1] = v.y; b[4 * i + 2] = v.z; b[4 * i + 3] = v.w;
        }
        uint32_t out[OUTB / 4];
        box_words<CN>(t, b, out);
        box_store<CN, STREAM>(drow + (long long)px0 * CN, out);
    } else box_tail<CN>(s0row, s1row, drow, px0, dst_w);   // right-edge tail group
}

// half a group (G/2 destination pixels) with its source bytes staged in shared memory (s0 / s1 = its first byte in each staged
// row): 24 (3 channels) or 16 bytes per row and thread -- 64-bit loads at a 24-byte lane pitch and 128-bit loads at a 16-byte
// pitch are both conflict-free -- and half the registers of a whole group, which the pipeline kernel's 32-register cap needs
template <int CN>
__device__ __forceinline__ void box_half_smem(const uint8_t* s0, const uint8_t* s1, uint8_t* d)
{
    constexpr int HG = BoxCfg<CN>::G / 2, NW = HG * 2 * CN / 4, OW = HG * CN / 4;
    uint32_t t[NW], b[NW], out[OW];
    if constexpr (CN == 3) {
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const uint2 u = reinterpret_cast<const uint2*>(s0)[i], v = reinterpret_cast<const uint2*>(s1)[i];
            t[2 * i] = u.x; t[2 * i + 1] = u.y; b[2 * i] = v.x; b[2 * i + 1] = v.y;
        }
    } else {
        const uint4 u = *reinterpret_cast<const uint4*>(s0), v = *reinterpret_cast<const uint4*>(s1);
        t[0] = u.x; t[1] = u.y; t[2] = u.z; t[3] = u.w; b[0] = v.x; b[1] = v.y; b[2] = v.z; b[3] = v.w;
    }
    box_words<CN, HG>(t, b, out);
    if constexpr (CN == 3) { uint32_t* dw = reinterpret_cast<uint32_t*>(d); dw[0] = out[0]; dw[1] = out[1]; dw[2] = out[2]; }
    else *reinterpret_cast<uint2*>(d) = make_uint2(out[0], out[1]);
}

template <int CN, bool STREAM>
__global__ void __launch_bounds__(128) box2x2_u8_kernel(const ResizeArgs a, int groups_per_row)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (g >= groups_per_row) return;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride + (long long)(2 * y) * a.src.step;
    uint8_t* drow = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y * a.dst.step;
    box_group<CN, STREAM>(src, src + a.src.step, drow, g, a.dst.w);
}

bool box2x2_u8_applicable(const ResizeArgs& a, int cn)
{
    if (!(cn == 1 || cn == 3 || cn == 4)) return false;
    if (a.src.w != 2 * a.dst.w || a.src.h != 2 * a.dst.h) return false;           // no ragged edges here
    if (((uintptr_t)a.src.p & 15) || (a.src.step & 15) || (a.src_frame_stride & 15)) return false;
    const int da = cn == 3 ? 8 : 16;
    if (((uintptr_t)a.dst.p % da) || (a.dst.step % da) || (a.dst_frame_stride % da)) return false;
    return true;
}

int launch_box2x2_u8(const ResizeArgs& a, int cn, int n_frames, cudaStream_t s)
{
    const int G = cn == 1 ? 16 : cn == 3 ? 8 : 4;
    const int groups = (a.dst.w + G - 1) / G;
    dim3 block(128), grid((groups + 127) / 128, a.dst.h, n_frames);
    // outputs nobody re-reads soon are written with evict-first (.cs) stores: measured on B200, write-back
    // cached stores leave ~100 MB of dirty lines in L2 per launch and cut steady-state DRAM throughput of this
    // kernel from 6.5 TB/s to 3.9 TB/s (profiles/r1_notes.md)
#define B200_BOX(CNV) (a.keep_in_l2 ? box2x2_u8_kernel<CNV, false><<<grid, block, 0, s>>>(a, groups) : box2x2_u8_kernel<CNV, true><<<grid, block, 0, s>>>(a, groups))
    if (cn == 1) B200_BOX(1); else if (cn == 3) B200_BOX(3); else B200_BOX(4);
#undef B200_BOX
    return (int)cudaGetLastError();
}

// =========================================================================== kx x ky box, 8U =====
// Integer-scale INTER_AREA other than the 2x2 special case (resize.cpp:3010-3026): D = saturate_cast<uchar>(sum * (1.f/area))
// with an exact integer window sum.  One thread = 4 destination pixels of one row: per source row it loads KX*CN aligned
// 32-bit words (the generic kernel issues KX*KY*CN byte loads per pixel), keeps vertical sums in packed 16-bit lanes, then
// adds KX lanes per output and applies the reference's single fp32 multiply + round-half-even.
template <int CN, int KX>
__global__ void __launch_bounds__(128) area_int_u8_kernel(const ResizeArgs a, int groups_per_row)
{
    constexpr int NWORD = KX * CN;                 // words per source row for 4 destination pixels
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (g >= groups_per_row) return;
    const int ky = a.iscale_y;
    const uint8_t* srow = a.src.p + (long long)blockIdx.z * a.src_frame_stride + (long long)(y * ky) * a.src.step;
    uint8_t* drow = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y * a.dst.step;
    const float scale = 1.f / (float)(KX * ky);    // resize.cpp:2969
    const int px0 = g * 4;
    if (px0 + 4 <= a.dst.w) {
        uint32_t ve[NWORD], vo[NWORD];
#pragma unroll
        for (int i = 0; i < NWORD; i++) ve[i] = vo[i] = 0;
        const uint32_t* p = reinterpret_cast<const uint32_t*>(srow + (long long)px0 * KX * CN);
        for (int r = 0; r < ky; r++) {
#pragma unroll
            for (int i = 0; i < NWORD; i++) {
                const uint32_t w = __ldg(p + i);
                ve[i] += __byte_perm(w, 0, 0x4240);
                vo[i] += __byte_perm(w, 0, 0x4341);
            }
            p = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(p) + a.src.step);
        }
        auto vs = [&](int k) -> uint32_t {
            const uint32_t w = (k & 1) ? vo[k >> 2] : ve[k >> 2];
            return (k & 2) ? (w >> 16) : (w & 0xFFFFu);
        };
        uint32_t out[CN];
#pragma unroll
        for (int wi = 0; wi < CN; wi++) {
            uint32_t word = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int ob = wi * 4 + j, q = ob / CN, c = ob % CN;
                uint32_t sum = 0;
#pragma unroll
                for (int t = 0; t < KX; t++) sum += vs((q * KX + t) * CN + c);
                word |= (uint32_t)sat_u8(rne_f(fmul(u2f(sum), scale))) << (8 * j);
            }
            out[wi] = word;
        }
        uint32_t* d = reinterpret_cast<uint32_t*>(drow + (long long)px0 * CN);
#pragma unroll
        for (int wi = 0; wi < CN; wi++) __stcs(d + wi, out[wi]);
    } else {                                       // right-edge tail group: per pixel
        for (int x = px0; x < a.dst.w; x++)
#pragma unroll
            for (int c = 0; c < CN; c++) {
                int sum = 0;
                for (int r = 0; r < ky; r++)
                    for (int t = 0; t < KX; t++) sum += srow[(long long)r * a.src.step + (x * KX + t) * CN + c];
                drow[x * CN + c] = (uint8_t)sat_u8(rne_f(fmul((float)sum, scale)));
            }
    }
}

bool area_int_u8_applicable(const ResizeArgs& a, int cn)
{
    if (!(cn == 1 || cn == 3 || cn == 4)) return false;
    if (a.iscale_x < 2 || a.iscale_x > 4 || a.iscale_y < 1 || a.iscale_y > 255) return false;
    if (a.iscale_x == 2 && a.iscale_y == 2) return false;                        // the (a+b+c+d+2)>>2 special case is not this formula
    if (a.src.w != a.iscale_x * a.dst.w || a.src.h != a.iscale_y * a.dst.h) return false;   // no ragged edges here
    if (((uintptr_t)a.src.p & 3) || (a.src.step & 3) || (a.src_frame_stride & 3)) return false;
    if (((uintptr_t)a.dst.p & 3) || (a.dst.step & 3) || (a.dst_frame_stride & 3)) return false;
    return true;
}

int launch_area_int_u8(const ResizeArgs& a, int cn, int n_frames, cudaStream_t s)
{
    const int groups = (a.dst.w + 3) / 4;
    dim3 block(128), grid((groups + 127) / 128, a.dst.h, n_frames);
#define B200_AREA(CNV, KXV) if (cn == CNV && a.iscale_x == KXV) area_int_u8_kernel<CNV, KXV><<<grid, block, 0, s>>>(a, groups);
    B200_AREA(1, 2) B200_AREA(1, 3) B200_AREA(1, 4)
    B200_AREA(3, 2) B200_AREA(3, 3) B200_AREA(3, 4)
    B200_AREA(4, 2) B200_AREA(4, 3) B200_AREA(4, 4)
#undef B200_AREA
    return (int)cudaGetLastError();
}

// =========================================================================== warpAffine via TMA ===
namespace {

struct AffineFastArgs {
    double M[6];
    const uint8_t* src; int src_step; long long src_frame_stride;
    int sw, sh, dw, dh;
    int boxw, boxh;            // staged box: bytes per row (multiple of 16), rows
    int x_shift;               // source tensor map: 0 = byte elements, 2 = 32-bit elements (boxes wider than 256 bytes)
    int border;
    int keep_in_l2;            // 0: evict-first store policy for the output
    int patch;                 // 0: TMA zero fill is already the border value (BORDER_CONSTANT, value 0)
    uint32_t cval;             // 4 packed border bytes
    uint8_t* dst; int dst_step; long long dst_frame_stride;   // destination frames (DIRECT stores only; the tensor map otherwise)
    int exp;                   // experiments (make TUNING=1 only): 1 = no TMA load, 2 = no ready wait, 4 = no interpolation, 8 = no box work, 16 = no TMA store
};

// one destination pixel: Xf/Yf = X0[y]+adelta[x] / Y0[y]+bdelta[x] (AB_BITS=10 fixed point, round_delta included);
// taps from the staged box at byte column bxb / row by0; exact integer blend (see Blend<uint8_t> in warp_kernels.cu)
// W64 (3 channels only): the tap bytes of a row come from two 64-bit loads + selects (2 conflict-free wavefronts each at a
// good pitch) instead of three 32-bit loads (fewer instructions, but ncu counts 2 wavefronts per LDS.32 whatever the pitch:
// 32 lanes along a line of slope > 1 word per lane cannot hit 32 distinct banks).  Measured on B200, 4K 8UC3 x 8 frames:
// stand-alone kernel 128 us (W64) vs 135 us; inside the frame pipeline 157 us (W64) vs 151 us.
template <int CN, bool W64>
__device__ __forceinline__ void warp_px_xy(const uint8_t* __restrict__ tile, int boxw, int X, int Y, int bxb, int by0,
                                           uint8_t* __restrict__ D);
template <int CN, bool W64>
__device__ __forceinline__ void warp_px(const uint8_t* __restrict__ tile, int boxw, int Xf, int Yf, int bxb, int by0,
                                        uint8_t* __restrict__ D)
{
    warp_px_xy<CN, W64>(tile, boxw, Xf >> 5, Yf >> 5, bxb, by0, D);      // imgwarp.cpp:2351-2358
}
// X, Y: source coordinate in 1/32 pixel units (INTER_BITS = 5)
template <int CN, bool W64>
__device__ __forceinline__ void warp_px_xy(const uint8_t* __restrict__ tile, int boxw, int X, int Y, int bxb, int by0,
                                           uint8_t* __restrict__ D)
{
    const uint32_t fx = X & 31, fy = Y & 31;
    const int ob = (X >> 5) * CN - bxb, oy = (Y >> 5) - by0;            // byte offset of the left tap in its row
    const int o = oy * boxw + ob;
    // the four 10-bit tap weights as two 16-bit pairs (top | bottom << 16): WL = (32-fx)*{32-fy, fy}, WR = fx*{32-fy, fy}.
    // One IDP.2A multiplies a (top, bottom) byte pair by such a pair and accumulates, so a channel costs two IDP.2A:
    //   out = (wTL*tl + wBL*bl + wTR*tr + wBR*br + 512) >> 10        (remapBilinear 8U: the 15-bit table = 32 * these)
    const uint32_t ayp = fy * 0xFFFFu + 32u;
    const uint32_t WR = fx * ayp, WL = (ayp << 5) - WR;
    if constexpr (CN == 3) {
        // the 6 tap bytes of a row start at byte (o & 7) of the aligned 16-byte window: two LDS.64 per row.  A 64-bit
        // shared load is served half a warp at a time, and 16 lanes along a slanted source line touch few enough
        // 8-byte words to be conflict-free at a good pitch -- three LDS.32 per row needed 1.9 wavefronts each (ncu)
        uint32_t tw0, tw1, tw2, bw0, bw1, bw2;
        if constexpr (W64) {
            const uint2* wt = reinterpret_cast<const uint2*>(tile + (o & ~7));
            const uint2* wb = reinterpret_cast<const uint2*>(tile + (o & ~7) + boxw);
            const uint2 t0 = wt[0], t1 = wt[1], b0 = wb[0], b1 = wb[1];
            const bool up = (o & 4) != 0;
            tw0 = up ? t0.y : t0.x; tw1 = up ? t1.x : t0.y; tw2 = up ? t1.y : t1.x;
            bw0 = up ? b0.y : b0.x; bw1 = up ? b1.x : b0.y; bw2 = up ? b1.y : b1.x;
        } else {
            // three aligned words per row (the 6 tap bytes start at byte o & 3 of the first)
            const uint32_t* wt = reinterpret_cast<const uint32_t*>(tile + (o & ~3));
            const uint32_t* wb = reinterpret_cast<const uint32_t*>(tile + (o & ~3) + boxw);
            tw0 = wt[0]; tw1 = wt[1]; tw2 = wt[2]; bw0 = wb[0]; bw1 = wb[1]; bw2 = wb[2];
        }
        const int sh = o << 3;                                           // funnel shifts use the low 5 bits: (o & 3) * 8
        const uint32_t tL = __funnelshift_r(tw0, tw1, sh), tH = __funnelshift_r(tw1, tw2, sh);
        const uint32_t bL = __funnelshift_r(bw0, bw1, sh), bH = __funnelshift_r(bw1, bw2, sh);
        // tap bytes: left pixel = bytes 0..2, right pixel = bytes 3..5 of the (tL,tH) / (bL,bH) windows
        const uint32_t p01 = __byte_perm(tL, bL, 0x5140);                // (t0,b0 | t1,b1): left taps of c0, c1
        const uint32_t p23 = __byte_perm(tL, bL, 0x7362);                // (t2,b2 | t3,b3): left tap of c2, right tap of c0
        const uint32_t p45 = __byte_perm(tH, bH, 0x5140);                // (t4,b4 | t5,b5): right taps of c1, c2
        const uint32_t s0 = __dp2a_hi(WR, p23, __dp2a_lo(WL, p01, 512u));
        const uint32_t s1 = __dp2a_lo(WR, p45, __dp2a_hi(WL, p01, 512u));
        const uint32_t s2 = __dp2a_hi(WR, p45, __dp2a_lo(WL, p23, 512u));
        D[0] = (uint8_t)(s0 >> 10); D[1] = (uint8_t)(s1 >> 10); D[2] = (uint8_t)(s2 >> 10);
    } else if constexpr (CN == 4) {
        const uint32_t* wt = reinterpret_cast<const uint32_t*>(tile + o);
        const uint32_t* wb = reinterpret_cast<const uint32_t*>(tile + o + boxw);
        const uint32_t p0 = wt[0], p1 = wt[1], q0 = wb[0], q1 = wb[1];
        const uint32_t l01 = __byte_perm(p0, q0, 0x5140), l23 = __byte_perm(p0, q0, 0x7362);   // left taps (top,bottom) of c0,c1 | c2,c3
        const uint32_t r01 = __byte_perm(p1, q1, 0x5140), r23 = __byte_perm(p1, q1, 0x7362);   // right taps
        const uint32_t s0 = __dp2a_lo(WR, r01, __dp2a_lo(WL, l01, 512u));
        const uint32_t s1 = __dp2a_hi(WR, r01, __dp2a_hi(WL, l01, 512u));
        const uint32_t s2 = __dp2a_lo(WR, r23, __dp2a_lo(WL, l23, 512u));
        const uint32_t s3 = __dp2a_hi(WR, r23, __dp2a_hi(WL, l23, 512u));
        *reinterpret_cast<uint32_t*>(D) = (s0 >> 10) | ((s1 >> 10) << 8) | ((s2 >> 10) << 16) | ((s3 >> 10) << 24);
    } else {
        const uint32_t* wt = reinterpret_cast<const uint32_t*>(tile + (o & ~3));
        const uint32_t* wb = reinterpret_cast<const uint32_t*>(tile + (o & ~3) + boxw);
        const int sh = o << 3;
        const uint32_t tL = __funnelshift_r(wt[0], wt[1], sh), bL = __funnelshift_r(wb[0], wb[1], sh);
        const uint32_t p = __byte_perm(tL, bL, 0x5140);                  // (t0,b0 | t1,b1)
        const uint32_t s0 = __dp2a_hi(WR, p, __dp2a_lo(WL, p, 512u));
        D[0] = (uint8_t)(s0 >> 10);
    }
}


// NT threads per CTA; 32 registers per thread so that 2048 threads per SM stay resident and hide each other's TMA wait
// One tile: (x0, y0) of destination frame `frame_dst`, source = frame `frame_src` of the source tensor.  `phase` = parity of
// the CTA's mbarrier (one use per tile).  When `ready` is given, thread 0 waits until *ready >= ready_count before it reads
// the source (frames produced earlier in the same launch), and `consumed` is incremented once the tile no longer needs it.
// A staged box that crosses the image edge: TMA zero-filled the outside; write the border pixels there.
// BORDER_CONSTANT -> cval; others -> src[borderInterpolate(y)][borderInterpolate(x)] (copy.cpp:905-950)
template <int CN, int NT, class Args>
__device__ __forceinline__ void patch_border(uint8_t* tile, const Args& a, int bxb, int by0, int frame)
{
    const uint8_t* src = a.src + (long long)frame * a.src_frame_stride;
    const int boxw = a.boxw, n = boxw * a.boxh;
    for (int i = threadIdx.x; i < n; i += NT) {
        const int row = i / boxw, col = i - row * boxw;
        const int gy = by0 + row, gb = bxb + col;
        const int px = (gb >= 0) ? gb / CN : -((-gb + CN - 1) / CN);
        const int c = gb - px * CN;
        if ((unsigned)gy < (unsigned)a.sh && (unsigned)px < (unsigned)a.sw) continue;
        uint8_t v;
        if (a.border == B_CONSTANT) v = (uint8_t)(a.cval >> (8 * c));
        else {
            const int yy = border_idx(gy, a.sh, a.border), xx = border_idx(px, a.sw, a.border);
            v = src[(long long)yy * a.src_step + xx * CN + c];
        }
        tile[i] = v;
    }
}

// 4-channel tiles of 4096 pixels write their result in two halves (see warp_tile); host and device must agree on it.  (Measured: 8UC4
// 127 -> 119 us at 4K x 8 frames; 8UC3 and 8UC1 gain nothing from the extra CTA per SM and pay the second barrier: 130-132 vs 129 us.)
__host__ __device__ constexpr int out_phases(int tw, int th, int cn) { return (tw * th >= 4096 && cn == 4) ? 2 : 1; }

// `after_load()` is run by thread 0 once the tile no longer reads the source frame from global memory.
struct NoOp { __device__ __forceinline__ void operator()() const {} };
// `after_issue()` is run by thread 0 between issuing the tile's load and waiting for it (time it would spend idle).
// DIRECT: results go straight to global memory (three byte stores per pixel for 3 channels) instead of through an output tile in
// shared memory and a bulk store: no tile-sized shared buffer, no barrier and no store drain at the end of a tile.
template <int CN, int TW, int TH, int NT, bool W64, class AfterLoad = NoOp, int BAR = 0, class AfterIssue = NoOp, bool DIRECT = false>
__device__ __forceinline__ void warp_tile(uint8_t* smem, const CUtensorMap& smap, const CUtensorMap& dmap, const AffineFastArgs& a,
                                          int x0, int y0, int frame_src, int frame_dst, uint32_t phase,
                                          const int* ready, int ready_count, AfterLoad after_load = AfterLoad(), uint64_t* bar_at = nullptr,
                                          AfterIssue after_issue = AfterIssue())
{
    // 4-channel 4096-pixel tiles go out in two halves through an output tile of half the height: 40 KB instead of 48 KB of shared
    // memory, 5 CTAs per SM instead of 4
    constexpr int PH = DIRECT ? 1 : out_phases(TW, TH, CN), OTH = TH / PH;
    constexpr int OUT_BYTES = DIRECT ? 0 : TW * CN * OTH;
    constexpr int ROWS_PER_PASS = NT / TW;
    constexpr int PASSES = OTH / ROWS_PER_PASS;                      // rows a thread owns per half
    uint8_t* out_tile = smem;                                        // [OTH][TW*CN]
    int* adelta = reinterpret_cast<int*>(smem + OUT_BYTES);          // [TW]
    int* bdelta = adelta + TW;
    int2* XY0 = reinterpret_cast<int2*>(bdelta + TW);                // [TH] (X0, Y0) per tile row
    int* org = reinterpret_cast<int*>(XY0 + TH);                     // box origin, crosses-edge flag
    uint64_t* bar = bar_at ? bar_at : reinterpret_cast<uint64_t*>(org + 4);   // the pipeline keeps it outside the region its strips reuse
    static_assert(PASSES % 2 == 0 && (TW * CN * OTH) % 16 == 0 && TW % 2 == 0, "table alignment");
    const long long ostride = DIRECT ? (long long)a.dst_step : (long long)(TW * CN);
    uint8_t* tile = smem + ((OUT_BYTES + (2 * TW + 2 * TH + 4) * 4 + 8 + 127) & ~127);   // [boxh][boxw] (+16 slack)

    const int tid = threadIdx.x;
    // the first look at the producer's counter is issued here and used after the coordinate terms: its L2 round trip (the strips
    // are normally long finished) overlaps them instead of preceding the tile's load
    int ready_seen = 0;
    if (ready && tid == 0) ready_seen = ld_acquire(ready);
    // per-tile coordinate terms, fp64, separately rounded (imgwarp.cpp:2266-2270, :2030-2031)
    if (tid < TW) {
        const double x = (double)(x0 + tid);
        adelta[tid] = rne_d(__dmul_rn(__dmul_rn(a.M[0], x), 1024.0));
        bdelta[tid] = rne_d(__dmul_rn(__dmul_rn(a.M[3], x), 1024.0));
    } else if (tid < TW + TH) {
        const int r = tid - TW;
        const double y = (double)(y0 + r);
        XY0[r] = make_int2(rne_d(__dmul_rn(__dadd_rn(__dmul_rn(a.M[1], y), a.M[2]), 1024.0)) + 16,
                           rne_d(__dmul_rn(__dadd_rn(__dmul_rn(a.M[4], y), a.M[5]), 1024.0)) + 16);
    }
    cta_sync<BAR, NT>();
    if (tid == 0) {
        // box origin = exact minimum over the (clipped) tile: both terms are monotone in their index
        const int cl = min(TW, a.dw - x0) - 1, rl = min(TH, a.dh - y0) - 1;
        const int bx0 = (min(XY0[0].x, XY0[rl].x) + min(adelta[0], adelta[cl])) >> 10;
        const int by0 = (min(XY0[0].y, XY0[rl].y) + min(bdelta[0], bdelta[cl])) >> 10;
        // measured on B200: the innermost TMA coordinate must be a multiple of 16 bytes (an unaligned one
        // raises "illegal instruction"), so the box starts at the 16-byte boundary at or below the left tap
        const int bxb = (bx0 * CN) & ~15;
        org[0] = bxb; org[1] = by0;
        org[2] = (bxb < 0 || by0 < 0 || bxb + a.boxw > a.sw * CN || by0 + a.boxh > a.sh);
        if (ready) {
            while (ready_seen < ready_count) { __nanosleep(100); ready_seen = ld_acquire(ready); }
            asm volatile("fence.proxy.async.global;" ::: "memory");     // generic-proxy stores of other CTAs -> this TMA read
        }
#ifdef B200_TUNING
        if (a.exp & 1) mbar_expect_tx(bar, 0); else
#endif
        {
        mbar_expect_tx(bar, (uint32_t)(a.boxw * a.boxh));
        tma_load_3d(tile, &smap, bar, bxb >> a.x_shift, by0, frame_src);
        }
        after_issue();
    }
    // only warp 0 polls the mbarrier; the other warps sleep at the CTA barrier instead of spinning
    if (tid < 32) mbar_wait(bar, phase);
    cta_sync<BAR, NT>();
    const int bxb = org[0], by0 = org[1];          // box origin: byte column (16-aligned), row
    const int boxw = a.boxw;

    if (a.patch && org[2]) {
        patch_border<CN, NT>(tile, a, bxb, by0, frame_src);
        cta_sync<BAR, NT>();
    }

    if (tid == 0) after_load();
    const int xl = tid % TW, rg = tid / TW;
    const int ad = adelta[xl], bd = bdelta[xl];
    const bool full = (x0 + TW <= a.dw) && (y0 + TH <= a.dh);
    uint8_t* const Dbase = DIRECT ? a.dst + (long long)frame_dst * a.dst_frame_stride + (long long)(y0 + rg * PASSES) * a.dst_step + (long long)(x0 + xl) * CN
                                  : out_tile + (rg * PASSES) * (TW * CN) + xl * CN;
#pragma unroll 1
    for (int ph = 0; ph < PH; ph++) {
        const int row0 = ph * OTH + rg * PASSES;                        // this thread's rows of this half: row0 .. row0 + PASSES - 1
        const int2* const xyrow = XY0 + row0;
#ifdef B200_TUNING
        if (a.exp & 4) {} else
#endif
        if (full) {                                                     // interior tile: no per-pixel bounds checks
            // (fetching the taps of 4 pixels before blending the first, with 64 registers, changed nothing: 129 us either way)
#pragma unroll
            for (int p = 0; p < PASSES; p += 2) {
                const int4 v = *reinterpret_cast<const int4*>(xyrow + p);
                warp_px<CN, W64>(tile, boxw, v.x + ad, v.y + bd, bxb, by0, Dbase + p * ostride);
                warp_px<CN, W64>(tile, boxw, v.z + ad, v.w + bd, bxb, by0, Dbase + (p + 1) * ostride);
            }
        } else if (x0 + xl < a.dw) {
#pragma unroll 2
            for (int p = 0; p < PASSES; p++) {
                if (y0 + row0 + p >= a.dh) break;
                const int2 v = xyrow[p];
                warp_px<CN, W64>(tile, boxw, v.x + ad, v.y + bd, bxb, by0, Dbase + p * ostride);
            }
        }
        if constexpr (DIRECT) return;
        // generic-proxy writes to out_tile must be visible to the async proxy before the bulk store
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        cta_sync<BAR, NT>();
#ifdef B200_TUNING
        if (a.exp & 16) return;
#endif
        // (rows below the image are clipped by the tensor map; tma_store_3d returns once the store has read the tile)
        if (tid == 0 && y0 + ph * OTH < a.dh) tma_store_3d(&dmap, out_tile, x0 * CN, y0 + ph * OTH, frame_dst, a.keep_in_l2 != 0);
        if (ph + 1 < PH) cta_sync<BAR, NT>();                           // the output tile is free for the second half
    }
}

template <int CN, int TW, int TH> struct FastLayout {
    static constexpr int OUT_BYTES = TW * CN * (TH / out_phases(TW, TH, CN));
    static constexpr int BAR_OFF = OUT_BYTES + (2 * TW + 2 * TH + 4) * 4;
};

template <int CN, int TW, int TH, int NT>
__global__ void __launch_bounds__(NT, 2048 / NT)
warp_affine_tma_u8(const __grid_constant__ CUtensorMap smap, const __grid_constant__ CUtensorMap dmap, const AffineFastArgs a)
{
    extern __shared__ __align__(128) uint8_t smem[];
    if (threadIdx.x == 0) mbar_init(reinterpret_cast<uint64_t*>(smem + FastLayout<CN, TW, TH>::BAR_OFF), 1);
    warp_tile<CN, TW, TH, NT, true>(smem, smap, dmap, a, blockIdx.x * TW, blockIdx.y * TH, blockIdx.z, blockIdx.z, 0, nullptr, 0);
}

// ================================================================= resize 2x2 -> warpAffine, one launch ==========
// The frame-stream pipeline as ONE persistent kernel.  CTAs draw work items from a queue in a fixed order --
// for every frame first its box-filter strips (SR rows of the half-size frame each), then its warpAffine tiles.  A tile
// waits until all strips of its frame are done (they precede it in the queue and never wait on a tile of the same or a
// later frame, so the wait terminates); the half-size frames live in a ring of R frames that stays in L2: a strip of
// frame f waits until every tile of frame f-R has read its box.  Compared with two batched launches the intermediate
// frames never travel to HBM and back, and the DRAM-bound strips overlap the issue-bound tiles on every SM.
struct PipeArgs {
    const uint8_t* src; long long src_frame_stride; int src_step;        // full-size frames
    uint8_t* mid; long long mid_frame_stride; int mid_step;              // ring of R half-size frames
    int mw, mh, ring;
    int n_frames, strips, strip_rows, ntx, nty, ahead;
    int full_groups, chunk_groups, chunks, stage_bytes;   // strips: staged row pairs (see the kernel)
    int* counters;             // [0] queue head, [1 + f] strips done of frame f, [1 + n + f] tiles that consumed frame f
};

enum { W_STRIP = 0, W_TILE = 1, W_EMPTY = 2, W_END = 3 };
// thread 0: draw the next work item and decode it into {kind | ring slot << 4, frame, strip index | x0, y0}
template <int TW, int TH>
__device__ __forceinline__ void pipe_fetch(const PipeArgs& p, int4* w)
{
    const int tiles = p.ntx * p.nty;
    const int per_frame = p.strips + tiles, head = p.ahead * p.strips;
    const int item = atomicAdd(p.counters, 1);
    int f, r;
    if (item >= head + per_frame * p.n_frames) { *w = make_int4(W_END, 0, 0, 0); return; }
    if (item < head) { f = item / p.strips; r = item - f * p.strips; }
    else {
        const int b = (item - head) / per_frame;
        r = (item - head) - b * per_frame;
        f = r < p.strips ? b + p.ahead : b;
    }
    if (f >= p.n_frames) { *w = make_int4(W_EMPTY, 0, 0, 0); return; }      // strips beyond the last frame
    const int slot = f % p.ring;
    if (r < p.strips) *w = make_int4(W_STRIP | (slot << 4), f, r, 0);
    else { const int t = r - p.strips, ty = t / p.ntx; *w = make_int4(W_TILE | (slot << 4), f, (t - ty * p.ntx) * TW, ty * TH); }
}

template <int CN, int TW, int TH, int NT, bool DIRECT = false>
__global__ void __launch_bounds__(NT, (CN == 1 ? 1536 : 2048) / NT)      // CN 1: 16-pixel box groups need 40 registers
pipe_box_warp_u8(const __grid_constant__ CUtensorMap smap, const __grid_constant__ CUtensorMap dmap, const AffineFastArgs a,
                 const __grid_constant__ PipeArgs p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ int4 s_work[2];                                    // {kind | ring slot << 4, frame, strip index | x0, y0}
    constexpr int G = BoxCfg<CN>::G;
    const int tid = threadIdx.x;
    __shared__ __align__(8) uint64_t s_sbar[3];                   // strips: one mbarrier per staged row pair; [2]: the tiles' box
    if (tid == 0) { mbar_init(&s_sbar[0], 1); mbar_init(&s_sbar[1], 1); mbar_init(&s_sbar[2], 1); }
    uint32_t sphase = 0;                                          // bit s: parity of stage s (the same in every thread)
    const int tiles = p.ntx * p.nty;
    int* strips_done = p.counters + 1;
    int* consumed = p.counters + 1 + p.n_frames;
    const int groups = (p.mw + G - 1) / G;
    uint32_t phase = 0;
    // Thread 0 draws the next item from the queue (an L2 atomic, ~1 us round trip) in the middle of the current one and
    // decodes it for the CTA.  Drawing it at the start held strips back behind a whole tile and cost more than it saved.
    // Queue order: strips of the first `ahead` frames, then per frame f: strips of frame f + ahead, tiles of frame f.
    // The strips of a frame are drawn `ahead` frames before its tiles, so a tile normally finds them finished; with the
    // plain order (strips f, tiles f) every CTA stalled at each frame boundary for the latency of the last strips.
    auto fetch = [&](int4* w) { pipe_fetch<TW, TH>(p, w); };
    if (tid == 0) fetch(&s_work[0]);
    for (int round = 0;; round++) {
        __syncthreads();                                          // s_work[round & 1] is written; the previous round's reads are over
        const int4 w = s_work[round & 1];
        int4* const next = &s_work[(round + 1) & 1];
        const int kind = w.x & 7, slot = w.x >> 4, f = w.y;
        if (kind == W_END) break;
        if (kind == W_EMPTY) { if (tid == 0) fetch(next); continue; }
        if (kind == W_STRIP) {
            // ---- box-filter strip w.z of frame f into its ring slot.  The source rows arrive by bulk copies (one thread issues
            // them, two row pairs in flight per CTA whatever the register budget is): with per-thread 16-byte loads under the
            // 32-register cap the strips moved 4.7 TB/s where the stand-alone box kernel moves 6.5 (profiles/r2_notes.md #7).
            const uint8_t* sf = p.src + (long long)f * p.src_frame_stride;
            uint8_t* mf = p.mid + (long long)slot * p.mid_frame_stride;
            const int y_lo = w.z * p.strip_rows, rows = min(p.strip_rows, p.mh - y_lo);
            const int units = rows * p.chunks;                    // unit = one chunk of one destination row = 2 staged source row pieces
            constexpr int GB = G * 2 * CN;                        // source bytes per group and row
            auto issue = [&](int u) {                             // thread 0
                const int r = u / p.chunks, c = u - r * p.chunks;
                const int g0 = c * p.chunk_groups, ng = min(p.chunk_groups, p.full_groups - g0);
                uint8_t* st = smem + (u & 1) * p.stage_bytes;
                const uint8_t* s0 = sf + (long long)(2 * (y_lo + r)) * p.src_step + (long long)g0 * GB;
                mbar_expect_tx(&s_sbar[u & 1], 2u * ng * GB);
                bulk_load_1d_stream(st, s0, ng * GB, &s_sbar[u & 1]);
                bulk_load_1d_stream(st + (p.stage_bytes >> 1), s0 + p.src_step, ng * GB, &s_sbar[u & 1]);
            };
            if (tid == 0) {                                       // the source does not depend on the ring: loads first, then the slot
                issue(0); if (units > 1) issue(1);
                fetch(next);
                if (f >= p.ring) while (ld_acquire(consumed + f - p.ring) < tiles) __nanosleep(100);
            }
            if (f >= p.ring) __syncthreads();
            for (int u = 0; u < units; u++) {
                const int s_ = u & 1, r = u / p.chunks, c = u - r * p.chunks;
                const int g0 = c * p.chunk_groups, ng = min(p.chunk_groups, p.full_groups - g0);
                // one warp polls the mbarrier, the others sleep at the CTA barrier: 256 polling threads took issue slots from the
                // tiles of the neighbouring CTAs (ncu: 115 M warp instructions per launch against 98 M before the strips were staged)
                if (tid < 32) mbar_wait(&s_sbar[s_], (sphase >> s_) & 1);
                __syncthreads();
                sphase ^= 1u << s_;
                const uint8_t* st = smem + s_ * p.stage_bytes;
                uint8_t* drow = mf + (long long)(y_lo + r) * p.mid_step + (long long)g0 * (G * CN);
                for (int hi = tid; hi < 2 * ng; hi += NT)           // one half group (G/2 pixels) per thread
                    box_half_smem<CN>(st + hi * (GB / 2), st + (p.stage_bytes >> 1) + hi * (GB / 2), drow + hi * (G * CN / 2));
                if (u + 2 < units) {
                    __syncthreads();                              // every thread has read stage s_
                    if (tid == 0) issue(u + 2);
                }
            }
            if (p.full_groups < groups)                           // ragged right edge: the last, partial group of each row from global memory
                for (int r = tid; r < rows; r += NT) {
                    const uint8_t* s0 = sf + (long long)(2 * (y_lo + r)) * p.src_step;
                    box_tail<CN>(s0, s0 + p.src_step, mf + (long long)(y_lo + r) * p.mid_step, p.full_groups * G, p.mw);
                }
            asm volatile("fence.proxy.async.global;" ::: "memory");    // this thread's stores vs. later TMA reads (per thread, not cumulative)
            __syncthreads();                                       // the CTA's stores happen-before thread 0's fence (cumulative)
            if (tid == 0) {
                __threadfence();
                atomicAdd(strips_done + f, 1);
            }
        } else {
            // ---- warpAffine tile at (w.z, w.w) of frame f
            // thread 0 draws the next item while the tile's box is in flight (it would only poll the mbarrier meanwhile); drawn
            // during the interpolation the queue's L2 round trips held warp 0 -- and with it the tile -- back
            auto after_load = [&]() {
                atomicAdd(consumed + f, 1);                    // frame f is no longer read by this tile
                fetch(next);                                   // draw the next item while this tile is interpolated
            };
            warp_tile<CN, TW, TH, NT, false, decltype(after_load), 0, NoOp, DIRECT>(smem, smap, dmap, a, w.z, w.w, slot, f, phase, strips_done + f, p.strips,
                                                                                      after_load, &s_sbar[2]);
            phase ^= 1;
        }
    }
}

#ifdef B200_TUNING      // an experiment that did not win (profiles/r2_notes.md section 7): kept for TUNING builds only
// ---------------------------------------------------------------------------------------------------------------------------
// The same pipeline with the two kinds of work on different WARPS of every CTA: 8 warps interpolate tiles (named barrier 1), a ninth
// warp filters strips on its own -- it draws from a strip queue, keeps two bulk copies of source rows in flight all the time and
// never holds a tile warp back.  In the form above a CTA that filters a strip parks all its warps on DRAM for the length of the
// strip (a quarter of the SM's warps on average), and the tiles, which are latency-bound, lose exactly that share.
// Two queues need an argument why nobody waits for ever whatever the device lets run at once: every resident CTA carries BOTH
// roles; queues are drawn in frame order; a strip warp waits only for tiles of an EARLIER frame (its ring slot), a tile group only
// for strips of its own frame, which were drawn before any strip of a later frame.  Take the lowest frame anybody waits for: what
// it waits for is held by running warps that wait for nothing lower, so it completes.
struct Pipe2Args {
    const uint8_t* src; long long src_frame_stride; int src_step;
    uint8_t* mid; long long mid_frame_stride; int mid_step;
    int mw, mh, ring, n_frames, strips, strip_rows, ntx, nty;
    int full_groups, chunk_groups, chunks, stage_bytes, tile_smem;
    int* counters;             // [0] strip queue head, [1] tile queue head, [2 + f] strips done of frame f, [2 + n + f] tiles that consumed frame f
};

template <int CN, int TW, int TH, bool DIRECT>
__global__ void __launch_bounds__(288, 7)
pipe2_box_warp_u8(const __grid_constant__ CUtensorMap smap, const __grid_constant__ CUtensorMap dmap, const AffineFastArgs a,
                  const __grid_constant__ Pipe2Args p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ int4 s_work[2];
    __shared__ __align__(8) uint64_t s_bar[3];                    // [0], [1]: the strip warp's stages; [2]: the tiles' box
    constexpr int G = BoxCfg<CN>::G, GB = G * 2 * CN, NT = 256;
    const int tid = threadIdx.x;
    if (tid == 0) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); mbar_init(&s_bar[2], 1); }
    __syncthreads();
    const int tiles = p.ntx * p.nty;
    int* strips_done = p.counters + 2;
    int* consumed = strips_done + p.n_frames;
    if (tid >= NT) {
        // ------------------------------------------------------------------ the strip warp
        const int lane = tid - NT;
        uint8_t* stage = smem + p.tile_smem;
        const int groups = (p.mw + G - 1) / G, total = p.strips * p.n_frames;
        uint32_t sphase = 0;
        for (;;) {
            int i = 0;
            if (lane == 0) i = atomicAdd(p.counters, 1);
            i = __shfl_sync(0xffffffffu, i, 0);
            if (i >= total) break;
            const int f = i / p.strips, r0 = i - f * p.strips, slot = f % p.ring;
            const uint8_t* sf = p.src + (long long)f * p.src_frame_stride;
            uint8_t* mf = p.mid + (long long)slot * p.mid_frame_stride;
            const int y_lo = r0 * p.strip_rows, rows = min(p.strip_rows, p.mh - y_lo);
            const int units = rows * p.chunks;
            auto issue = [&](int u) {                             // lane 0
                const int r = u / p.chunks, c = u - r * p.chunks;
                const int g0 = c * p.chunk_groups, ng = min(p.chunk_groups, p.full_groups - g0);
                uint8_t* st = stage + (u & 1) * p.stage_bytes;
                const uint8_t* s0 = sf + (long long)(2 * (y_lo + r)) * p.src_step + (long long)g0 * GB;
                mbar_expect_tx(&s_bar[u & 1], 2u * ng * GB);
                bulk_load_1d_stream(st, s0, ng * GB, &s_bar[u & 1]);
                bulk_load_1d_stream(st + (p.stage_bytes >> 1), s0 + p.src_step, ng * GB, &s_bar[u & 1]);
            };
            if (lane == 0) {
                issue(0); if (units > 1) issue(1);
                if (f >= p.ring) while (ld_acquire(consumed + f - p.ring) < tiles) __nanosleep(200);
            }
            __syncwarp();
            for (int u = 0; u < units; u++) {
                const int s_ = u & 1, r = u / p.chunks, c = u - r * p.chunks;
                const int g0 = c * p.chunk_groups, ng = min(p.chunk_groups, p.full_groups - g0);
                mbar_wait(&s_bar[s_], (sphase >> s_) & 1);
                sphase ^= 1u << s_;
                const uint8_t* st = stage + s_ * p.stage_bytes;
                uint8_t* drow = mf + (long long)(y_lo + r) * p.mid_step + (long long)g0 * (G * CN);
                for (int hi = lane; hi < 2 * ng; hi += 32)
                    box_half_smem<CN>(st + hi * (GB / 2), st + (p.stage_bytes >> 1) + hi * (GB / 2), drow + hi * (G * CN / 2));
                __syncwarp();                                     // every lane has read stage s_
                if (lane == 0 && u + 2 < units) issue(u + 2);
            }
            if (p.full_groups < groups)
                for (int r = lane; r < rows; r += 32) {
                    const uint8_t* s0 = sf + (long long)(2 * (y_lo + r)) * p.src_step;
                    box_tail<CN>(s0, s0 + p.src_step, mf + (long long)(y_lo + r) * p.mid_step, p.full_groups * G, p.mw);
                }
            asm volatile("fence.proxy.async.global;" ::: "memory");    // this lane's stores vs. later TMA reads of other CTAs
            __syncwarp();
            if (lane == 0) { __threadfence(); atomicAdd(strips_done + f, 1); }
        }
        return;
    }
    // ---------------------------------------------------------------------- the eight tile warps
    const int total_t = tiles * p.n_frames;
    auto fetch = [&](int4* w) {
        const int i = atomicAdd(p.counters + 1, 1);
        if (i >= total_t) { *w = make_int4(W_END, 0, 0, 0); return; }
        const int f = i / tiles, t = i - f * tiles, ty = t / p.ntx;
        *w = make_int4(W_TILE | ((f % p.ring) << 4), f, (t - ty * p.ntx) * TW, ty * TH);
    };
    if (tid == 0) fetch(&s_work[0]);
    uint32_t phase = 0;
    for (int round = 0;; round++) {
        cta_sync<1, NT>();
        const int4 w = s_work[round & 1];
        int4* const next = &s_work[(round + 1) & 1];
        if ((w.x & 7) == W_END) break;
        const int slot = w.x >> 4, f = w.y;
        auto after_load = [&]() { atomicAdd(consumed + f, 1); fetch(next); };
        warp_tile<CN, TW, TH, NT, false, decltype(after_load), 1, NoOp, DIRECT>(smem, smap, dmap, a, w.z, w.w, slot, f, phase, strips_done + f, p.strips,
                                                                                  after_load, &s_bar[2]);
        phase ^= 1;
    }
}
#endif

// ================================================================= warpPerspective, 8U, TMA-staged ===============
// Same staging as warp_affine_tma_u8; the coordinates are the reference's block-relative fp64 form (persp_coords.cuh).
// A projective map is monotone along every row and column of a tile in which W keeps its sign, so the source box of a tile
// is spanned by the coordinates of its four corner pixels (+-1 pixel for the per-pixel rounding); the host has evaluated
// every tile the same way and sized the TMA box for the largest one.
struct PerspFastArgs {
    double M[9];
    const uint8_t* src; int src_step; long long src_frame_stride;
    int sw, sh, dw, dh;
    int boxw, boxh;
    int border, keep_in_l2, patch;
    uint32_t cval;
    int bw0, bw0_shift;        // coordinate block width of the reference (a power of two >= 32 here)
};

template <int CN, int TW, int TH, int NT>
__global__ void __launch_bounds__(NT, 1536 / NT)
warp_persp_tma_u8(const __grid_constant__ CUtensorMap smap, const __grid_constant__ CUtensorMap dmap, const PerspFastArgs a)
{
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int OUT_BYTES = TW * CN * TH;
    constexpr int ROWS_PER_PASS = NT / TW, PASSES = TH / ROWS_PER_PASS;
    constexpr int NBLK = TW / 32 + 1;                                  // coordinate blocks a tile can touch (bw0 >= 32)
    uint8_t* out_tile = smem;                                          // [TH][TW*CN]
    PerspRow* prow = reinterpret_cast<PerspRow*>(smem + OUT_BYTES);    // [TH][NBLK] row terms per coordinate block
    int2* corner = reinterpret_cast<int2*>(prow + TH * NBLK);          // 4 corner coordinates
    int* org = reinterpret_cast<int*>(corner + 4);
    uint64_t* bar = reinterpret_cast<uint64_t*>(org + 4);
    uint8_t* tile = smem + ((OUT_BYTES + TH * NBLK * (int)sizeof(PerspRow) + 4 * 8 + 16 + 8 + 127) & ~127);
    static_assert(OUT_BYTES % 8 == 0, "row-term table alignment");
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH, frame = blockIdx.z;
    const int sft = a.bw0_shift, blk0 = x0 >> sft;
    if (tid == 0) mbar_init(bar, 1);
    if (tid < TH * NBLK) {
        const int r = tid / NBLK, b = tid - r * NBLK;
        prow[tid] = persp_row_terms(a, (blk0 + b) << sft, y0 + r);
    }
    __syncthreads();
    auto coord = [&](int xl, int r) {                                 // exact (X, Y) of tile pixel (xl, r), 1/32 px units
        const int x = x0 + xl, blk = x >> sft, xb = blk << sft;
        return persp_xy<true>(a, prow[r * NBLK + (blk - blk0)], x - xb, min(a.bw0, a.dw - xb));
    };
    const int cl = min(TW, a.dw - x0) - 1, rl = min(TH, a.dh - y0) - 1;
    if (tid < 4) corner[tid] = coord((tid & 1) ? cl : 0, (tid & 2) ? rl : 0);
    __syncthreads();
    if (tid == 0) {
        const int mx = min(min(corner[0].x, corner[1].x), min(corner[2].x, corner[3].x));
        const int my = min(min(corner[0].y, corner[1].y), min(corner[2].y, corner[3].y));
        const int bx0 = (mx >> 5) - 1, by0 = (my >> 5) - 1;            // one pixel of slack for the per-pixel rounding
        const int bxb = (bx0 * CN) & ~15;                              // TMA: innermost coordinate 16-byte aligned
        org[0] = bxb; org[1] = by0;
        org[2] = (bxb < 0 || by0 < 0 || bxb + a.boxw > a.sw * CN || by0 + a.boxh > a.sh);
        mbar_expect_tx(bar, (uint32_t)(a.boxw * a.boxh));
        tma_load_3d(tile, &smap, bar, bxb, by0, frame);
    }
    if (tid < 32) mbar_wait(bar, 0);
    __syncthreads();
    const int bxb = org[0], by0 = org[1];
    const int boxw = a.boxw;
    if (a.patch && org[2]) {
        patch_border<CN, NT>(tile, a, bxb, by0, frame);
        __syncthreads();
    }
    const int xl = tid % TW, rg = tid / TW;
    uint8_t* const Dbase = out_tile + (rg * PASSES) * (TW * CN) + xl * CN;
    if (x0 + xl < a.dw) {
#pragma unroll 2
        for (int p = 0; p < PASSES; p++) {
            const int r = rg * PASSES + p;
            if (y0 + r >= a.dh) break;
            const int2 c = coord(xl, r);
            warp_px_xy<CN, true>(tile, boxw, c.x, c.y, bxb, by0, Dbase + p * (TW * CN));
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) tma_store_3d(&dmap, out_tile, x0 * CN, y0, frame, a.keep_in_l2 != 0);
}

template <int CN, int TW, int TH, int NT>
cudaError_t launch_fast(const CUtensorMap& sm, const CUtensorMap& dm, const AffineFastArgs& fa, int n_frames, size_t smem, cudaStream_t s)
{
    dim3 grid((fa.dw + TW - 1) / TW, (fa.dh + TH - 1) / TH, n_frames);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(warp_affine_tma_u8<CN, TW, TH, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    warp_affine_tma_u8<CN, TW, TH, NT><<<grid, NT, smem, s>>>(sm, dm, fa);
    return cudaGetLastError();
}

} // namespace

// Shared-memory row pitch of the staged box (= TMA box width, a multiple of 16 bytes).  A warp gathers along a
// slanted source line, so the pitch decides the bank pattern: rows that differ by k map to banks shifted by
// k*pitch/4.  The host simulates the lanes of a warp for each admissible pitch and takes the cheapest.
// Model, checked against ncu's l1tex wavefront counters on B200 (profiles/r1_notes.md):
//   32-bit loads (cn 1, 4): one wavefront per distinct word per bank over the 32 lanes;
//   64-bit loads (cn 3):    each half warp separately, one wavefront per distinct 8-byte word per bank pair
//                           (measured 2.08 wavefronts per LDS.64 at pitch 192 and 3.65 at pitch 160 for the bench
//                           matrix; the model says 2.0 and 4.0).
// Returns the mean wavefronts per load instruction relative to the conflict-free count (1.0 = no conflicts).
static double gather_cost(int bw, int cn, double m0, double m3, bool w64 = true)
{
    long cost = 0, ideal = 0;
    for (int ph = 0; ph < 8; ph++) {
        const double fx0 = ph * 0.37 + 0.11, fy0 = ph * 0.29 + 0.07;
        int o[32];
        for (int l = 0; l < 32; l++) {
            const int sx = (int)floor(fx0 + m0 * l + 64.0), sy = (int)floor(fy0 + m3 * l + 64.0);
            o[l] = sy * bw + sx * cn;
        }
        const bool wide = cn == 3 && w64;
        const int unit = wide ? 8 : 4, nbank = wide ? 16 : 32, group = wide ? 16 : 32, nwd = (cn == 3 && !w64) ? 3 : 2;
        for (int wd = 0; wd < nwd; wd++)
            for (int g0 = 0; g0 < 32; g0 += group) {
                int cnt[32] = {0}; int seen[32][8]; int nseen[32] = {0};
                for (int l = g0; l < g0 + group; l++) {
                    const int w = o[l] / unit + wd, b = w % nbank;
                    bool dup = false;
                    for (int k = 0; k < nseen[b]; k++) if (seen[b][k] == w) dup = true;
                    if (!dup && nseen[b] < 8) { seen[b][nseen[b]++] = w; cnt[b]++; }
                }
                int mx = 1;
                for (int b = 0; b < nbank; b++) mx = cnt[b] > mx ? cnt[b] : mx;
                cost += mx; ideal += 1;
            }
    }
    return (double)cost / (double)ideal;
}

struct FastPlan { int tw, th, nt, boxw, boxh; size_t smem; };

// Tile shape, CTA size and pitch for one call.  Candidates in order of preference (most pixels per thread first: the
// per-tile set-up is ~90 instructions per thread); a later candidate wins only if its gather is >5 % cheaper.
// Remembered per host thread (the simulation costs tens of microseconds).
#ifdef B200_TUNING
static const char* tuning_env(const char* name) { return getenv(name); }     // experiments only (make TUNING=1)
#else
static const char* tuning_env(const char*) { return nullptr; }
#endif

// the source image as a tensor map of 32-bit elements when its rows are whole words: boxes up to 1024 bytes wide instead of
// 256 (the box limit is 256 ELEMENTS per dimension), which is what lets a 3-channel tile be 64 pixels wide
static bool rows_are_words(const Image& src, int cn) { return ((long long)src.w * cn) % 4 == 0; }

static bool plan_affine(const WarpArgs& a, int cn, int n_frames, bool w64, FastPlan& out)
{
    const int max_boxw = (rows_are_words(a.src, cn) ? 1024 : 256) + (w64 ? 0 : 1);   // (+1: the two fetch variants plan separately)
    struct Key { int cn, dw, dh, nf, max_boxw; double m0, m1, m3, m4; FastPlan plan; bool ok; };
    static thread_local Key cache[4];
    static thread_local int next = 0;
    for (const Key& k : cache)
        if (k.cn == cn && k.dw == a.dst.w && k.dh == a.dst.h && k.nf == n_frames && k.max_boxw == max_boxw && k.m0 == a.M[0] && k.m1 == a.M[1] &&
            k.m3 == a.M[3] && k.m4 == a.M[4]) {
            out = k.plan; return k.ok;
        }
    struct Shape { int tw, th, nt; };
    // stand-alone kernel: 16 pixels per thread first; the frame pipeline measured best with 64x32 tiles on 256 threads
    // (2048 resident threads also feed its DRAM-bound box stage): 151 us per 64 frames vs 168 us with 64x64 tiles
    const Shape shapes_w64[] = {{64, 64, 256}, {64, 32, 256}, {32, 64, 256}, {32, 32, 128}, {64, 16, 128}, {32, 16, 64}};
    const Shape shapes_w32[] = {{64, 32, 256}, {32, 64, 256}, {64, 64, 256}, {32, 32, 128}, {64, 16, 128}, {32, 16, 64}};
    const Shape* shapes = w64 ? shapes_w64 : shapes_w32;
    int e_tw = 0, e_th = 0, e_pitch = 0, e_nt = 256;
    if (const char* e = tuning_env("B200_EXP_SHAPE")) sscanf(e, "%d,%d,%d,%d", &e_tw, &e_th, &e_pitch, &e_nt);   // "tw,th,pitch,nt"
    const DeviceLimits lim = device_limits();
    const size_t smem_cap = std::min<size_t>((size_t)lim.smem_per_block_optin, 56 * 1024);
    bool found = false; double best = 0.0;
    for (int si = 0; si < 6; si++) {
        const Shape& sh = shapes[si];
        if (e_tw && (sh.tw != e_tw || sh.th != e_th || sh.nt != e_nt)) continue;
        // the largest tiles amortise the per-tile set-up best but need enough tiles to fill every SM (8 CTAs each) twice
        if (sh.tw * sh.th >= 2048 &&
            (long long)((a.dst.w + sh.tw - 1) / sh.tw) * ((a.dst.h + sh.th - 1) / sh.th) * n_frames < 16LL * lim.sms) continue;
        // extent of the source box of a tw x th tile (+1 right/bottom tap, +1 rounding, +1 floor)
        const int ext_x = (int)ceil(fabs(a.M[0]) * (sh.tw - 1) + fabs(a.M[1]) * (sh.th - 1)) + 3;
        const int ext_y = (int)ceil(fabs(a.M[3]) * (sh.tw - 1) + fabs(a.M[4]) * (sh.th - 1)) + 3;
        const int min_boxw = (ext_x * cn + 15 + 15) & ~15, boxh = ext_y;  // +15: start is rounded down to 16 bytes
        if (min_boxw > (max_boxw & ~1) || boxh > 256) continue;
        int boxw = min_boxw; double cost = -1.0;
        for (int bw = min_boxw; bw <= (max_boxw & ~1) && bw <= min_boxw + 128; bw += 16) {
            const double c = gather_cost(bw, cn, a.M[0], a.M[3], w64) + (bw - min_boxw) * 1e-4;   // tie: narrower box
            if (cost < 0 || c < cost) { cost = c; boxw = bw; }
        }
        if (e_pitch >= min_boxw && e_pitch <= (max_boxw & ~1)) boxw = e_pitch;
        const size_t head = ((size_t)(sh.tw * cn * (sh.th / out_phases(sh.tw, sh.th, cn)) + (2 * sh.tw + 2 * sh.th + 4) * 4 + 8 + 127) & ~(size_t)127);
        const size_t smem = head + (size_t)boxw * boxh + 16;
        if (smem > smem_cap) continue;
        if (!found || cost < best * 0.95) {
            found = true; best = cost;
            out = FastPlan{sh.tw, sh.th, sh.nt, boxw, boxh, smem};
        }
    }
    Key& k = cache[next];
    next = (next + 1) & 3;
    k = Key{cn, a.dst.w, a.dst.h, n_frames, max_boxw, a.M[0], a.M[1], a.M[3], a.M[4], out, found};
    return found;
}

// tensor map of the source for a planned box: 32-bit elements when the rows are whole words (see rows_are_words)
static bool make_src_map(CUtensorMap* m, int* x_shift, const Image& src, int cn, int frames, long long frame_stride, const FastPlan& pl)
{
    if (rows_are_words(src, cn)) {
        *x_shift = 2;
        return make_tensor_map(m, 4, src.p, (long long)src.w * cn / 4, src.h, frames, src.step, frame_stride, pl.boxw / 4, pl.boxh);
    }
    *x_shift = 0;
    return make_byte_map(m, src.p, (long long)src.w * cn, src.h, frames, src.step, frame_stride, pl.boxw, pl.boxh);
}

// returns 0 on launch, -1 when the fast path does not apply (caller uses the generic kernel), >0 = cudaError
int try_launch_warp_affine_tma_u8(const WarpArgs& a, int cn, int n_frames, cudaStream_t stream)
{
    if (!a.linear || !(cn == 1 || cn == 3 || cn == 4) || a.border == B_TRANSPARENT) return -1;
    // TMA needs 16-byte aligned bases / strides
    if (((uintptr_t)a.src.p & 15) || (a.src.step & 15) || (a.src_frame_stride & 15)) return -1;
    if (((uintptr_t)a.dst.p & 15) || (a.dst.step & 15) || (a.dst_frame_stride & 15)) return -1;
    for (int i = 0; i < 6; i++) if (!std::isfinite(a.M[i])) return -1;
    // the kernel skips the int16 saturation of the coordinates and relies on int32 headroom: require every
    // source coordinate reached from the dst rectangle to be far inside both (imgwarp.cpp:2301-2358)
    const double ex = fabs(a.M[0]) * a.dst.w + fabs(a.M[1]) * a.dst.h + fabs(a.M[2]);
    const double ey = fabs(a.M[3]) * a.dst.w + fabs(a.M[4]) * a.dst.h + fabs(a.M[5]);
    if (!(ex < 30000.0 && ey < 30000.0)) return -1;

    FastPlan pl;
    if (!plan_affine(a, cn, n_frames, true, pl)) return -1;
    CUtensorMap smap, dmap;
    AffineFastArgs fa;
    fa.exp = 0;
    if (!make_src_map(&smap, &fa.x_shift, a.src, cn, n_frames, a.src_frame_stride, pl)) return -1;
    if (!make_byte_map(&dmap, a.dst.p, (long long)a.dst.w * cn, a.dst.h, n_frames, a.dst.step, a.dst_frame_stride, pl.tw * cn, pl.th / out_phases(pl.tw, pl.th, cn))) return -1;
    memcpy(fa.M, a.M, sizeof(fa.M));
    fa.src = a.src.p; fa.src_step = a.src.step; fa.src_frame_stride = a.src_frame_stride;
    fa.sw = a.src.w; fa.sh = a.src.h; fa.dw = a.dst.w; fa.dh = a.dst.h;
    fa.dst = a.dst.p; fa.dst_step = a.dst.step; fa.dst_frame_stride = a.dst_frame_stride;
    fa.boxw = pl.boxw; fa.boxh = pl.boxh; fa.border = a.border; fa.keep_in_l2 = a.keep_in_l2;
    memcpy(&fa.cval, a.cval, 4);
    fa.patch = !(a.border == B_CONSTANT && fa.cval == 0);
    cudaError_t e;
#define B200_FAST_CASE(TWV, THV, NTV) if (pl.tw == TWV && pl.th == THV && pl.nt == NTV) { \
        e = cn == 1 ? launch_fast<1, TWV, THV, NTV>(smap, dmap, fa, n_frames, pl.smem, stream) \
          : cn == 3 ? launch_fast<3, TWV, THV, NTV>(smap, dmap, fa, n_frames, pl.smem, stream) \
                    : launch_fast<4, TWV, THV, NTV>(smap, dmap, fa, n_frames, pl.smem, stream); return (int)e; }
    B200_FAST_CASE(64, 64, 256) B200_FAST_CASE(64, 32, 256) B200_FAST_CASE(32, 64, 256)
    B200_FAST_CASE(32, 32, 128) B200_FAST_CASE(64, 16, 128) B200_FAST_CASE(32, 16, 64)
#undef B200_FAST_CASE
    return -1;
}

// resize (exact 2x2 box) -> warpAffine over n_frames frames in one launch (see pipe_box_warp_u8).  `mid` is the caller's
// scratch for the half-size frames: its first `ring` frames hold the ring, the queue counters sit behind them.
// Returns 0 = launched, -1 = not applicable (caller runs the two batched kernels), >0 = cudaError.
int try_launch_pipe_box_warp_u8(const ResizeArgs& ra, const WarpArgs& wa, int cn, int n_frames, size_t mid_bytes, cudaStream_t stream)
{
    if (const char* e = tuning_env("B200_PIPE_FUSED")) { if (!atoi(e)) return -1; }
    if (ra.force_generic || wa.force_generic) return -1;
    if (!(cn == 1 || cn == 3 || cn == 4) || wa.interp != 1 || wa.border == B_TRANSPARENT) return -1;
    if (!box2x2_u8_applicable(ra, cn)) return -1;
    if (ra.dst.p != wa.src.p || ra.dst.step != wa.src.step || ra.dst_frame_stride != wa.src_frame_stride) return -1;
    if (((uintptr_t)wa.src.p & 15) || (wa.src.step & 15) || (wa.src_frame_stride & 15)) return -1;
    if (((uintptr_t)wa.dst.p & 15) || (wa.dst.step & 15) || (wa.dst_frame_stride & 15)) return -1;
    for (int i = 0; i < 6; i++) if (!std::isfinite(wa.M[i])) return -1;
    const double ex = fabs(wa.M[0]) * wa.dst.w + fabs(wa.M[1]) * wa.dst.h + fabs(wa.M[2]);
    const double ey = fabs(wa.M[3]) * wa.dst.w + fabs(wa.M[4]) * wa.dst.h + fabs(wa.M[5]);
    if (!(ex < 30000.0 && ey < 30000.0)) return -1;
    FastPlan pl;
    if (!plan_affine(wa, cn, n_frames, false, pl)) return -1;
    // ring: enough frames that strips rarely wait for tiles, few enough to stay in L2 (126 MB) beside the streams
    const long long mfs = wa.src_frame_stride;
    int exp_ring_mb = 24, exp_sr = 4, exp_ahead = 10;          // measured optimum on B200 (profiles/r1_notes.md)
    if (const char* e = tuning_env("B200_EXP_RING_MB")) exp_ring_mb = atoi(e);
    if (const char* e = tuning_env("B200_EXP_SR")) exp_sr = atoi(e);
    if (const char* e = tuning_env("B200_EXP_AHEAD")) exp_ahead = atoi(e);
    int ring = (int)(((long long)exp_ring_mb << 20) / (mfs > 0 ? mfs : 1));
    if (ring > 32) ring = 32;
    if (ring < 4) return -1;
    const size_t counter_bytes = (size_t)(1 + 2 * n_frames) * sizeof(int);
    const size_t counters_off = ((size_t)ring * mfs + 255) & ~(size_t)255;
    if (n_frames <= ring || counters_off + counter_bytes > mid_bytes) return -1;

    PipeArgs p;
    p.src = ra.src.p; p.src_frame_stride = ra.src_frame_stride; p.src_step = ra.src.step;
    p.mid = ra.dst.p; p.mid_frame_stride = mfs; p.mid_step = ra.dst.step;
    p.mw = ra.dst.w; p.mh = ra.dst.h; p.ring = ring; p.n_frames = n_frames;
    p.strip_rows = exp_sr; p.strips = (p.mh + p.strip_rows - 1) / p.strip_rows;
    p.ahead = exp_ahead < ring - 1 ? exp_ahead : ring - 2;
    if (p.ahead < 0) p.ahead = 0;
    if (p.ahead > n_frames) p.ahead = n_frames;
#ifdef B200_TUNING
    if (const char* e2 = tuning_env("B200_EXP_PIPE2")) {
        if (atoi(e2) && pl.tw == 64 && pl.th == 32 && pl.nt == 256) {
            const int G = cn == 1 ? 16 : cn == 3 ? 8 : 4, GB = G * 2 * cn;
            int stage_row = 1920;
            if (const char* e3 = tuning_env("B200_EXP_STAGE_ROW")) stage_row = atoi(e3);
            Pipe2Args q;
            q.src = p.src; q.src_frame_stride = p.src_frame_stride; q.src_step = p.src_step;
            q.mid = p.mid; q.mid_frame_stride = p.mid_frame_stride; q.mid_step = p.mid_step;
            q.mw = p.mw; q.mh = p.mh; q.ring = p.ring; q.n_frames = p.n_frames; q.strips = p.strips; q.strip_rows = p.strip_rows;
            q.ntx = (wa.dst.w + pl.tw - 1) / pl.tw; q.nty = (wa.dst.h + pl.th - 1) / pl.th;
            q.full_groups = p.mw / G;
            if (q.full_groups < 1) return -1;
            q.chunk_groups = q.full_groups < stage_row / GB ? q.full_groups : stage_row / GB;
            if (q.chunk_groups < 1) q.chunk_groups = 1;
            q.chunks = (q.full_groups + q.chunk_groups - 1) / q.chunk_groups;
            q.stage_bytes = 2 * ((q.chunk_groups * GB + 127) & ~127);
            int direct = 0;
            if (const char* e5 = tuning_env("B200_EXP_DIRECT")) direct = atoi(e5);
            q.tile_smem = (int)((pl.smem - (direct ? (size_t)pl.tw * cn * pl.th : 0) + 127) & ~(size_t)127);
            const size_t smem2 = (size_t)q.tile_smem + 2 * (size_t)q.stage_bytes;
            const size_t cb2 = (size_t)(2 + 2 * n_frames) * sizeof(int);
            if (counters_off + cb2 > mid_bytes) return -1;
            q.counters = reinterpret_cast<int*>(ra.dst.p + counters_off);
            CUtensorMap smap2, dmap2;
            AffineFastArgs fb;
            if (!make_src_map(&smap2, &fb.x_shift, wa.src, cn, ring, mfs, pl)) return -1;
            if (!make_byte_map(&dmap2, wa.dst.p, (long long)wa.dst.w * cn, wa.dst.h, n_frames, wa.dst.step, wa.dst_frame_stride, pl.tw * cn, pl.th / out_phases(pl.tw, pl.th, cn))) return -1;
            memcpy(fb.M, wa.M, sizeof(fb.M));
            fb.src = wa.src.p; fb.src_step = wa.src.step; fb.src_frame_stride = mfs;
            fb.sw = wa.src.w; fb.sh = wa.src.h; fb.dw = wa.dst.w; fb.dh = wa.dst.h;
            fb.dst = wa.dst.p; fb.dst_step = wa.dst.step; fb.dst_frame_stride = wa.dst_frame_stride;
            fb.boxw = pl.boxw; fb.boxh = pl.boxh; fb.border = wa.border; fb.keep_in_l2 = wa.keep_in_l2;
            memcpy(&fb.cval, wa.cval, 4);
            fb.patch = !(wa.border == B_CONSTANT && fb.cval == 0);
            fb.exp = 0;
            cudaError_t e4 = cudaMemsetAsync(q.counters, 0, cb2, stream);
            if (e4 != cudaSuccess) return (int)e4;
            const DeviceLimits lim2 = device_limits();
            int per_sm = (int)((size_t)lim2.smem_per_sm / (smem2 + 1024));
            if (per_sm > 7) per_sm = 7;
            if (per_sm < 1) return -1;
            const int grid2 = lim2.sms * per_sm;
#define B200_PIPE2(CNV) { if (direct) { if (smem2 > 48 * 1024) cudaFuncSetAttribute(pipe2_box_warp_u8<CNV, 64, 32, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2); \
                                         pipe2_box_warp_u8<CNV, 64, 32, true><<<grid2, 288, smem2, stream>>>(smap2, dmap2, fb, q); } \
                          else { if (smem2 > 48 * 1024) cudaFuncSetAttribute(pipe2_box_warp_u8<CNV, 64, 32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2); \
                                 pipe2_box_warp_u8<CNV, 64, 32, false><<<grid2, 288, smem2, stream>>>(smap2, dmap2, fb, q); } }
            if (cn == 1) B200_PIPE2(1) else if (cn == 3) B200_PIPE2(3) else B200_PIPE2(4)
#undef B200_PIPE2
            return (int)cudaGetLastError();
        }
    }
#endif
    bool pipe_direct = false;
    {   // strips: row pairs staged in shared memory, at most STAGE_ROW bytes of a source row per stage, two stages
        const int G = cn == 1 ? 16 : cn == 3 ? 8 : 4, GB = G * 2 * cn, STAGE_ROW = 5760;
        p.full_groups = p.mw / G;
        if (p.full_groups < 1) return -1;
        p.chunk_groups = p.full_groups < STAGE_ROW / GB ? p.full_groups : STAGE_ROW / GB;
        p.chunks = (p.full_groups + p.chunk_groups - 1) / p.chunk_groups;
        p.stage_bytes = 2 * ((p.chunk_groups * GB + 127) & ~127);
#ifdef B200_TUNING
        if (const char* e5 = tuning_env("B200_EXP_DIRECT")) if (atoi(e5)) { pipe_direct = true; pl.smem -= (size_t)pl.tw * cn * pl.th; }
#endif
        if ((size_t)2 * p.stage_bytes > pl.smem) pl.smem = (size_t)2 * p.stage_bytes;
    }
    p.ntx = (wa.dst.w + pl.tw - 1) / pl.tw; p.nty = (wa.dst.h + pl.th - 1) / pl.th;
    if ((long long)(p.strips + p.ntx * p.nty) * (n_frames + p.ahead) > 0x7fffffffLL) return -1;
    p.counters = reinterpret_cast<int*>(ra.dst.p + counters_off);
    CUtensorMap smap, dmap;
    AffineFastArgs fa;
    if (!make_src_map(&smap, &fa.x_shift, wa.src, cn, ring, mfs, pl)) return -1;
    if (!make_byte_map(&dmap, wa.dst.p, (long long)wa.dst.w * cn, wa.dst.h, n_frames, wa.dst.step, wa.dst_frame_stride, pl.tw * cn, pl.th / out_phases(pl.tw, pl.th, cn))) return -1;
    memcpy(fa.M, wa.M, sizeof(fa.M));
    fa.src = wa.src.p; fa.src_step = wa.src.step; fa.src_frame_stride = mfs;
    fa.sw = wa.src.w; fa.sh = wa.src.h; fa.dw = wa.dst.w; fa.dh = wa.dst.h;
    fa.dst = wa.dst.p; fa.dst_step = wa.dst.step; fa.dst_frame_stride = wa.dst_frame_stride;
    fa.boxw = pl.boxw; fa.boxh = pl.boxh; fa.border = wa.border; fa.keep_in_l2 = wa.keep_in_l2;
    memcpy(&fa.cval, wa.cval, 4);
    fa.patch = !(wa.border == B_CONSTANT && fa.cval == 0);
    fa.exp = 0;
    if (const char* e = tuning_env("B200_EXP_FLAGS")) fa.exp = atoi(e);
    cudaError_t e = cudaMemsetAsync(p.counters, 0, counter_bytes, stream);
    if (e != cudaSuccess) return (int)e;
    const int ctas_per_sm = (cn == 1 ? 1536 : 2048) / pl.nt;
    const DeviceLimits lim = device_limits();
    const int by_smem = (int)((size_t)lim.smem_per_sm / (pl.smem + 1024));
    const int grid = lim.sms * (ctas_per_sm < by_smem ? ctas_per_sm : by_smem);
#define B200_PIPE_LAUNCH(K) { if (pl.smem > 48 * 1024) { cudaError_t e_ = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem); \
                                                          if (e_ != cudaSuccess) return (int)e_; } \
                              K<<<grid, pl.nt, pl.smem, stream>>>(smap, dmap, fa, p); }
#ifdef B200_TUNING
#define B200_PIPE_DIRECT(TWV, THV, NTV) if (pipe_direct && cn == 3 && pl.tw == TWV && pl.th == THV && pl.nt == NTV) { \
        B200_PIPE_LAUNCH((pipe_box_warp_u8<3, TWV, THV, NTV, true>)) return (int)cudaGetLastError(); }
    B200_PIPE_DIRECT(64, 64, 256) B200_PIPE_DIRECT(64, 32, 256)
#undef B200_PIPE_DIRECT
#endif
    (void)pipe_direct;
#define B200_PIPE_CASE(TWV, THV, NTV) if (pl.tw == TWV && pl.th == THV && pl.nt == NTV) { \
        if (cn == 1) B200_PIPE_LAUNCH((pipe_box_warp_u8<1, TWV, THV, NTV>)) \
        else if (cn == 3) B200_PIPE_LAUNCH((pipe_box_warp_u8<3, TWV, THV, NTV>)) \
        else B200_PIPE_LAUNCH((pipe_box_warp_u8<4, TWV, THV, NTV>)) \
        return (int)cudaGetLastError(); }
    B200_PIPE_CASE(64, 64, 256) B200_PIPE_CASE(64, 32, 256) B200_PIPE_CASE(32, 64, 256) B200_PIPE_CASE(32, 32, 128)
#undef B200_PIPE_CASE
#undef B200_PIPE_LAUNCH
    return -1;
}

// ---- warpPerspective fast path: host side ---------------------------------------------------------------------------
// the reference's coordinate of destination pixel (x, y) in fp64, as warpPerspectiveBlockline evaluates it (imgwarp.cpp:2712-2714,
// :2803-2821); false when W is zero / not finite there
static bool persp_host_xy(const double* M, int bw0, int x, int y, double& X, double& Y, double& W)
{
    const int xb = (x / bw0) * bw0, x1 = x - xb;
    const double X0 = M[0] * xb + M[1] * y + M[2], Y0 = M[3] * xb + M[4] * y + M[5], W0 = M[6] * xb + M[7] * y + M[8];
    W = W0 + M[6] * x1;
    if (!(fabs(W) > 1e-30 && fabs(W) < 1e30)) return false;
    const double q = 32.0 / W;
    X = nearbyint((X0 + M[0] * x1) * q); Y = nearbyint((Y0 + M[3] * x1) * q);
    return std::isfinite(X) && std::isfinite(Y);
}

struct PerspPlan { int tw, th, nt, boxw, boxh; size_t smem; bool ok; };

// Largest source box over all tiles of a shape, evaluated exactly like the kernel does (corner pixels, +-1 pixel slack).
// Remembered per host thread: a video stream warps every frame with the same matrix.
static bool plan_persp(const WarpArgs& a, int cn, PerspPlan& out)
{
    struct Key { double M[9]; int dw, dh, sw, sh, cn; PerspPlan plan; };
    static thread_local Key cache[2];
    static thread_local int next = 0;
    for (const Key& k : cache)
        if (k.cn == cn && k.dw == a.dst.w && k.dh == a.dst.h && k.sw == a.src.w && k.sh == a.src.h && !memcmp(k.M, a.M, sizeof(k.M))) {
            out = k.plan; return out.ok;
        }
    out = PerspPlan{0, 0, 0, 0, 0, 0, false};
    struct Shape { int tw, th, nt; };
    const Shape shapes[] = {{32, 64, 256}, {64, 32, 256}, {32, 32, 128}, {32, 16, 64}};
    const int dw = a.dst.w, dh = a.dst.h;
    for (const Shape& sh : shapes) {
        const long long tiles = (long long)((dw + sh.tw - 1) / sh.tw) * ((dh + sh.th - 1) / sh.th);
        if (sh.tw * sh.th >= 2048 && tiles < 16LL * device_limits().sms) continue;
        int ext_x = 0, ext_y = 0; bool good = true; int wsign = 0;
        for (int y0 = 0; y0 < dh && good; y0 += sh.th)
            for (int x0 = 0; x0 < dw && good; x0 += sh.tw) {
                const int cl = std::min(sh.tw, dw - x0) - 1, rl = std::min(sh.th, dh - y0) - 1;
                double mnx = 0, mxx = 0, mny = 0, mxy = 0;
                for (int c = 0; c < 4; c++) {
                    double X, Y, W;
                    if (!persp_host_xy(a.M, a.bw0, x0 + ((c & 1) ? cl : 0), y0 + ((c & 2) ? rl : 0), X, Y, W)) { good = false; break; }
                    const int sg = W > 0 ? 1 : -1;
                    if (!wsign) wsign = sg;
                    if (sg != wsign || fabs(X) > 30000.0 * 32 || fabs(Y) > 30000.0 * 32) { good = false; break; }
                    const double sx = floor(X / 32), sy = floor(Y / 32);
                    if (c == 0) { mnx = mxx = sx; mny = mxy = sy; }
                    else { mnx = std::min(mnx, sx); mxx = std::max(mxx, sx); mny = std::min(mny, sy); mxy = std::max(mxy, sy); }
                }
                if (!good) break;
                ext_x = std::max(ext_x, (int)(mxx - mnx) + 4);     // +1 right tap, +1 below origin, +1 above, +1 count
                ext_y = std::max(ext_y, (int)(mxy - mny) + 4);
            }
        if (!good) break;                                         // W changes sign or coordinates out of range: generic kernel
        const int min_boxw = (ext_x * cn + 15 + 15) & ~15, boxh = ext_y;
        if (min_boxw > 256 || boxh > 256) continue;
        // local slope of the source position along a destination row, for the bank model
        double X0, Y0, X1, Y1, W;
        double m0 = 1.0, m3 = 0.0;
        if (persp_host_xy(a.M, a.bw0, dw / 2, dh / 2, X0, Y0, W) && persp_host_xy(a.M, a.bw0, std::min(dw / 2 + 16, dw - 1), dh / 2, X1, Y1, W) && dw > 32) {
            m0 = (X1 - X0) / 32 / 16; m3 = (Y1 - Y0) / 32 / 16;
        }
        int boxw = min_boxw; double cost = -1.0;
        for (int bw = min_boxw; bw <= 256; bw += 16) {
            const double c = gather_cost(bw, cn, m0, m3) + (bw - min_boxw) * 1e-4;
            if (cost < 0 || c < cost) { cost = c; boxw = bw; }
        }
        const int nblk = sh.tw / 32 + 1;
        const size_t head = ((size_t)(sh.tw * cn * sh.th + sh.th * nblk * 24 + 4 * 8 + 16 + 8 + 127) & ~(size_t)127);
        const size_t smem = head + (size_t)boxw * boxh + 16;
        if (smem > 40 * 1024) continue;
        out = PerspPlan{sh.tw, sh.th, sh.nt, boxw, boxh, smem, true};
        break;
    }
    Key& k = cache[next]; next ^= 1;
    memcpy(k.M, a.M, sizeof(k.M)); k.dw = a.dst.w; k.dh = a.dst.h; k.sw = a.src.w; k.sh = a.src.h; k.cn = cn; k.plan = out;
    return out.ok;
}

// returns 0 on launch, -1 when the fast path does not apply (caller uses the generic kernel), >0 = cudaError
int try_launch_warp_persp_tma_u8(const WarpArgs& a, int cn, int n_frames, cudaStream_t stream)
{
    // Measured at 4K (profiles/r1_notes.md): 8UC1 41.8 us staged vs 47.2 us generic, but 8UC3 71 vs 53 and 8UC4 77 vs 53 -- the
    // fp64 coordinate derivation (~110 instructions per pixel) dominates either way and the staged kernel holds fewer
    // warps per SM, so only single-channel images take this path.
    if (cn != 1 && !tuning_env("B200_PERSP_TMA_ALL")) return -1;
    if (a.interp != 1 || !(cn == 1 || cn == 3 || cn == 4) || a.border == B_TRANSPARENT || a.bw0_shift < 5) return -1;
    if (((uintptr_t)a.src.p & 15) || (a.src.step & 15) || (a.src_frame_stride & 15)) return -1;
    if (((uintptr_t)a.dst.p & 15) || (a.dst.step & 15) || (a.dst_frame_stride & 15)) return -1;
    for (int i = 0; i < 9; i++) if (!std::isfinite(a.M[i])) return -1;
    PerspPlan pl;
    if (!plan_persp(a, cn, pl)) return -1;
    CUtensorMap smap, dmap;
    if (!make_byte_map(&smap, a.src.p, (long long)a.src.w * cn, a.src.h, n_frames, a.src.step, a.src_frame_stride, pl.boxw, pl.boxh)) return -1;
    if (!make_byte_map(&dmap, a.dst.p, (long long)a.dst.w * cn, a.dst.h, n_frames, a.dst.step, a.dst_frame_stride, pl.tw * cn, pl.th)) return -1;
    PerspFastArgs fa;
    memcpy(fa.M, a.M, sizeof(fa.M));
    fa.src = a.src.p; fa.src_step = a.src.step; fa.src_frame_stride = a.src_frame_stride;
    fa.sw = a.src.w; fa.sh = a.src.h; fa.dw = a.dst.w; fa.dh = a.dst.h;
    fa.boxw = pl.boxw; fa.boxh = pl.boxh; fa.border = a.border; fa.keep_in_l2 = a.keep_in_l2;
    memcpy(&fa.cval, a.cval, 4);
    fa.patch = !(a.border == B_CONSTANT && fa.cval == 0);
    fa.bw0 = a.bw0; fa.bw0_shift = a.bw0_shift;
    dim3 grid((fa.dw + pl.tw - 1) / pl.tw, (fa.dh + pl.th - 1) / pl.th, n_frames);
#define B200_PERSP_CASE(TWV, THV, NTV) if (pl.tw == TWV && pl.th == THV && pl.nt == NTV) { \
        if (cn == 1) warp_persp_tma_u8<1, TWV, THV, NTV><<<grid, NTV, pl.smem, stream>>>(smap, dmap, fa); \
        else if (cn == 3) warp_persp_tma_u8<3, TWV, THV, NTV><<<grid, NTV, pl.smem, stream>>>(smap, dmap, fa); \
        else warp_persp_tma_u8<4, TWV, THV, NTV><<<grid, NTV, pl.smem, stream>>>(smap, dmap, fa); \
        return (int)cudaGetLastError(); }
    B200_PERSP_CASE(32, 64, 256) B200_PERSP_CASE(64, 32, 256) B200_PERSP_CASE(32, 32, 128) B200_PERSP_CASE(32, 16, 64)
#undef B200_PERSP_CASE
    return -1;
}

} // namespace b200
