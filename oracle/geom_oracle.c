/*
 * TEST INFRASTRUCTURE ONLY -- this file is the parity ORACLE, never the product path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load it.
 *
 * Plain-C, single-threaded restatement of the arithmetic of the reference CPU path for
 *   cv::warpAffine / cv::warpPerspective / cv::remap / cv::resize
 * (INTER_NEAREST, INTER_LINEAR, INTER_AREA; CV_8U/16U/32F (+16S); BORDER_CONSTANT,
 * REPLICATE, REFLECT, WRAP, REFLECT_101), cv::convertMaps, cv::initUndistortRectifyMap
 * (standard model as its AVX2 dispatch variant computes it, and fisheye), dnn::blobFromImages.  Every function cites the reference
 * file:line it follows (paths relative to /root/reference/modules).
 *
 * Parity status: PINNED.  tests/test_oracle_pinned.py checks this restatement
 *   (a) against the reference's own known-answer tests (test_imgwarp.cpp:1241-1314,
 *       :1441-1479, :1494-1507, :1564-1588, test_imgwarp_strict.cpp:1101-1152) and
 *   (b) bit-for-bit against oracle/_ref/libcvref.so, i.e. the reference's own sources
 *       compiled here by oracle/build_ref.sh, plus committed fixtures in tests/golden/
 *       generated from that library by tests/golden/make_golden.py.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (see oracle/Makefile).  -ffp-contract=off
 * matters: the reference's default build (baseline SSE3) has no FMA, every fp op rounds.
 *
 * Entry points use the argument lists of the HAL replacement hooks
 * (imgproc/src/hal_replacement.hpp:369,387,428,483-487,506-509,530-534).
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <limits.h>
#include <math.h>
#include <float.h>

typedef unsigned char uchar;
typedef unsigned short ushort;

/* core/include/opencv2/core/hal/interface.h:70-100 */
enum { D_8U = 0, D_8S = 1, D_16U = 2, D_16S = 3, D_32S = 4, D_32F = 5, D_64F = 6 };
#define MAT_DEPTH(t) ((t) & 7)
#define MAT_CN(t) ((((t) >> 3) & 511) + 1)
/* imgproc/include/opencv2/imgproc.hpp:249-292 ; core/base.hpp:268-281 */
enum { I_NEAREST = 0, I_LINEAR = 1, I_CUBIC = 2, I_AREA = 3, I_LANCZOS4 = 4, I_LINEAR_EXACT = 5, I_NEAREST_EXACT = 6 };
enum { B_CONSTANT = 0, B_REPLICATE = 1, B_REFLECT = 2, B_WRAP = 3, B_REFLECT_101 = 4, B_TRANSPARENT = 5 };
enum { WARP_RELATIVE_MAP = 32 };
#define INTER_BITS 5
#define INTER_TAB_SIZE 32
#define INTER_TAB_SIZE2 1024
#define INTER_REMAP_COEF_BITS 15
#define INTER_REMAP_COEF_SCALE (1 << 15)
#define INTER_RESIZE_COEF_BITS 11
#define INTER_RESIZE_COEF_SCALE 2048

enum { ORC_OK = 0, ORC_NOT_IMPLEMENTED = 1, ORC_ERR = -1 };

/* ---- rounding / saturation: core/fast_math.hpp:199-262,330-386 ; core/saturate.hpp:101-151 ---- */
/* cvRound == cvtsd2si / cvtss2si: nearest-even, "integer indefinite" 0x80000000 when out of range or NaN */
static int rne_d(double v) { double r = nearbyint(v); if (!(r >= -2147483648.0 && r <= 2147483647.0)) return INT_MIN; return (int)r; }
static int rne_f(float v)  { float r = nearbyintf(v); if (!(r >= -2147483648.0f && r < 2147483648.0f)) return INT_MIN; return (int)r; }
/* (int)value on x86 == cvttsd2si: same indefinite value when out of range */
static int trunc_d(double v) { if (!(v > -2147483649.0 && v < 2147483648.0)) return INT_MIN; return (int)v; }
static int trunc_f(float v)  { if (!(v >= -2147483648.0f && v < 2147483648.0f)) return INT_MIN; return (int)v; }
static int floor_d(double v) { int i = trunc_d(v); return i - (i > v); }   /* fast_math.hpp:252-254 */
static int floor_f(float v)  { int i = trunc_f(v); return i - (i > v); }   /* fast_math.hpp:365-386 */
static int ceil_d(double v)  { int i = trunc_d(v); return i + (i < v); }   /* fast_math.hpp:284-286 */

static uchar  sat_u8(int v)  { return (uchar)((unsigned)v <= 255 ? v : v > 0 ? 255 : 0); }          /* saturate.hpp:103 */
static ushort sat_u16(int v) { return (ushort)((unsigned)v <= 65535 ? v : v > 0 ? 65535 : 0); }     /* :123 */
static short  sat_s16(int v) { return (short)((unsigned)(v - SHRT_MIN) <= (unsigned)USHRT_MAX ? v : v > 0 ? SHRT_MAX : SHRT_MIN); } /* :131 */
static int clipi(int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; }                  /* imgwarp.cpp:259-262 */

/* core/src/copy.cpp:905-950 */
static int border_interpolate(int p, int len, int borderType)
{
    if ((unsigned)p < (unsigned)len) return p;
    if (borderType == B_REPLICATE) return p < 0 ? 0 : len - 1;
    if (borderType == B_REFLECT || borderType == B_REFLECT_101) {
        int delta = borderType == B_REFLECT_101;
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
    return -1; /* BORDER_CONSTANT */
}

/* ---- interpolation tables: imgwarp.cpp:66-240 (INTER_LINEAR part of initInterTab2D) ---- */
static float  BilinearTab_f[INTER_TAB_SIZE2][4];
static short  BilinearTab_i[INTER_TAB_SIZE2 + 2][4];   /* +2: the sum fix-up at :190-206 reads past the entry */
static uchar  NNDeltaTab_i[INTER_TAB_SIZE2][2];
static int    tabs_ready = 0;

static void init_tabs(void)
{
    if (tabs_ready) return;
    float tab1[INTER_TAB_SIZE * 2];
    const float scale = 1.f / INTER_TAB_SIZE;
    for (int i = 0; i < INTER_TAB_SIZE; i++) { float x = i * scale; tab1[i*2] = 1.f - x; tab1[i*2+1] = x; } /* :83-87,:137-141 */
    memset(BilinearTab_i, 0, sizeof(BilinearTab_i));
    const int ksize = 2;
    for (int i = 0; i < INTER_TAB_SIZE; i++)
        for (int j = 0; j < INTER_TAB_SIZE; j++) {
            float* tab = BilinearTab_f[i*INTER_TAB_SIZE + j];
            short* itab = BilinearTab_i[i*INTER_TAB_SIZE + j];
            int isum = 0;
            NNDeltaTab_i[i*INTER_TAB_SIZE + j][0] = j >= INTER_TAB_SIZE/2;   /* :176-177 */
            NNDeltaTab_i[i*INTER_TAB_SIZE + j][1] = i >= INTER_TAB_SIZE/2;
            for (int k1 = 0; k1 < ksize; k1++) {
                float vy = tab1[i*ksize + k1];
                for (int k2 = 0; k2 < ksize; k2++) {
                    float v = vy * tab1[j*ksize + k2];
                    tab[k1*ksize + k2] = v;
                    isum += itab[k1*ksize + k2] = sat_s16(rne_f(v * INTER_REMAP_COEF_SCALE));
                }
            }
            if (isum != INTER_REMAP_COEF_SCALE) {                             /* :190-206 (literal) */
                int diff = isum - INTER_REMAP_COEF_SCALE;
                int ksize2 = ksize/2, Mk1 = ksize2, Mk2 = ksize2, mk1 = ksize2, mk2 = ksize2;
                for (int k1 = ksize2; k1 < ksize2 + 2; k1++)
                    for (int k2 = ksize2; k2 < ksize2 + 2; k2++) {
                        if (itab[k1*ksize + k2] < itab[mk1*ksize + mk2]) mk1 = k1, mk2 = k2;
                        else if (itab[k1*ksize + k2] > itab[Mk1*ksize + Mk2]) Mk1 = k1, Mk2 = k2;
                    }
                if (diff < 0) itab[Mk1*ksize + Mk2] = (short)(itab[Mk1*ksize + Mk2] - diff);
                else itab[mk1*ksize + mk2] = (short)(itab[mk1*ksize + mk2] - diff);
            }
        }
    tabs_ready = 1;
}

/* exported for tests (table pin: entry 0 == {32767,0,0,1}) */
const short* orc_bilinear_tab_i(void) { init_tabs(); return &BilinearTab_i[0][0]; }
const float* orc_bilinear_tab_f(void) { init_tabs(); return &BilinearTab_f[0][0]; }

/* ---- INTER_CUBIC / INTER_LANCZOS4 tables: interpolateCubic :89-97, interpolateLanczos4 :99-127,
 *      initInterTab1D :129-151, initInterTab2D :154-206 (float product table, int16 table with the sum fix-up) ---- */
static float  BicubicTab_f[INTER_TAB_SIZE2][16];
static short  BicubicTab_i[INTER_TAB_SIZE2][16];
static float  Lanczos4Tab_f[INTER_TAB_SIZE2][64];
static short  Lanczos4Tab_i[INTER_TAB_SIZE2][64];
static int    hq_tabs_ready = 0;

static void coeffs_cubic(float x, float* c)
{
    const float A = -0.75f;
    c[0] = ((A*(x + 1) - 5*A)*(x + 1) + 8*A)*(x + 1) - 4*A;
    c[1] = ((A + 2)*x - (A + 3))*x*x + 1;
    c[2] = ((A + 2)*(1 - x) - (A + 3))*(1 - x)*(1 - x) + 1;
    c[3] = 1.f - c[0] - c[1] - c[2];
}

static void coeffs_lanczos4(float x, float* c)
{
    static const double s45 = 0.70710678118654752440084436210485;
    static const double PI = 3.1415926535897932384626433832795;
    const double cs[8][2] = {{1, 0}, {-s45, -s45}, {0, 1}, {s45, -s45}, {-1, 0}, {s45, s45}, {0, -1}, {-s45, s45}};
    if (x < FLT_EPSILON) {
        for (int i = 0; i < 8; i++) c[i] = 0;
        c[3] = 1;
        return;
    }
    float sum = 0;
    double y0 = -(x + 3)*PI*0.25, s0 = sin(y0), c0 = cos(y0);
    for (int i = 0; i < 8; i++) {
        double y = -(x + 3 - i)*PI*0.25;
        c[i] = (float)((cs[i][0]*s0 + cs[i][1]*c0)/(y*y));
        sum += c[i];
    }
    sum = 1.f/sum;
    for (int i = 0; i < 8; i++) c[i] *= sum;
}

static void init_tab2d(int ksize, float* tab, short* itab)
{
    float tab1[INTER_TAB_SIZE * 8];
    const float scale = 1.f / INTER_TAB_SIZE;
    for (int i = 0; i < INTER_TAB_SIZE; i++) {
        if (ksize == 4) coeffs_cubic(i*scale, tab1 + i*4);
        else coeffs_lanczos4(i*scale, tab1 + i*8);
    }
    for (int i = 0; i < INTER_TAB_SIZE; i++)
        for (int j = 0; j < INTER_TAB_SIZE; j++, tab += ksize*ksize, itab += ksize*ksize) {
            int isum = 0;
            for (int k1 = 0; k1 < ksize; k1++) {
                float vy = tab1[i*ksize + k1];
                for (int k2 = 0; k2 < ksize; k2++) {
                    float v = vy * tab1[j*ksize + k2];
                    tab[k1*ksize + k2] = v;
                    isum += itab[k1*ksize + k2] = sat_s16(rne_f(v * INTER_REMAP_COEF_SCALE));
                }
            }
            if (isum != INTER_REMAP_COEF_SCALE) {                             /* :190-206 */
                int diff = isum - INTER_REMAP_COEF_SCALE;
                int ksize2 = ksize/2, Mk1 = ksize2, Mk2 = ksize2, mk1 = ksize2, mk2 = ksize2;
                for (int k1 = ksize2; k1 < ksize2 + 2; k1++)
                    for (int k2 = ksize2; k2 < ksize2 + 2; k2++) {
                        if (itab[k1*ksize + k2] < itab[mk1*ksize + mk2]) mk1 = k1, mk2 = k2;
                        else if (itab[k1*ksize + k2] > itab[Mk1*ksize + Mk2]) Mk1 = k1, Mk2 = k2;
                    }
                if (diff < 0) itab[Mk1*ksize + Mk2] = (short)(itab[Mk1*ksize + Mk2] - diff);
                else itab[mk1*ksize + mk2] = (short)(itab[mk1*ksize + mk2] - diff);
            }
        }
}

static void init_hq_tabs(void)
{
    if (hq_tabs_ready) return;
    init_tab2d(4, &BicubicTab_f[0][0], &BicubicTab_i[0][0]);
    init_tab2d(8, &Lanczos4Tab_f[0][0], &Lanczos4Tab_i[0][0]);
    hq_tabs_ready = 1;
}

/* exported for tests: ksize 4 -> bicubic, 8 -> Lanczos4 */
const short* orc_interp_tab_i(int ksize) { init_hq_tabs(); return ksize == 4 ? &BicubicTab_i[0][0] : &Lanczos4Tab_i[0][0]; }
const float* orc_interp_tab_f(int ksize) { init_hq_tabs(); return ksize == 4 ? &BicubicTab_f[0][0] : &Lanczos4Tab_f[0][0]; }

/* ------------------------------------------------------------------------------------------
 * Sampling of one destination pixel given the decoded fixed-point coordinate.
 * NEAREST:  remapNearest<T>            imgwarp.cpp:268-369  (pure element copy, elemSize bytes)
 * LINEAR:   remapBilinear<...>         imgwarp.cpp:614-843
 *           8U : FixedPtCast<int,uchar,15> over BilinearTab_i      (:250-257, :1542)
 *           16U/16S/32F: fp32  ((v0*w0 + v1*w1) + v2*w2) + v3*w3  with BilinearTab_f (:675,:795,:837)
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    const uchar* src; size_t sstep; int sw, sh;
    int depth, cn, esz1;      /* esz1 = bytes per channel */
    int borderType;
    uchar cval[64];           /* border pixel, already saturate_cast<T>(borderValue[k&3]) (:278-279,:630-631) */
} Sampler;

static void sampler_init(Sampler* s, int src_type, const uchar* src, size_t sstep, int sw, int sh,
                         int borderType, const double bv[4])
{
    s->src = src; s->sstep = sstep; s->sw = sw; s->sh = sh;
    s->depth = MAT_DEPTH(src_type); s->cn = MAT_CN(src_type);
    s->esz1 = s->depth <= D_8S ? 1 : s->depth <= D_16S ? 2 : s->depth <= D_32F ? 4 : 8;
    s->borderType = borderType;
    for (int k = 0; k < s->cn && k < 16; k++) {
        double v = bv ? bv[k & 3] : 0.0;
        switch (s->depth) {
        case D_8U:  s->cval[k] = sat_u8(rne_d(v)); break;
        case D_16U: ((ushort*)s->cval)[k] = sat_u16(rne_d(v)); break;
        case D_16S: ((short*)s->cval)[k] = sat_s16(rne_d(v)); break;
        case D_32F: ((float*)s->cval)[k] = (float)v; break;
        case D_64F: if (k < 8) ((double*)s->cval)[k] = v; break;      /* saturate_cast<double> */
        default: break;
        }
    }
}

static void sample_nearest(const Sampler* s, int sx, int sy, uchar* D)
{
    const int esz = s->esz1 * s->cn;
    const uchar* S;
    if ((unsigned)sx < (unsigned)s->sw && (unsigned)sy < (unsigned)s->sh)
        S = s->src + (size_t)sy * s->sstep + (size_t)sx * esz;
    else if (s->borderType == B_REPLICATE) {
        sx = clipi(sx, 0, s->sw); sy = clipi(sy, 0, s->sh);
        S = s->src + (size_t)sy * s->sstep + (size_t)sx * esz;
    } else if (s->borderType == B_CONSTANT)
        S = s->cval;
    else if (s->borderType == B_TRANSPARENT)
        return;
    else {
        sx = border_interpolate(sx, s->sw, s->borderType);
        sy = border_interpolate(sy, s->sh, s->borderType);
        S = s->src + (size_t)sy * s->sstep + (size_t)sx * esz;
    }
    memcpy(D, S, esz);
}

static float ld_f(int depth, const uchar* p, int i)
{
    return depth == D_16U ? (float)((const ushort*)p)[i] : depth == D_16S ? (float)((const short*)p)[i] : ((const float*)p)[i];
}
static void st_f(int depth, uchar* D, int k, float t)
{
    if (depth == D_16U) ((ushort*)D)[k] = sat_u16(rne_f(t));
    else if (depth == D_16S) ((short*)D)[k] = sat_s16(rne_f(t));
    else ((float*)D)[k] = t;
}

static void sample_linear(const Sampler* s, int sx, int sy, int a, uchar* D)
{
    const int cn = s->cn, esz = s->esz1 * cn;
    const uchar *p0, *p1, *p2, *p3;
    /* inlier test (:634-638,:650-652); the 8UC3 SIMD variant uses W-2, which selects a different
       code path but never a different value */
    if ((unsigned)sx < (unsigned)(s->sw - 1 > 0 ? s->sw - 1 : 0) && (unsigned)sy < (unsigned)(s->sh - 1 > 0 ? s->sh - 1 : 0)) {
        p0 = s->src + (size_t)sy * s->sstep + (size_t)sx * esz;
        p1 = p0 + esz; p2 = p0 + s->sstep; p3 = p2 + esz;
    } else {
        if (s->borderType == B_TRANSPARENT && s->depth == D_64F) return;   /* not restated; callers decline this pair */
        if (s->borderType == B_TRANSPARENT) {           /* :726-755 */
            /* a point on the last row / column lacks neighbours: the available taps are renormalised;
               a point outside the image leaves the destination untouched */
            if (!(sx >= 0 && sx <= s->sw - 1 && sy >= 0 && sy <= s->sh - 1)) return;
            const int a1 = sx < s->sw - 1, a2 = sy < s->sh - 1;
            const uchar* S = s->src + (size_t)sy * s->sstep + (size_t)sx * esz;
            if (s->depth == D_8U) {
                const short* w = BilinearTab_i[a];
                int w_tot = w[0] + (a1 ? w[1] : 0) + (a2 ? w[2] : 0) + (a1 && a2 ? w[3] : 0);
                if (w_tot == 0) return;
                const int w_tot_ini = w[0] + w[1] + w[2] + w[3];
                for (int k = 0; k < cn; k++) {
                    int t0 = S[k] * w[0];
                    if (a1) t0 += S[k + cn] * w[1];
                    if (a2) t0 += S[s->sstep + k] * w[2];
                    if (a1 && a2) t0 += S[s->sstep + k + cn] * w[3];
                    t0 = trunc_f(t0 * (float)w_tot_ini / w_tot);
                    D[k] = sat_u8((t0 + (1 << (INTER_REMAP_COEF_BITS - 1))) >> INTER_REMAP_COEF_BITS);
                }
            } else {
                const float* w = BilinearTab_f[a];
                float w_tot = 0;
                w_tot += w[0]; if (a1) w_tot += w[1]; if (a2) w_tot += w[2]; if (a1 && a2) w_tot += w[3];
                if (w_tot == 0.f) return;
                const float w_tot_ini = w[0] + w[1] + w[2] + w[3];
                for (int k = 0; k < cn; k++) {
                    float t0 = 0;
                    t0 += ld_f(s->depth, S, k) * w[0];
                    if (a1) t0 += ld_f(s->depth, S, k + cn) * w[1];
                    if (a2) t0 += ld_f(s->depth, S + s->sstep, k) * w[2];
                    if (a1 && a2) t0 += ld_f(s->depth, S + s->sstep, k + cn) * w[3];
                    t0 = t0 * w_tot_ini / w_tot;
                    st_f(s->depth, D, k, t0);
                }
            }
            return;
        }
        if (s->borderType == B_CONSTANT && (sx >= s->sw || sx + 1 < 0 || sy >= s->sh || sy + 1 < 0)) { /* :755-763,:807-813 */
            memcpy(D, s->cval, esz);
            return;
        }
        int sx0, sx1, sy0, sy1;
        if (s->borderType == B_REPLICATE) {            /* :771-782 */
            sx0 = clipi(sx, 0, s->sw); sx1 = clipi(sx + 1, 0, s->sw);
            sy0 = clipi(sy, 0, s->sh); sy1 = clipi(sy + 1, 0, s->sh);
        } else {                                       /* :784-793 */
            sx0 = border_interpolate(sx, s->sw, s->borderType);
            sx1 = border_interpolate(sx + 1, s->sw, s->borderType);
            sy0 = border_interpolate(sy, s->sh, s->borderType);
            sy1 = border_interpolate(sy + 1, s->sh, s->borderType);
        }
        p0 = sx0 >= 0 && sy0 >= 0 ? s->src + (size_t)sy0 * s->sstep + (size_t)sx0 * esz : s->cval;
        p1 = sx1 >= 0 && sy0 >= 0 ? s->src + (size_t)sy0 * s->sstep + (size_t)sx1 * esz : s->cval;
        p2 = sx0 >= 0 && sy1 >= 0 ? s->src + (size_t)sy1 * s->sstep + (size_t)sx0 * esz : s->cval;
        p3 = sx1 >= 0 && sy1 >= 0 ? s->src + (size_t)sy1 * s->sstep + (size_t)sx1 * esz : s->cval;
    }
    if (s->depth == D_64F) {
        /* remapBilinear<Cast<double,double>, RemapNoVec, float> (imgwarp.cpp:1546): float table weights, WT = double */
        const float* w = BilinearTab_f[a];
        for (int k = 0; k < cn; k++)
            ((double*)D)[k] = ((const double*)p0)[k]*w[0] + ((const double*)p1)[k]*w[1] + ((const double*)p2)[k]*w[2] + ((const double*)p3)[k]*w[3];
        return;
    }
    if (s->depth == D_8U) {
        const short* w = BilinearTab_i[a];
        for (int k = 0; k < cn; k++) {
            int t = p0[k]*w[0] + p1[k]*w[1] + p2[k]*w[2] + p3[k]*w[3];
            D[k] = sat_u8((t + (1 << (INTER_REMAP_COEF_BITS - 1))) >> INTER_REMAP_COEF_BITS);
        }
    } else {
        const float* w = BilinearTab_f[a];
        for (int k = 0; k < cn; k++) {
            float v0, v1, v2, v3;
            if (s->depth == D_16U) { v0 = ((const ushort*)p0)[k]; v1 = ((const ushort*)p1)[k]; v2 = ((const ushort*)p2)[k]; v3 = ((const ushort*)p3)[k]; }
            else if (s->depth == D_16S) { v0 = ((const short*)p0)[k]; v1 = ((const short*)p1)[k]; v2 = ((const short*)p2)[k]; v3 = ((const short*)p3)[k]; }
            else { v0 = ((const float*)p0)[k]; v1 = ((const float*)p1)[k]; v2 = ((const float*)p2)[k]; v3 = ((const float*)p3)[k]; }
            float t = v0*w[0] + v1*w[1] + v2*w[2] + v3*w[3];     /* left-to-right, each op rounded */
            if (s->depth == D_16U) ((ushort*)D)[k] = sat_u16(rne_f(t));
            else if (s->depth == D_16S) ((short*)D)[k] = sat_s16(rne_f(t));
            else ((float*)D)[k] = t;
        }
    }
}

/* remapBicubic :846-949 (ksize 4, origin -1) and remapLanczos4 :952-1059 (ksize 8, origin -3).
 * 8U: int sums over the int16 table, FixedPtCast<int,uchar,15>; 16U/16S/32F: fp32, products added left to right,
 * one row at a time (sum += row expression), border pixels as cv + sum (S - cv)*w in tap order. */
static void sample_hq(const Sampler* s, int sx, int sy, int a, uchar* D, int ksize)
{
    const int cn = s->cn, esz = s->esz1 * cn, depth = s->depth;
    const int org = ksize/2 - 1;
    const short* wi = ksize == 4 ? BicubicTab_i[a] : Lanczos4Tab_i[a];
    const float* wf = ksize == 4 ? BicubicTab_f[a] : Lanczos4Tab_f[a];
    sx -= org; sy -= org;
    const unsigned width1 = s->sw - (ksize - 1) > 0 ? s->sw - (ksize - 1) : 0;
    const unsigned height1 = s->sh - (ksize - 1) > 0 ? s->sh - (ksize - 1) : 0;
    if ((unsigned)sx < width1 && (unsigned)sy < height1) {
        const uchar* S0 = s->src + (size_t)sy * s->sstep + (size_t)sx * esz;
        for (int k = 0; k < cn; k++) {
            if (depth == D_8U) {
                int sum = 0;
                for (int r = 0; r < ksize; r++)
                    for (int c = 0; c < ksize; c++) sum += S0[r * s->sstep + c*cn + k] * wi[r*ksize + c];
                D[k] = sat_u8((sum + (1 << (INTER_REMAP_COEF_BITS - 1))) >> INTER_REMAP_COEF_BITS);
            } else if (depth == D_64F) {                 /* Cast<double,double>, AT = float: double products and sums (:1564, :1588) */
                double sum = 0.;
                for (int r = 0; r < ksize; r++) {
                    const double* S = (const double*)(S0 + r * s->sstep);
                    double row = S[k] * wf[r*ksize];
                    for (int c = 1; c < ksize; c++) row = row + S[c*cn + k] * wf[r*ksize + c];
                    sum = (r == 0 && ksize == 4) ? row : sum + row;
                }
                ((double*)D)[k] = sum;
            } else {
                float sum = 0.f;
                for (int r = 0; r < ksize; r++) {
                    const uchar* S = S0 + r * s->sstep;
                    float row = ld_f(depth, S, k) * wf[r*ksize];
                    for (int c = 1; c < ksize; c++) row = row + ld_f(depth, S, c*cn + k) * wf[r*ksize + c];
                    /* bicubic: the first row initialises sum (:885); Lanczos4 starts from 0 and adds (:991-994) */
                    sum = (r == 0 && ksize == 4) ? row : sum + row;
                }
                st_f(depth, D, k, sum);
            }
        }
        return;
    }
    /* BORDER_TRANSPARENT: untouched when the centre sample lies outside, otherwise taps by REFLECT_101 (:862,:897-900) */
    if (s->borderType == B_TRANSPARENT && ((unsigned)(sx + org) >= (unsigned)s->sw || (unsigned)(sy + org) >= (unsigned)s->sh)) return;
    const int bt1 = s->borderType == B_TRANSPARENT ? B_REFLECT_101 : s->borderType;
    if (bt1 == B_CONSTANT && (sx >= s->sw || sx + ksize <= 0 || sy >= s->sh || sy + ksize <= 0)) {   /* :905-912 */
        memcpy(D, s->cval, esz);
        return;
    }
    int x[8], y[8];
    for (int i = 0; i < ksize; i++) {
        x[i] = border_interpolate(sx + i, s->sw, bt1);
        y[i] = border_interpolate(sy + i, s->sh, bt1);
    }
    for (int k = 0; k < cn; k++) {
        if (depth == D_8U) {
            int cv = s->cval[k], sum = cv * INTER_REMAP_COEF_SCALE;
            for (int i = 0; i < ksize; i++) {
                if (y[i] < 0) continue;
                const uchar* S = s->src + (size_t)y[i] * s->sstep;
                for (int c = 0; c < ksize; c++)
                    if (x[c] >= 0) sum += (S[x[c]*cn + k] - cv) * wi[i*ksize + c];
            }
            D[k] = sat_u8((sum + (1 << (INTER_REMAP_COEF_BITS - 1))) >> INTER_REMAP_COEF_BITS);
        } else if (depth == D_64F) {
            double cv = ((const double*)s->cval)[k], sum = cv * 1;
            for (int i = 0; i < ksize; i++) {
                if (y[i] < 0) continue;
                const double* S = (const double*)(s->src + (size_t)y[i] * s->sstep);
                for (int c = 0; c < ksize; c++)
                    if (x[c] >= 0) sum += (S[x[c]*cn + k] - cv) * wf[i*ksize + c];
            }
            ((double*)D)[k] = sum;
        } else {
            float cv = ld_f(depth, s->cval, k), sum = cv * 1;
            for (int i = 0; i < ksize; i++) {
                if (y[i] < 0) continue;
                const uchar* S = s->src + (size_t)y[i] * s->sstep;
                for (int c = 0; c < ksize; c++)
                    if (x[c] >= 0) sum += (ld_f(depth, S, x[c]*cn + k) - cv) * wf[i*ksize + c];
            }
            st_f(depth, D, k, sum);
        }
    }
}

/* one destination pixel for any non-nearest interpolation */
static void sample_interp(const Sampler* s, int interpolation, int sx, int sy, int a, uchar* D)
{
    if (interpolation == I_LINEAR) sample_linear(s, sx, sy, a, D);
    else sample_hq(s, sx, sy, a, D, interpolation == I_CUBIC ? 4 : 8);
}

static int check_type(int src_type, int interpolation)
{
    int depth = MAT_DEPTH(src_type), cn = MAT_CN(src_type);
    if (interpolation == I_NEAREST) return (depth <= D_64F && cn * 8 <= 64) ? 0 : 1;
    if (depth == D_64F) return ((interpolation == I_LINEAR || interpolation == I_CUBIC || interpolation == I_LANCZOS4) && cn <= 4) ? 0 : 1;
    if (!(depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F)) return 1;
    if (cn > 16) return 1;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * warpAffine: hal::warpAffine imgwarp.cpp:2250-2277, WarpAffineInvoker :2004-2048,
 * warpAffineBlockline[NN] :2280-2361.  M is the inverse (dst->src) matrix, as the hook gets it.
 * ------------------------------------------------------------------------------------------ */
int orc_warpAffine(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                   uchar* dst, size_t dst_step, int dw, int dh, const double M[6],
                   int interpolation, int borderType, const double bv[4])
{
    if (interpolation == I_AREA) interpolation = I_LINEAR;          /* :2395 (done above the hook) */
    if (interpolation != I_NEAREST && interpolation != I_LINEAR && interpolation != I_CUBIC && interpolation != I_LANCZOS4) return ORC_NOT_IMPLEMENTED;
    if (check_type(src_type, interpolation)) return ORC_NOT_IMPLEMENTED;
    init_tabs(); init_hq_tabs();
    Sampler s; sampler_init(&s, src_type, src, src_step, sw, sh, borderType, bv);
    const int esz = s.esz1 * s.cn;
    const int AB_BITS = 10, AB_SCALE = 1 << AB_BITS;
    const int round_delta = interpolation == I_NEAREST ? AB_SCALE/2 : AB_SCALE/INTER_TAB_SIZE/2;   /* :2011 */
    int* adelta = (int*)malloc(sizeof(int) * 2 * (size_t)(dw > 0 ? dw : 1));
    int* bdelta = adelta + dw;
    for (int x = 0; x < dw; x++) {                                   /* :2266-2270 */
        adelta[x] = rne_d(M[0]*x*AB_SCALE);
        bdelta[x] = rne_d(M[3]*x*AB_SCALE);
    }
    for (int y = 0; y < dh; y++) {
        /* :2030-2031 ; int adds wrap (CV_DISABLE_UBSAN :2279,:2310) -> do them in unsigned */
        int X0 = (int)((unsigned)rne_d((M[1]*y + M[2])*AB_SCALE) + (unsigned)round_delta);
        int Y0 = (int)((unsigned)rne_d((M[4]*y + M[5])*AB_SCALE) + (unsigned)round_delta);
        uchar* D = dst + (size_t)y * dst_step;
        for (int x = 0; x < dw; x++, D += esz) {
            int X = (int)((unsigned)X0 + (unsigned)adelta[x]);
            int Y = (int)((unsigned)Y0 + (unsigned)bdelta[x]);
            if (interpolation == I_NEAREST) {                        /* :2301-2307 */
                sample_nearest(&s, sat_s16(X >> AB_BITS), sat_s16(Y >> AB_BITS), D);
            } else {                                                 /* :2351-2358 */
                X >>= (AB_BITS - INTER_BITS); Y >>= (AB_BITS - INTER_BITS);
                int a = (Y & (INTER_TAB_SIZE-1))*INTER_TAB_SIZE + (X & (INTER_TAB_SIZE-1));
                sample_interp(&s, interpolation, sat_s16(X >> INTER_BITS), sat_s16(Y >> INTER_BITS), a, D);
            }
        }
    }
    free(adelta);
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * warpPerspective: hal::warpPerspective imgwarp.cpp:2744-2756, WarpPerspectiveInvoker :2689-2731,
 * warpPerspectiveBlockline[NN] :2758-2824 and the SSE4.1 line class imgwarp.sse4_1.cpp:176-460
 * (the variant an x86-64 host with SSE4.1 runs).  Coordinates are evaluated per 64-px wide block
 * relative to the block origin; the vector body covers the first (bw & ~15) pixels of a block
 * and the scalar tail the rest -- they differ only for NaN products (min/max operand order).
 * ------------------------------------------------------------------------------------------ */
int orc_warpPerspective(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                        uchar* dst, size_t dst_step, int dw, int dh, const double M[9],
                        int interpolation, int borderType, const double bv[4])
{
    if (interpolation == I_AREA) interpolation = I_LINEAR;          /* :2855 */
    if (interpolation != I_NEAREST && interpolation != I_LINEAR && interpolation != I_CUBIC && interpolation != I_LANCZOS4) return ORC_NOT_IMPLEMENTED;
    if (check_type(src_type, interpolation)) return ORC_NOT_IMPLEMENTED;
    init_tabs(); init_hq_tabs();
    Sampler s; sampler_init(&s, src_type, src, src_step, sw, sh, borderType, bv);
    const int esz = s.esz1 * s.cn;
    const int BLOCK_SZ = 32;
    int bh0 = BLOCK_SZ/2 < dh ? BLOCK_SZ/2 : dh;                     /* :2695-2697 */
    int bw0 = BLOCK_SZ*BLOCK_SZ/bh0 < dw ? BLOCK_SZ*BLOCK_SZ/bh0 : dw;
    for (int y = 0; y < dh; y++) {
        uchar* Drow = dst + (size_t)y * dst_step;
        for (int x = 0; x < dw; x += bw0) {
            int bw = bw0 < dw - x ? bw0 : dw - x;
            double X0 = M[0]*x + M[1]*y + M[2];                      /* :2712-2714 */
            double Y0 = M[3]*x + M[4]*y + M[5];
            double W0 = M[6]*x + M[7]*y + M[8];
            int nvec = bw & ~15;
            for (int x1 = 0; x1 < bw; x1++) {
                uchar* D = Drow + (size_t)(x + x1) * esz;
                double W = W0 + M[6]*x1;
                double num = interpolation == I_NEAREST ? 1. : (double)INTER_TAB_SIZE;
                W = W != 0 ? num / W : 0;                            /* NaN != 0 -> NaN, as both variants */
                double fX = (X0 + M[0]*x1)*W, fY = (Y0 + M[3]*x1)*W;
                if (x1 < nvec) {
                    /* _mm_max_pd(min, _mm_min_pd(max, v)): a NaN v propagates -> cvtpd2dq gives INT_MIN */
                    fX = fX != fX ? fX : (fX > (double)INT_MAX ? (double)INT_MAX : fX); fX = fX != fX ? fX : (fX < (double)INT_MIN ? (double)INT_MIN : fX);
                    fY = fY != fY ? fY : (fY > (double)INT_MAX ? (double)INT_MAX : fY); fY = fY != fY ? fY : (fY < (double)INT_MIN ? (double)INT_MIN : fY);
                } else {
                    /* std::max((double)INT_MIN, std::min((double)INT_MAX, v)): NaN v -> INT_MAX */
                    fX = fX < (double)INT_MAX ? fX : (double)INT_MAX; fX = (double)INT_MIN < fX ? fX : (double)INT_MIN;
                    fY = fY < (double)INT_MAX ? fY : (double)INT_MAX; fY = (double)INT_MIN < fY ? fY : (double)INT_MIN;
                }
                int X = rne_d(fX), Y = rne_d(fY);
                if (interpolation == I_NEAREST)
                    sample_nearest(&s, sat_s16(X), sat_s16(Y), D);
                else {
                    int a = (Y & (INTER_TAB_SIZE-1))*INTER_TAB_SIZE + (X & (INTER_TAB_SIZE-1));
                    sample_interp(&s, interpolation, sat_s16(X >> INTER_BITS), sat_s16(Y >> INTER_BITS), a, D);
                }
            }
        }
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * remap: cv::remap imgwarp.cpp:1519-1679, RemapInvoker map decoding :1094-1265.
 * kind 0: mapx,mapy CV_32FC1 (cv_hal_remap32f) ; 1: CV_32FC2 (cv_hal_remap32fc2) ;
 * kind 2: CV_16SC2 [+ CV_16UC1] (cv_hal_remap16s).
 * ------------------------------------------------------------------------------------------ */
static int remap_impl(int kind, int src_type, const uchar* src, size_t src_step, int sw, int sh,
                      uchar* dst, size_t dst_step, int dw, int dh,
                      const void* map1, size_t map1_step, const void* map2, size_t map2_step,
                      int interpolation, int borderType, const double bv[4])
{
    int relative = (interpolation & WARP_RELATIVE_MAP) != 0;        /* :1525 */
    interpolation &= ~WARP_RELATIVE_MAP;                             /* :1624-1626 */
    if (interpolation == I_AREA) interpolation = I_LINEAR;
    if (interpolation != I_NEAREST && interpolation != I_LINEAR && interpolation != I_CUBIC && interpolation != I_LANCZOS4) return ORC_NOT_IMPLEMENTED;
    if (check_type(src_type, interpolation)) return ORC_NOT_IMPLEMENTED;
    if (kind == 2 && interpolation != I_NEAREST && !map2) return ORC_NOT_IMPLEMENTED; /* reference reads the 16SC2 map as floats here */
    init_tabs(); init_hq_tabs();
    Sampler s; sampler_init(&s, src_type, src, src_step, sw, sh, borderType, bv);
    const int esz = s.esz1 * s.cn;
    for (int y = 0; y < dh; y++) {
        uchar* D = dst + (size_t)y * dst_step;
        const uchar* m1 = (const uchar*)map1 + (size_t)y * map1_step;
        const uchar* m2 = map2 ? (const uchar*)map2 + (size_t)y * map2_step : 0;
        for (int x = 0; x < dw; x++, D += esz) {
            int sx, sy, a = 0;
            if (kind == 2) {
                sx = ((const short*)m1)[x*2]; sy = ((const short*)m1)[x*2+1];
                if (interpolation == I_NEAREST) {
                    if (m2) {                                        /* :1109-1121 (short wrap) */
                        int aa = ((const ushort*)m2)[x] & (INTER_TAB_SIZE2-1);
                        sx = (short)(sx + NNDeltaTab_i[aa][0]); sy = (short)(sy + NNDeltaTab_i[aa][1]);
                    }
                } else a = ((const ushort*)m2)[x] & (INTER_TAB_SIZE2-1);   /* :1168-1185 */
            } else {
                float fx = kind == 0 ? ((const float*)m1)[x] : ((const float*)m1)[x*2];
                float fy = kind == 0 ? ((const float*)m2)[x] : ((const float*)m1)[x*2+1];
                if (interpolation == I_NEAREST) {                    /* :1123-1156: saturate_cast<short>(float) */
                    sx = sat_s16(rne_f(fx)); sy = sat_s16(rne_f(fy));
                } else {                                             /* :1186-1263 */
                    int isx = rne_f(fx * (float)INTER_TAB_SIZE), isy = rne_f(fy * (float)INTER_TAB_SIZE);
                    a = (isy & (INTER_TAB_SIZE-1))*INTER_TAB_SIZE + (isx & (INTER_TAB_SIZE-1));
                    sx = sat_s16(isx >> INTER_BITS); sy = sat_s16(isy >> INTER_BITS);
                }
            }
            if (relative) { sx += x; sy += y; }                     /* :293-299,:647-652 */
            if (interpolation == I_NEAREST) sample_nearest(&s, sx, sy, D);
            else sample_interp(&s, interpolation, sx, sy, a, D);
        }
    }
    return ORC_OK;
}

int orc_remap32f(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                 uchar* dst, size_t dst_step, int dw, int dh,
                 float* mapx, size_t mapx_step, float* mapy, size_t mapy_step,
                 int interpolation, int borderType, const double bv[4])
{ return remap_impl(0, src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, mapx, mapx_step, mapy, mapy_step, interpolation, borderType, bv); }

int orc_remap32fc2(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                   uchar* dst, size_t dst_step, int dw, int dh,
                   float* map, size_t map_step, int interpolation, int borderType, const double bv[4])
{ return remap_impl(1, src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, map, map_step, 0, 0, interpolation, borderType, bv); }

int orc_remap16s(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                 uchar* dst, size_t dst_step, int dw, int dh,
                 short* mapx, size_t mapx_step, ushort* mapy, size_t mapy_step,
                 int interpolation, int borderType, const double bv[4])
{ return remap_impl(2, src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, mapx, mapx_step, mapy, mapy_step, interpolation, borderType, bv); }

/* ------------------------------------------------------------------------------------------
 * cv::fisheye::initUndistortRectifyMap, calib3d/src/fisheye.cpp:577-631: the per-pixel loop, given f = (K00, K11),
 * c = (K02, K12), the distortion k[4] and iR = (P*R)^-1 (the reference's SVD-based inverse is the caller's business).
 * The row accumulators _x/_y/_w advance by REPEATED addition (:627-629), so column j is not _x0 + j*iR00 in floating
 * point; atan / sqrt are the host libm's, as in the reference.  m1type: CV_16SC2 (11) + CV_16UC1, or CV_32FC1 (5) twice.
 * ------------------------------------------------------------------------------------------ */
int orc_fisheye_initUndistortRectifyMap(const double* K9, const double* D4, const double* iR, int w, int h, int m1type,
                                        void* map1, size_t step1, void* map2, size_t step2)
{
    if (m1type <= 0) m1type = 11;
    if (m1type != 11 && m1type != 5) return ORC_NOT_IMPLEMENTED;
    const double f0 = K9[0], f1 = K9[4], c0 = K9[2], c1 = K9[5];
    double k[4] = {0, 0, 0, 0};
    if (D4) { k[0] = D4[0]; k[1] = D4[1]; k[2] = D4[2]; k[3] = D4[3]; }
    for (int i = 0; i < h; i++) {
        float* m1f = (float*)((uchar*)map1 + i * step1);
        float* m2f = (float*)((uchar*)map2 + i * step2);
        short* m1 = (short*)m1f; ushort* m2 = (ushort*)m2f;
        double _x = i * iR[1] + iR[2], _y = i * iR[4] + iR[5], _w = i * iR[7] + iR[8];
        for (int j = 0; j < w; j++) {
            double u, v;
            if (_w <= 0) {
                u = (_x > 0) ? -INFINITY : INFINITY;
                v = (_y > 0) ? -INFINITY : INFINITY;
            } else {
                double x = _x / _w, y = _y / _w;
                double r = sqrt(x * x + y * y);
                double theta = atan(r);
                double theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
                double theta_d = theta * (1 + k[0] * theta2 + k[1] * theta4 + k[2] * theta6 + k[3] * theta8);
                double scale = (r == 0) ? 1.0 : theta_d / r;
                u = f0 * x * scale + c0;
                v = f1 * y * scale + c1;
            }
            if (m1type == 11) {
                int iu = rne_d(u * 32.0), iv = rne_d(v * 32.0);      /* saturate_cast<int>(double) = cvRound */
                m1[j * 2] = (short)(iu >> 5); m1[j * 2 + 1] = (short)(iv >> 5);
                m2[j] = (ushort)((iv & 31) * 32 + (iu & 31));
            } else { m1f[j] = (float)u; m2f[j] = (float)v; }
            _x += iR[0]; _y += iR[3]; _w += iR[6];
        }
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * cv::initUndistortRectifyMap, calib3d/src/undistort.simd.hpp:102-277, as a stock x86 build executes it: the AVX2 dispatch
 * variant.  A row is walked in chunks of 8 pixels by the vector body (4 fp64 lanes x 2, v_muladd = FMA3: one rounding), the
 * last width % 8 pixels by the scalar tail (every operation rounded; this tree compiles with -ffp-contract=off).  They differ
 * in rounding, in how the row accumulators advance (body: += 8*ir per chunk plus the table ir*i; tail: += ir per pixel) and in the
 * CV_16SC2 pack (body saturates to int16, tail wraps).  Inputs: K (fx, fy, u0, v0 are read), dist = k1 k2 p1 p2 k3 k4 k5 k6 s1 s2
 * s3 s4, ir = (newK(:,0:3)*R)^-1 and the tilt matrix, both computed by the caller.  m1type 11 (CV_16SC2 + CV_16UC1), 5
 * (2 x CV_32FC1) or 13 (CV_32FC2).
 * ------------------------------------------------------------------------------------------ */
static void undist_store(int m1type, int body, int j, double u, double v, float* m1f, float* m2f)
{
    short* m1 = (short*)m1f; ushort* m2 = (ushort*)m2f;
    if (m1type == 11) {
        int iu = rne_d(u * 32.0), iv = rne_d(v * 32.0);                /* v_round / saturate_cast<int>: cvtpd2dq */
        if (body) { m1[j*2] = sat_s16(iu >> 5); m1[j*2+1] = sat_s16(iv >> 5); }      /* v_pack: packs_epi32 */
        else { m1[j*2] = (short)(iu >> 5); m1[j*2+1] = (short)(iv >> 5); }
        m2[j] = (ushort)((iv & 31) * 32 + (iu & 31));
    } else if (m1type == 5) { m1f[j] = (float)u; m2f[j] = (float)v; }
    else { m1f[j*2] = (float)u; m1f[j*2+1] = (float)v; }
}

int orc_initUndistortRectifyMap(const double* K9, const double* dist12, const double* ir, const double* tilt, int w, int h, int m1type,
                                void* map1, size_t step1, void* map2, size_t step2)
{
    if (m1type <= 0) m1type = 11;
    if (m1type != 11 && m1type != 5 && m1type != 13) return ORC_NOT_IMPLEMENTED;
    const double fx = K9[0], fy = K9[4], u0 = K9[2], v0 = K9[5];
    const double k1 = dist12[0], k2 = dist12[1], p1 = dist12[2], p2 = dist12[3], k3 = dist12[4], k4 = dist12[5], k5 = dist12[6],
                 k6 = dist12[7], s1 = dist12[8], s2 = dist12[9], s3 = dist12[10], s4 = dist12[11];
    double s_x[8], s_y[8], s_w[8];
    for (int i = 0; i < 8; i++) { s_x[i] = ir[0] * i; s_y[i] = ir[3] * i; s_w[i] = ir[6] * i; }      /* :93-98 */
    for (int i = 0; i < h; i++) {
        float* m1f = (float*)((uchar*)map1 + i * step1);
        float* m2f = map2 ? (float*)((uchar*)map2 + i * step2) : 0;
        double _x = i*ir[1] + ir[2], _y = i*ir[4] + ir[5], _w = i*ir[7] + ir[8];
        int j = 0;
        for (; j <= w - 8; j += 8, _x += 8 * ir[0], _y += 8 * ir[3], _w += 8 * ir[6]) {               /* :128 */
            for (int l = 0; l < 8; l++) {
                double m = 1.0 / (_w + s_w[l]);
                double x = (_x + s_x[l]) * m, y = (_y + s_y[l]) * m;
                double xd = x * x, yd = y * y, r2 = xd + yd;
                double kr = fma(fma(fma(k3, r2, k2), r2, k1), r2, 1.0) / fma(fma(fma(k6, r2, k5), r2, k4), r2, 1.0);
                xd = fma(2.0, xd, r2); yd = fma(2.0, yd, r2);
                double _2xy = (x * y) * 2.0;
                x = x * kr; y = y * kr;
                xd = fma(xd, p2, x); yd = fma(yd, p1, y);
                xd = fma(p1, _2xy, xd); yd = fma(p2, _2xy, yd);
                double r4 = r2 * r2;
                xd = fma(s1, r2, fma(s2, r4, xd));
                yd = fma(s3, r2, fma(s4, r4, yd));
                double X = fma(tilt[0], xd, fma(tilt[1], yd, tilt[2]));
                double Y = fma(tilt[3], xd, fma(tilt[4], yd, tilt[5]));
                double Z = fma(tilt[6], xd, fma(tilt[7], yd, tilt[8]));
                double inv = (Z == 0.0) ? 1.0 : 1.0 / Z;
                double u = fma(fx * inv, X, u0), v = fma(fy * inv, Y, v0);
                undist_store(m1type, 1, j + l, u, v, m1f, m2f);
            }
        }
        for (; j < w; j++, _x += ir[0], _y += ir[3], _w += ir[6]) {                                    /* :242-272 */
            double ww = 1./_w, x = _x*ww, y = _y*ww;
            double x2 = x*x, y2 = y*y;
            double r2 = x2 + y2, _2xy = 2*x*y;
            double kr = (1 + ((k3*r2 + k2)*r2 + k1)*r2)/(1 + ((k6*r2 + k5)*r2 + k4)*r2);
            double xd = (x*kr + p1*_2xy + p2*(r2 + 2*x2) + s1*r2+s2*r2*r2);
            double yd = (y*kr + p1*(r2 + 2*y2) + p2*_2xy + s3*r2+s4*r2*r2);
            double X = 0 + tilt[0]*xd; X += tilt[1]*yd; X += tilt[2]*1;                               /* Matx33d * Vec3d */
            double Y = 0 + tilt[3]*xd; Y += tilt[4]*yd; Y += tilt[5]*1;
            double Z = 0 + tilt[6]*xd; Z += tilt[7]*yd; Z += tilt[8]*1;
            double invProj = Z ? 1./Z : 1;
            double u = fx*invProj*X + u0, v = fy*invProj*Y + v0;
            undist_store(m1type, 0, j, u, v, m1f, m2f);
        }
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * cv::pyrDown, imgproc/src/pyramids.cpp:880-1036 (pyrDown_ + PyrDownInvoker) as the default x86 build runs it (universal
 * intrinsics at 128 bits, 4 float lanes, no FMA).  CV_8U / CV_16U / CV_16S accumulate in int and end with (v + 128) >> 8
 * (FixPtCast, :51-56) -- exact in any order.  CV_32F (FltCast, :58-63) is order sensitive, and the reference uses two orders:
 *   horizontal, vector body (:322-404: cn 1 and 4 in steps of 4 elements while 4 remain, cn 2 in steps of 8 while 8 remain,
 *   cn 3 in steps of 3 while 4 remain; all start after the first pixel): s2*6 + ((s1 + s3)*4 + (s0 + s4));
 *   horizontal, scalar (left border pixel, tail, right border, :955-1021): ((s2*6 + (s1 + s3)*4) + s0) + s4;
 *   vertical, vector body (:494-514, the first 4*floor(n/4) elements of a row): (((r1 + r3) + r2)*4 + ((r0 + r4) + (r2 + r2))) * (1/256);
 *   vertical, scalar tail (:1033-1034): (((r2*6 + (r1 + r3)*4) + r0) + r4) * (1/256).
 * Border taps follow borderInterpolate (tabL / tabR, :897-906; rows :947).  BORDER_CONSTANT is rejected (:1270). */
static int pyr_simd_h_count(int cn, int W)     /* elements the vector body of PyrDownVecH<float,float,cn> handles out of W */
{
    if (W < 0) W = 0;
    if (cn == 1 || cn == 4) return (W / 4) * 4;
    if (cn == 2) return (W / 8) * 8;
    if (cn == 3) return W >= 4 ? ((W - 4) / 3 + 1) * 3 : 0;
    return 0;
}

int orc_pyrDown(int type, const uchar* src, size_t src_step, int sw, int sh, uchar* dst, size_t dst_step, int dw, int dh, int borderType)
{
    const int depth = MAT_DEPTH(type), cn = MAT_CN(type);
    borderType &= ~16;                                     /* BORDER_ISOLATED */
    if (borderType == B_CONSTANT || cn > 4) return 1;
    if (!(depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F)) return 1;
    if (sw <= 0 || sh <= 0 || abs(dw*2 - sw) > 2 || abs(dh*2 - sh) > 2) return 1;
    int width0 = (sw - 2 - 1)/2 + 1; if (width0 > dw) width0 = dw;
    const int n = dw*cn, simd_h = pyr_simd_h_count(cn, width0*cn - cn), simd_v = (n/4)*4;
    int* ibuf = (int*)malloc(sizeof(int)*5*(size_t)n); float* fbuf = (float*)ibuf;
    if (!ibuf) return -1;
    for (int y = 0; y < dh; y++) {
        for (int k = 0; k < 5; k++) {
            const int sy = border_interpolate(2*y - 2 + k, sh, borderType);
            const uchar* srow = src + (size_t)sy*src_step;
            for (int x = 0; x < n; x++) {
                const int px = x/cn, c = x - px*cn;
                int t[5];
                for (int j = 0; j < 5; j++) {
                    int sx = 2*px - 2 + j;                                                /* interior taps are in range */
                    if (px < 1) sx = border_interpolate(sx, sw, borderType);             /* tabL, :899 */
                    /* tabR, :900 + :1015-1021: the table is indexed by (pixel - width0) + tap, so a SECOND right-hand pixel (only a
                     * non-default dsize has one) takes its taps one source pixel to the right of the first one's, not two */
                    else if (px >= width0) sx = border_interpolate(2*width0 - 2 + (px - width0) + j, sw, borderType);
                    t[j] = sx*cn + c;
                }
                if (depth == D_32F) {
                    const float* s = (const float*)srow;
                    const float s0 = s[t[0]], s1 = s[t[1]], s2 = s[t[2]], s3 = s[t[3]], s4 = s[t[4]];
                    const int vec = px >= 1 && px < width0 && (x - cn) < simd_h;
                    fbuf[k*n + x] = vec ? s2*6.f + ((s1 + s3)*4.f + (s0 + s4)) : ((s2*6.f + (s1 + s3)*4.f) + s0) + s4;
                } else {
                    int v[5];
                    for (int j = 0; j < 5; j++)
                        v[j] = depth == D_8U ? srow[t[j]] : depth == D_16U ? ((const ushort*)srow)[t[j]] : ((const short*)srow)[t[j]];
                    ibuf[k*n + x] = v[2]*6 + (v[1] + v[3])*4 + v[0] + v[4];
                }
            }
        }
        uchar* drow = dst + (size_t)y*dst_step;
        for (int x = 0; x < n; x++) {
            if (depth == D_32F) {
                const float r0 = fbuf[x], r1 = fbuf[n + x], r2 = fbuf[2*n + x], r3 = fbuf[3*n + x], r4 = fbuf[4*n + x];
                ((float*)drow)[x] = x < simd_v ? (((r1 + r3) + r2)*4.f + ((r0 + r4) + (r2 + r2))) * (1.f/256)
                                               : (((r2*6.f + (r1 + r3)*4.f) + r0) + r4) * (1.f/256);
            } else {
                const int v = (ibuf[2*n + x]*6 + (ibuf[n + x] + ibuf[3*n + x])*4 + ibuf[x] + ibuf[4*n + x] + 128) >> 8;
                if (depth == D_8U) drow[x] = (uchar)v; else if (depth == D_16U) ((ushort*)drow)[x] = (ushort)v; else ((short*)drow)[x] = (short)v;
            }
        }
    }
    free(ibuf);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * cv::pyrUp, imgproc/src/pyramids.cpp:1036-1155 (pyrUp_) + :1374-1416: every source pixel becomes a 2x2 block, filtered with the
 * even / odd phases of the 5-tap binomial kernel (1 6 1 and 4 4 per axis), result / 64.
 *   horizontal (:1074-1111, scalar in the reference): even column 2i = (s[i-1] + s[i]*6) + s[i+1], first = s[0]*6 + s[1]*2,
 *   last = s[W-2] + s[W-1]*7; odd column 2i+1 = (s[i] + s[i+1])*4, last = s[W-1]*8; a one-pixel row = s*8 twice; a destination one
 *   column wider than 2W repeats the last odd column (:1099-1102).
 *   rows (:1075): source row of filtered row sy = borderInterpolate(2*sy, 2*H, REFLECT_101)/2 -- row 1 above the top, H-1 below.
 *   vertical (:1121-1140): even row 2y = cast((h[y-1] + h[y]*6) + h[y+1]), odd row 2y+1 = cast((h[y] + h[y+1])*4); a destination
 *   one row taller than 2H repeats row 2H-2 (:1144-1153); one row shorter drops the last odd row.
 * cast = (v + 32) >> 6 for CV_8U / CV_16S / CV_16U (FixPtCast, :51-56; the packed vector bodies :586-706 never saturate on these
 * ranges), v*(1/64) for CV_32F / CV_64F (FltCast).  The float vector body (:708-726, the first 4*floor(n/4) elements of a row pair)
 * writes the odd row as (h[y] + h[y+1])*(1/16) -- the same value unless the *4 overflows -- and the even row in the scalar order.
 * A one-pixel-wide source with a 3-column destination reads an unwritten buffer cell in the reference (:1080-1085 `continue`s
 * before :1099): not restated. */
int orc_pyrUp(int type, const uchar* src, size_t src_step, int sw, int sh, uchar* dst, size_t dst_step, int dw, int dh, int borderType)
{
    const int depth = MAT_DEPTH(type), cn = MAT_CN(type);
    if (borderType != B_REFLECT_101 || cn > 4) return 1;                                          /* :1378 CV_Assert(BORDER_DEFAULT) */
    if (!(depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F || depth == D_64F)) return 1;
    if (sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0 || abs(dw - sw*2) != dw % 2 || abs(dh - sh*2) != dh % 2) return 1;   /* :1055-1056 */
    if (sw == 1 && dw == 3) return 1;
    const int n = dw*cn, simd_v = depth == D_32F ? (n/4)*4 : 0;
    double* hbuf = (double*)malloc(sizeof(double)*3*(size_t)n);       /* int / float / double accumulators, all exact in a double */
    if (!hbuf) return -1;
    for (int y = 0; y < sh; y++) {
        for (int k = 0; k < 3; k++) {
            const int sy = border_interpolate((y - 1 + k)*2, sh*2, B_REFLECT_101)/2;
            const uchar* srow = src + (size_t)sy*src_step;
            for (int x = 0; x < n; x++) {
                int p = x/cn; const int c = x - p*cn;
                if (p >= 2*sw) p = 2*sw - 1;
                const int i = p >> 1, im = i > 0 ? i - 1 : 0, ip = i < sw - 1 ? i + 1 : sw - 1;
                double a, b, d;                                        /* s[i-1], s[i], s[i+1] */
                #define PU_LD(j) (depth == D_8U ? (double)srow[(j)*cn + c] : depth == D_16U ? (double)((const ushort*)srow)[(j)*cn + c] : \
                                  depth == D_16S ? (double)((const short*)srow)[(j)*cn + c] : depth == D_32F ? (double)((const float*)srow)[(j)*cn + c] : \
                                  ((const double*)srow)[(j)*cn + c])
                a = PU_LD(im); b = PU_LD(i); d = PU_LD(ip);
                #undef PU_LD
                double v;
                if (depth == D_32F) {
                    const float fa = (float)a, fb = (float)b, fd = (float)d;
                    if (sw == 1) v = fb*8.f;
                    else if (!(p & 1)) v = i == 0 ? (float)(fb*6.f) + (float)(fd*2.f) : i == sw - 1 ? fa + (float)(fb*7.f) : (float)(fa + (float)(fb*6.f)) + fd;
                    else v = i == sw - 1 ? fb*8.f : (float)(fb + fd)*4.f;
                    v = (float)v;
                } else {                                               /* int (exact) or double, same expression order */
                    if (sw == 1) v = b*8;
                    else if (!(p & 1)) v = i == 0 ? b*6 + d*2 : i == sw - 1 ? a + b*7 : (a + b*6) + d;
                    else v = i == sw - 1 ? b*8 : (b + d)*4;
                }
                hbuf[k*n + x] = v;
            }
        }
        for (int half = 0; half < 2; half++) {
            const int Y = 2*y + half;
            if (Y >= dh) break;
            uchar* drow = dst + (size_t)Y*dst_step;
            for (int x = 0; x < n; x++) {
                const double h0 = hbuf[x], h1 = hbuf[n + x], h2 = hbuf[2*n + x];
                if (depth == D_32F) {
                    const float f0 = (float)h0, f1 = (float)h1, f2 = (float)h2;
                    const int vec = x < simd_v && 2*y + 1 < dh;        /* PyrUpVecV only; PyrUpVecVOneRow has no odd row */
                    ((float*)drow)[x] = !half ? (float)((float)(f0 + (float)(f1*6.f)) + f2) * (1.f/64)
                                              : vec ? (1.f/16) * (float)(f1 + f2) : (float)((float)(f1 + f2)*4.f) * (1.f/64);
                } else if (depth == D_64F) {
                    ((double*)drow)[x] = !half ? ((h0 + h1*6) + h2) * (1./64) : ((h1 + h2)*4) * (1./64);
                } else {
                    const int v = ((int)(!half ? h0 + h1*6 + h2 : (h1 + h2)*4) + 32) >> 6;
                    if (depth == D_8U) drow[x] = (uchar)v; else if (depth == D_16U) ((ushort*)drow)[x] = (ushort)v; else ((short*)drow)[x] = (short)v;
                }
            }
        }
    }
    free(hbuf);
    if (dh > sh*2) memcpy(dst + (size_t)(sh*2)*dst_step, dst + (size_t)(sh*2 - 2)*dst_step, (size_t)n * (depth == D_8U ? 1 : depth == D_32F ? 4 : depth == D_64F ? 8 : 2));
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * cv::warpPolar, forward direction (imgproc/src/imgwarp.cpp:3163-3224): the CV_32FC1 maps of a dw x dh polar image -- column = radius,
 * row = angle.  rho table (:3190-3204): (float)(rho * (maxRadius / dw)), or (float)(exp(rho * (log(maxRadius) / dw)) - 1) with
 * WARP_POLAR_LOG; per row (:3208-3210) cos / sin of (2*pi / dh) * phi in double from the host libm; per entry (:3216-3220)
 * (float)(rho_table[rho] * cp + center.x), (float)(rho_table[rho] * sp + center.y), multiply and add rounded separately in double.
 * The image itself is cv::remap over these maps with BORDER_CONSTANT (WARP_FILL_OUTLIERS) or BORDER_TRANSPARENT (:3223).
 * WARP_INVERSE_MAP (:3225-3281: copyMakeBorder + cartToPolar + cv::log) is not restated. */
int orc_warpPolar_maps(int dw, int dh, float cx, float cy, double maxRadius, int semilog, float* mapx, size_t xstep, float* mapy, size_t ystep)
{
    if (dw <= 0 || dh <= 0) return 1;
    const double Kangle = 6.283185307179586476925286766559 / dh;                 /* CV_2PI */
    float* rhos = (float*)malloc(sizeof(float)*(size_t)dw);
    if (!rhos) return -1;
    if (semilog) { const double Kmag = log(maxRadius) / dw; for (int r = 0; r < dw; r++) rhos[r] = (float)(exp(r * Kmag) - 1.0); }
    else { const double Kmag = maxRadius / dw; for (int r = 0; r < dw; r++) rhos[r] = (float)(r * Kmag); }
    for (int phi = 0; phi < dh; phi++) {
        const double KKy = Kangle * phi, cp = cos(KKy), sp = sin(KKy);
        float* mx = (float*)((uchar*)mapx + (size_t)phi*xstep); float* my = (float*)((uchar*)mapy + (size_t)phi*ystep);
        for (int r = 0; r < dw; r++) {
            const double x = rhos[r] * cp + cx, y = rhos[r] * sp + cy;
            mx[r] = (float)x; my[r] = (float)y;
        }
    }
    free(rhos);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * The map loops of the stitching module's rotation warpers: RotationWarperBase<P>::buildMaps / PlaneWarper::buildMaps
 * (stitching/include/opencv2/stitching/detail/warpers_inl.hpp:74-99, stitching/src/warpers.cpp:192-243) with
 * P::mapBackward for the plane (:235-248), spherical (:263-281) and cylindrical (:296-314) projectors -- float arithmetic in the
 * reference's order, sinf / cosf from the host libm.  kind 0 / 1 / 2; k_rinv, t, scale = the projector's members after
 * setCameraParams (warpers.cpp:128-161); (tl_x, tl_y) + size = the result ROI the warper's detectResultRoi found on the host. */
int orc_stitch_buildMaps(int kind, const float* k, const float* t, float scale, int tl_x, int tl_y, int w, int h,
                         float* xmap, size_t xstep, float* ymap, size_t ystep)
{
    if (kind < 0 || kind > 2) return ORC_NOT_IMPLEMENTED;
    const float pi = (float)3.1415926535897932384626433832795;
    for (int j = 0; j < h; j++) {
        float* xr = (float*)((uchar*)xmap + j*xstep); float* yr = (float*)((uchar*)ymap + j*ystep);
        for (int i = 0; i < w; i++) {
            float u = (float)(tl_x + i), v = (float)(tl_y + j), x, y, z;
            if (kind == 0) {
                u = u / scale - t[0]; v = v / scale - t[1];
                x = k[0]*u + k[1]*v + k[2]*(1 - t[2]);
                y = k[3]*u + k[4]*v + k[5]*(1 - t[2]);
                z = k[6]*u + k[7]*v + k[8]*(1 - t[2]);
                x /= z; y /= z;
            } else {
                u /= scale; v /= scale;
                float x_, y_, z_;
                if (kind == 1) { float sinv = sinf(pi - v); x_ = sinv * sinf(u); y_ = cosf(pi - v); z_ = sinv * cosf(u); }
                else { x_ = sinf(u); y_ = v; z_ = cosf(u); }
                x = k[0]*x_ + k[1]*y_ + k[2]*z_;
                y = k[3]*x_ + k[4]*y_ + k[5]*z_;
                z = k[6]*x_ + k[7]*y_ + k[8]*z_;
                if (z > 0) { x /= z; y /= z; } else x = y = -1;
            }
            xr[i] = x; yr[i] = y;
        }
    }
    return ORC_OK;
}

/* cv::Mat::inv(DECOMP_LU) of a 3x3 CV_64F matrix = cv::invert's closed form, core/src/lapack.cpp:711-713 (det3), :946-968 */
static int inv3(const double* m, double* t)
{
    double d = m[0]*(m[4]*m[8] - m[5]*m[7]) - m[1]*(m[3]*m[8] - m[5]*m[6]) + m[2]*(m[3]*m[7] - m[4]*m[6]);
    if (d == 0.) { for (int i = 0; i < 9; i++) t[i] = 0.; return 0; }
    d = 1./d;
    t[0] = (m[4]*m[8] - m[5]*m[7]) * d;  t[1] = (m[2]*m[7] - m[1]*m[8]) * d;  t[2] = (m[1]*m[5] - m[2]*m[4]) * d;
    t[3] = (m[5]*m[6] - m[3]*m[8]) * d;  t[4] = (m[0]*m[8] - m[2]*m[6]) * d;  t[5] = (m[2]*m[3] - m[0]*m[5]) * d;
    t[6] = (m[3]*m[7] - m[4]*m[6]) * d;  t[7] = (m[1]*m[6] - m[0]*m[7]) * d;  t[8] = (m[0]*m[4] - m[1]*m[3]) * d;
    return 1;
}

static void matx33_mul(const double* a, const double* b, double* c)   /* core/matx.inl.hpp:471-481: s = 0; s += a(i,k)*b(k,j) */
{
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += a[i*3+k] * b[k*3+j];
            c[i*3+j] = s;
        }
}

/* the two matrices cv::initUndistortRectifyMap derives before its pixel loop (calib3d/src/undistort.dispatch.cpp:123, :155-157):
 * iR = (newK(:,0:3) * R)^-1 -- the product by cv::gemm's 3x3 form (core/src/matmul.simd.hpp:1470-1481: a0*b0 + a1*b3 + a2*b6,
 * then *alpha(1) + 0*beta(0)), the inverse by inv3 -- and the tilt projection matrix of (tauX, tauY)
 * (calib3d/src/distortion_model.hpp:75-94; identity when both are 0, :154).  R9 may be NULL (identity). */
int orc_undistort_params(const double* newK9, const double* R9, double tauX, double tauY, double* iR, double* tilt)
{
    static const double I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    const double* b = R9 ? R9 : I3;
    double p[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double t = newK9[i*3]*b[j] + newK9[i*3+1]*b[3+j] + newK9[i*3+2]*b[6+j];
            p[i*3+j] = t*1.0 + 0.0*0.0;
        }
    inv3(p, iR);
    if (tilt) {
        memcpy(tilt, I3, sizeof I3);
        if (tauX != 0 || tauY != 0) {
            double cx = cos(tauX), sx = sin(tauX), cy = cos(tauY), sy = sin(tauY);
            double rx[9] = {1, 0, 0, 0, cx, sx, 0, -sx, cx}, ry[9] = {cy, 0, -sy, 0, 1, 0, sy, 0, cy}, rxy[9];
            matx33_mul(ry, rx, rxy);
            double pz[9] = {rxy[8], 0, -rxy[2], 0, rxy[8], -rxy[5], 0, 0, 1};
            matx33_mul(pz, rxy, tilt);
        }
    }
    return 0;
}

/* cv::undistort, calib3d/src/undistort.dispatch.cpp:284-328: the image is taken in stripes of max(1, 4096 / cols) rows; each stripe
 * builds its own CV_16SC2 map from a new camera matrix whose v0 is lowered by the stripe's first row (so every stripe inverts its
 * own matrix and starts its row accumulators from 0 -- NOT the entries the whole-image map has) and remaps INTER_LINEAR /
 * BORDER_CONSTANT(0).  newK9 may be NULL (the camera matrix); map1 (w*4 bytes per row) / map2 (w*2) receive the stripes' maps. */
int orc_undistort(int type, const uchar* src, size_t src_step, int w, int h, uchar* dst, size_t dst_step,
                  const double* K9, const double* dist12, const double* newK9, const double* tilt,
                  short* map1, size_t step1, ushort* map2, size_t step2)
{
    int s0 = 4096 / (w > 1 ? w : 1); if (s0 < 1) s0 = 1; if (s0 > h) s0 = h;
    double Ar[9], iR[9];
    const double bv[4] = {0, 0, 0, 0};
    memcpy(Ar, newK9 ? newK9 : K9, sizeof Ar);
    const double v0 = Ar[5];
    for (int y = 0; y < h; y += s0) {
        int ss = s0 < h - y ? s0 : h - y;
        Ar[5] = v0 - y;
        orc_undistort_params(Ar, NULL, 0, 0, iR, NULL);
        int rc = orc_initUndistortRectifyMap(K9, dist12, iR, tilt, w, ss, 11, (float*)((uchar*)map1 + y*step1), step1, (float*)((uchar*)map2 + y*step2), step2);
        if (rc) return rc;
    }
    return orc_remap16s(type, src, src_step, w, h, dst, dst_step, w, h, map1, step1, map2, step2, 1, 0, bv);
}

/* ------------------------------------------------------------------------------------------
 * cv::convertMaps imgwarp.cpp:1682-1986 (the scalar tails; the SIMD bodies compute the same values)
 * ------------------------------------------------------------------------------------------ */
int orc_convertMaps(int m1type, const void* map1, size_t s1, int m2type, const void* map2, size_t s2,
                    int dtype, void* dst1, size_t ds1, void* dst2, size_t ds2, int w, int h, int nn)
{
    enum { T_16UC1 = 2, T_16SC1 = 3, T_32FC1 = 5, T_16SC2 = 11, T_32FC2 = 13 };
    if (!map2) m2type = -1;
    if (m2type == T_16SC2) { const void* t = map1; map1 = map2; map2 = t; size_t ts = s1; s1 = s2; s2 = ts; int tt = m1type; m1type = m2type; m2type = tt; }
    if (!((m1type == T_16SC2 && (nn || m2type == T_16UC1 || m2type == T_16SC1)) ||
          (m1type == T_32FC1 && m2type == T_32FC1) || (m1type == T_32FC2 && m2type == -1))) return ORC_NOT_IMPLEMENTED;
    if (dtype <= 0) dtype = m1type == T_16SC2 ? T_32FC2 : T_16SC2;
    if (dtype != T_16SC2 && dtype != T_32FC1 && dtype != T_32FC2) return ORC_NOT_IMPLEMENTED;
    const int want2 = !nn && dtype != T_32FC2;
    if (want2 && !dst2) return ORC_NOT_IMPLEMENTED;
    if (m1type == T_16SC2 && dtype == T_32FC1 && nn) return ORC_NOT_IMPLEMENTED;
    const int have_frac = m1type == T_16SC2 && (m2type == T_16UC1 || m2type == T_16SC1);
    const float scale = 1.f / INTER_TAB_SIZE;
    for (int y = 0; y < h; y++) {
        const float* a1f = (const float*)((const uchar*)map1 + y * s1);
        const float* a2f = map2 ? (const float*)((const uchar*)map2 + y * s2) : 0;
        const short* a1 = (const short*)a1f; const ushort* a2 = (const ushort*)a2f;
        float* d1f = (float*)((uchar*)dst1 + y * ds1);
        float* d2f = want2 ? (float*)((uchar*)dst2 + y * ds2) : 0;
        short* d1 = (short*)d1f; ushort* d2 = (ushort*)d2f;
        for (int x = 0; x < w; x++) {
            if (m1type == dtype) {                                             /* :1718-1725 */
                if (m1type == T_16SC2) { d1[x*2] = a1[x*2]; d1[x*2+1] = a1[x*2+1]; if (want2 && m2type == T_16UC1) d2[x] = a2[x]; }
                else if (m1type == T_32FC1) { d1f[x] = a1f[x]; if (want2) d2f[x] = a2f[x]; }
                else { d1f[x*2] = a1f[x*2]; d1f[x*2+1] = a1f[x*2+1]; }
            } else if (m1type == T_32FC1 && dtype == T_32FC2) { d1f[x*2] = a1f[x]; d1f[x*2+1] = a2f[x]; }
            else if (m1type == T_32FC2 && dtype == T_32FC1) { d1f[x] = a1f[x*2]; if (want2) d2f[x] = a1f[x*2+1]; }   /* nn: the y plane is dropped */
            else if (dtype == T_16SC2) {
                float fx = m1type == T_32FC1 ? a1f[x] : a1f[x*2], fy = m1type == T_32FC1 ? a2f[x] : a1f[x*2+1];
                if (nn) { d1[x*2] = sat_s16(rne_f(fx)); d1[x*2+1] = sat_s16(rne_f(fy)); }              /* :1794-1798 */
                else {                                                                                /* :1836-1842 */
                    int ix = rne_f(fx * (float)INTER_TAB_SIZE), iy = rne_f(fy * (float)INTER_TAB_SIZE);
                    d1[x*2] = sat_s16(ix >> INTER_BITS); d1[x*2+1] = sat_s16(iy >> INTER_BITS);
                    d2[x] = (ushort)((iy & (INTER_TAB_SIZE-1))*INTER_TAB_SIZE + (ix & (INTER_TAB_SIZE-1)));
                }
            } else {                                                                                  /* :1945-1950,:1974-1979 */
                int fxy = (have_frac && !(nn && dtype == T_32FC2)) ? a2[x] & (INTER_TAB_SIZE2-1) : 0;
                float fx = a1[x*2] + (fxy & (INTER_TAB_SIZE-1))*scale, fy = a1[x*2+1] + (fxy >> INTER_BITS)*scale;
                if (dtype == T_32FC1) { d1f[x] = fx; d2f[x] = fy; } else { d1f[x*2] = fx; d1f[x*2+1] = fy; }
            }
        }
    }
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * resize: cv::hal::resize resize.cpp:3634-4000
 * ------------------------------------------------------------------------------------------ */
#define LD(depth, p, i) ((depth) == D_8U ? (float)((const uchar*)(p))[i] : (depth) == D_16U ? (float)((const ushort*)(p))[i] : \
                         (depth) == D_16S ? (float)((const short*)(p))[i] : ((const float*)(p))[i])
#define LDI(depth, p, i) ((depth) == D_8U ? (int)((const uchar*)(p))[i] : (depth) == D_16U ? (int)((const ushort*)(p))[i] : (int)((const short*)(p))[i])

static void store_sat_f(int depth, uchar* D, int i, float v)  /* saturate_cast<T>(float) */
{
    switch (depth) {
    case D_8U:  D[i] = sat_u8(rne_f(v)); break;
    case D_16U: ((ushort*)D)[i] = sat_u16(rne_f(v)); break;
    case D_16S: ((short*)D)[i] = sat_s16(rne_f(v)); break;
    default:    ((float*)D)[i] = v; break;
    }
}

/* resizeNN resize.cpp:1121-1172, resizeNNInvoker :1039-1080 */
static void resize_nn(int esz, const uchar* src, size_t src_step, int sw, int sh,
                      uchar* dst, size_t dst_step, int dw, int dh, double fx, double fy)
{
    double ifx = 1./fx, ify = 1./fy;
    for (int y = 0; y < dh; y++) {
        int sy = floor_d(y*ify); if (sy > sh - 1) sy = sh - 1;
        const uchar* S = src + (size_t)sy * src_step;
        uchar* D = dst + (size_t)y * dst_step;
        for (int x = 0; x < dw; x++) {
            int sx = floor_d(x*ifx); if (sx > sw - 1) sx = sw - 1;
            memcpy(D + (size_t)x*esz, S + (size_t)sx*esz, esz);
        }
    }
}

/* resizeAreaFast_Invoker resize.cpp:2919-3059 ; ResizeAreaFastVec :2937-2987 ; ResizeAreaFastVec_SIMD_32f :2857-2908 */
static void resize_area_fast(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                             uchar* dst, size_t dst_step, int dw, int dh, int scale_x, int scale_y)
{
    const int area = scale_x * scale_y;
    const float scale = 1.f / area;
    const int dwidth1 = (sw / scale_x) * cn;
    const int dwidth = dw * cn, swidth = sw * cn;
    const int esz1 = depth == D_8U ? 1 : (depth == D_32F ? 4 : 2);
    const int is_int_wt = depth == D_8U;                           /* areafast_tab :3745-3755: WT=int for 8U, float otherwise */
    const int fast2x2 = scale_x == 2 && scale_y == 2 && (cn == 1 || cn == 3 || cn == 4);
    for (int dy = 0; dy < dh; dy++) {
        uchar* D = dst + (size_t)dy * dst_step;
        int sy0 = dy * scale_y;
        int w = sy0 + scale_y <= sh ? dwidth1 : 0;
        if (sy0 >= sh) { memset(D, 0, (size_t)dwidth * esz1); continue; }
        const uchar* Srow = src + (size_t)sy0 * src_step;
        int dx = 0;
        if (depth != D_32F && fast2x2) {
            /* (a+b+c+d+2)>>2 in int for 8U/16U/16S (:2950-2980), SIMD bodies identical */
            const uchar* Snext = Srow + src_step;
            for (; dx < w; dx++) {
                int c = dx % cn, px = dx / cn, idx = px*2*cn + c;
                int v = (LDI(depth, Srow, idx) + LDI(depth, Srow, idx + cn) + LDI(depth, Snext, idx) + LDI(depth, Snext, idx + cn) + 2) >> 2;
                if (depth == D_8U) D[dx] = (uchar)v; else if (depth == D_16U) ((ushort*)D)[dx] = (ushort)v; else ((short*)D)[dx] = (short)v;
            }
        } else if (depth == D_32F && scale_x == 2 && scale_y == 2 && (cn == 1 || cn == 4)) {
            /* SIMD body (4 lanes at the SSE3 baseline): ((r00+r01)+(r10+r11))*0.25f for dx < (w & ~3) */
            const float* S0 = (const float*)Srow; const float* S1 = (const float*)(Srow + src_step);
            int wv = w & ~3;
            for (; dx < wv; dx++) {
                int c = dx % cn, px = dx / cn, idx = px*2*cn + c;
                ((float*)D)[dx] = ((S0[idx] + S0[idx + cn]) + (S1[idx] + S1[idx + cn])) * 0.25f;
            }
        }
        for (; dx < w; dx++) {                                      /* generic body :3010-3026 */
            int c = dx % cn, px = dx / cn;
            const uchar* S = Srow + (size_t)(px*scale_x*cn + c) * esz1;  /* xofs[dx] = scale_x*px*cn + c (:3836-3842) */
            int k = 0;
            if (is_int_wt) {
                int sum = 0;
                for (int sy = 0; sy < scale_y; sy++) for (int sx = 0; sx < scale_x; sx++)
                    sum += S[(size_t)sy*src_step + (size_t)sx*cn];
                D[dx] = sat_u8(rne_f((float)sum * scale));
            } else {
                float sum = 0;
                /* ofs[k] enumerates (sy,sx) row-major (:3832-3834); unrolled by 4 (CV_ENABLE_UNROLLED):
                   the 4-tuple is added in int for 16U/16S (integer promotion) and in float for 32F */
                #define OFS(kk) ((size_t)((kk)/scale_x)*src_step + (size_t)(((kk)%scale_x)*cn)*esz1)
                for (; k <= area - 4; k += 4) {
                    if (depth == D_32F)
                        sum += *(const float*)(S + OFS(k)) + *(const float*)(S + OFS(k+1)) + *(const float*)(S + OFS(k+2)) + *(const float*)(S + OFS(k+3));
                    else
                        sum += (float)(LDI(depth, S + OFS(k), 0) + LDI(depth, S + OFS(k+1), 0) + LDI(depth, S + OFS(k+2), 0) + LDI(depth, S + OFS(k+3), 0));
                }
                for (; k < area; k++)
                    sum += LD(depth, S + OFS(k), 0);
                #undef OFS
                store_sat_f(depth, D, dx, sum * scale);
            }
        }
        for (; dx < dwidth; dx++) {                                 /* ragged right/bottom edge :3028-3050 */
            int c = dx % cn, px = dx / cn;
            int sx0 = px*scale_x*cn + c, count = 0;
            float fsum = 0; int isum = 0;
            for (int sy = 0; sy < scale_y; sy++) {
                if (sy0 + sy >= sh) break;
                const uchar* S = src + (size_t)(sy0 + sy) * src_step;
                for (int sx = 0; sx < scale_x*cn; sx += cn) {
                    if (sx0 + sx >= swidth) break;
                    if (is_int_wt) isum += S[sx0 + sx]; else fsum += LD(depth, S, sx0 + sx);
                    count++;
                }
            }
            float q = (is_int_wt ? (float)isum : fsum) / count;
            store_sat_f(depth, D, dx, q);
        }
    }
}

/* computeResizeAreaTab resize.cpp:3334-3373 */
typedef struct { int si, di; float alpha; } DecimateAlpha;
static int compute_area_tab(int ssize, int dsize, int cn, double scale, DecimateAlpha* tab)
{
    int k = 0;
    for (int dx = 0; dx < dsize; dx++) {
        double fsx1 = dx * scale, fsx2 = fsx1 + scale;
        double cellWidth = scale < ssize - fsx1 ? scale : ssize - fsx1;
        int sx1 = ceil_d(fsx1), sx2 = floor_d(fsx2);
        sx2 = sx2 < ssize - 1 ? sx2 : ssize - 1;
        sx1 = sx1 < sx2 ? sx1 : sx2;
        if (sx1 - fsx1 > 1e-3) { tab[k].di = dx*cn; tab[k].si = (sx1 - 1)*cn; tab[k++].alpha = (float)((sx1 - fsx1) / cellWidth); }
        for (int sx = sx1; sx < sx2; sx++) { tab[k].di = dx*cn; tab[k].si = sx*cn; tab[k++].alpha = (float)(1.0 / cellWidth); }
        if (fsx2 - sx2 > 1e-3) {
            double m = fsx2 - sx2 < 1. ? fsx2 - sx2 : 1.; m = m < cellWidth ? m : cellWidth;
            tab[k].di = dx*cn; tab[k].si = sx2*cn; tab[k++].alpha = (float)(m / cellWidth);
        }
    }
    return k;
}

/* ResizeArea_Invoker resize.cpp:3183-3304 (WT=float for 8U/16U/16S/32F, area_tab :3757-3762) */
static int resize_area(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                       uchar* dst, size_t dst_step, int dw, int dh, double scale_x, double scale_y)
{
    DecimateAlpha* xtab = (DecimateAlpha*)malloc(sizeof(DecimateAlpha) * 2 * (size_t)(sw + sh));
    DecimateAlpha* ytab = xtab + sw*2;
    int xtab_size = compute_area_tab(sw, dw, cn, scale_x, xtab);
    int ytab_size = compute_area_tab(sh, dh, 1, scale_y, ytab);
    const int dwidth = dw * cn;
    float* buf = (float*)malloc(sizeof(float) * 2 * (size_t)dwidth);
    float* sum = buf + dwidth;
    int prev_dy = ytab[0].di;
    for (int dx = 0; dx < dwidth; dx++) sum[dx] = 0;
    for (int j = 0; j < ytab_size; j++) {
        float beta = ytab[j].alpha; int dy = ytab[j].di, sy = ytab[j].si;
        const uchar* S = src + (size_t)sy * src_step;
        for (int dx = 0; dx < dwidth; dx++) buf[dx] = 0;
        for (int k = 0; k < xtab_size; k++) {
            int sxn = xtab[k].si, dxn = xtab[k].di; float alpha = xtab[k].alpha;
            for (int c = 0; c < cn; c++) buf[dxn + c] = buf[dxn + c] + LD(depth, S, sxn + c)*alpha;
        }
        if (dy != prev_dy) {
            uchar* D = dst + (size_t)prev_dy * dst_step;
            for (int dx = 0; dx < dwidth; dx++) store_sat_f(depth, D, dx, sum[dx]);
            for (int dx = 0; dx < dwidth; dx++) sum[dx] = beta * buf[dx];
            prev_dy = dy;
        } else
            for (int dx = 0; dx < dwidth; dx++) sum[dx] += beta * buf[dx];
    }
    {
        uchar* D = dst + (size_t)prev_dy * dst_step;
        for (int dx = 0; dx < dwidth; dx++) store_sat_f(depth, D, dx, sum[dx]);
    }
    free(buf); free(xtab);
    return ORC_OK;
}

/* generic 2-tap path: coefficient tables resize.cpp:3876-3996, HResizeLinear :1877-1928,
   VResizeLinear :1931-1989 (8U specialisation :1961-1989), row choice resizeGeneric_Invoker :2205-2232 */
static int resize_linear(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                         uchar* dst, size_t dst_step, int dw, int dh,
                         double inv_scale_x, double inv_scale_y, int area_mode)
{
    const double scale_x = 1./inv_scale_x, scale_y = 1./inv_scale_y;
    const int fixpt = depth == D_8U;
    int xmax = dw;
    int* xofs = (int*)malloc(sizeof(int) * (size_t)dw);
    float* alpha = (float*)malloc(sizeof(float) * 2 * (size_t)dw);
    short* ialpha = (short*)malloc(sizeof(short) * 2 * (size_t)dw);
    for (int dx = 0; dx < dw; dx++) {
        float fx; int sx;
        if (!area_mode) { fx = (float)((dx + 0.5)*scale_x - 0.5); sx = floor_f(fx); fx -= sx; }
        else { sx = floor_d(dx*scale_x); fx = (float)((dx + 1) - (sx + 1)*inv_scale_x); fx = fx <= 0 ? 0.f : fx - floor_f(fx); }
        if (sx < 0) { fx = 0; sx = 0; }                              /* :3918-3923 (ksize2-1 == 0) */
        if (sx + 1 >= sw) { xmax = xmax < dx ? xmax : dx; if (sx >= sw - 1) { fx = 0; sx = sw - 1; } }  /* :3925-3930 */
        xofs[dx] = sx;
        float c0 = 1.f - fx, c1 = fx;
        alpha[dx*2] = c0; alpha[dx*2+1] = c1;
        ialpha[dx*2] = sat_s16(rne_f(c0 * INTER_RESIZE_COEF_SCALE)); ialpha[dx*2+1] = sat_s16(rne_f(c1 * INTER_RESIZE_COEF_SCALE));
    }
    const int dwidth = dw * cn;
    if (depth == D_64F) {
        /* HResizeLinear<double,double,float> / VResizeLinear<double,double,float,Cast<double,double>> (resize.cpp:3759-3769 table):
           float coefficients, double products and sums */
        double* drow0 = (double*)malloc(sizeof(double) * 2 * (size_t)dwidth); double* drow1 = drow0 + dwidth;
        for (int dy = 0; dy < dh; dy++) {
            float fy; int sy;
            if (!area_mode) { fy = (float)((dy + 0.5)*scale_y - 0.5); sy = floor_f(fy); fy -= sy; }
            else { sy = floor_d(dy*scale_y); fy = (float)((dy + 1) - (sy + 1)*inv_scale_y); fy = fy <= 0 ? 0.f : fy - floor_f(fy); }
            float b0 = 1.f - fy, b1 = fy;
            const double* S0 = (const double*)(src + (size_t)clipi(sy, 0, sh) * src_step);
            const double* S1 = (const double*)(src + (size_t)clipi(sy + 1, 0, sh) * src_step);
            double* D = (double*)(dst + (size_t)dy * dst_step);
            for (int dx = 0; dx < dw; dx++) for (int c = 0; c < cn; c++) {
                int sxi = xofs[dx]*cn + c, o = dx*cn + c;
                if (dx < xmax) {
                    drow0[o] = S0[sxi]*alpha[dx*2] + S0[sxi + cn]*alpha[dx*2+1];
                    drow1[o] = S1[sxi]*alpha[dx*2] + S1[sxi + cn]*alpha[dx*2+1];
                } else { drow0[o] = S0[sxi]*1.f; drow1[o] = S1[sxi]*1.f; }
            }
            for (int o = 0; o < dwidth; o++) D[o] = drow0[o]*b0 + drow1[o]*b1;
        }
        free(drow0); free(ialpha); free(alpha); free(xofs);
        return ORC_OK;
    }
    float* frow0 = (float*)malloc(sizeof(float) * 2 * (size_t)dwidth); float* frow1 = frow0 + dwidth;
    int* irow0 = (int*)frow0; int* irow1 = (int*)frow1;
    for (int dy = 0; dy < dh; dy++) {
        float fy; int sy;
        if (!area_mode) { fy = (float)((dy + 0.5)*scale_y - 0.5); sy = floor_f(fy); fy -= sy; }
        else { sy = floor_d(dy*scale_y); fy = (float)((dy + 1) - (sy + 1)*inv_scale_y); fy = fy <= 0 ? 0.f : fy - floor_f(fy); }
        float b0 = 1.f - fy, b1 = fy;
        short ib0 = sat_s16(rne_f(b0 * INTER_RESIZE_COEF_SCALE)), ib1 = sat_s16(rne_f(b1 * INTER_RESIZE_COEF_SCALE));
        const uchar* S0 = src + (size_t)clipi(sy, 0, sh) * src_step;       /* :2211 with ksize2 == 1 */
        const uchar* S1 = src + (size_t)clipi(sy + 1, 0, sh) * src_step;
        uchar* D = dst + (size_t)dy * dst_step;
        for (int dx = 0; dx < dw; dx++) for (int c = 0; c < cn; c++) {
            int sxi = xofs[dx]*cn + c, o = dx*cn + c;
            if (fixpt) {
                if (dx < xmax) {
                    irow0[o] = S0[sxi]*ialpha[dx*2] + S0[sxi + cn]*ialpha[dx*2+1];
                    irow1[o] = S1[sxi]*ialpha[dx*2] + S1[sxi + cn]*ialpha[dx*2+1];
                } else { irow0[o] = S0[sxi]*INTER_RESIZE_COEF_SCALE; irow1[o] = S1[sxi]*INTER_RESIZE_COEF_SCALE; }
            } else {
                if (dx < xmax) {
                    frow0[o] = LD(depth, S0, sxi)*alpha[dx*2] + LD(depth, S0, sxi + cn)*alpha[dx*2+1];
                    frow1[o] = LD(depth, S1, sxi)*alpha[dx*2] + LD(depth, S1, sxi + cn)*alpha[dx*2+1];
                } else { frow0[o] = LD(depth, S0, sxi)*1.f; frow1[o] = LD(depth, S1, sxi)*1.f; }
            }
        }
        for (int o = 0; o < dwidth; o++) {
            if (fixpt)
                D[o] = (uchar)((((ib0 * (irow0[o] >> 4)) >> 16) + ((ib1 * (irow1[o] >> 4)) >> 16) + 2) >> 2);
            else
                store_sat_f(depth, D, o, frow0[o]*b0 + frow1[o]*b1);
        }
    }
    free(frow0); free(ialpha); free(alpha); free(xofs);
    return ORC_OK;
}

/* CV_64F INTER_AREA: resizeAreaFast_<double, double, ResizeAreaFastNoVec> (no vector body; the window sum in double, four values at a
 * time, times the FLOAT scale 1.f/area; the ragged edge divides (float)sum by the count, resize.cpp:3010-3050) and
 * resizeArea_<double, double> (float table weights, double products and sums, :3183-3304). */
static void resize_area_fast_f64(int cn, const uchar* src, size_t src_step, int sw, int sh,
                                 uchar* dst, size_t dst_step, int dw, int dh, int scale_x, int scale_y)
{
    const int area = scale_x * scale_y;
    const float scale = 1.f / area;
    const int dwidth1 = (sw / scale_x) * cn, dwidth = dw * cn, swidth = sw * cn;
    for (int dy = 0; dy < dh; dy++) {
        double* D = (double*)(dst + (size_t)dy * dst_step);
        int sy0 = dy * scale_y;
        int w = sy0 + scale_y <= sh ? dwidth1 : 0;
        if (sy0 >= sh) { for (int dx = 0; dx < dwidth; dx++) D[dx] = 0; continue; }
        int dx = 0;
        for (; dx < w; dx++) {
            int c = dx % cn, px = dx / cn;
            const uchar* S = src + (size_t)sy0 * src_step + (size_t)(px*scale_x*cn + c) * 8;
            #define AT(kk) (*(const double*)(S + (size_t)((kk)/scale_x)*src_step + (size_t)(((kk)%scale_x)*cn)*8))
            double sum = 0; int k = 0;
            for (; k <= area - 4; k += 4) sum += AT(k) + AT(k+1) + AT(k+2) + AT(k+3);
            for (; k < area; k++) sum += AT(k);
            #undef AT
            D[dx] = sum * scale;
        }
        for (; dx < dwidth; dx++) {
            int c = dx % cn, px = dx / cn, sx0 = px*scale_x*cn + c, count = 0;
            double sum = 0;
            for (int sy = 0; sy < scale_y; sy++) {
                if (sy0 + sy >= sh) break;
                const double* S = (const double*)(src + (size_t)(sy0 + sy) * src_step);
                for (int sx = 0; sx < scale_x*cn; sx += cn) {
                    if (sx0 + sx >= swidth) break;
                    sum += S[sx0 + sx]; count++;
                }
            }
            D[dx] = (double)((float)sum / count);
        }
    }
}

static int resize_area_f64(int cn, const uchar* src, size_t src_step, int sw, int sh,
                           uchar* dst, size_t dst_step, int dw, int dh, double scale_x, double scale_y)
{
    DecimateAlpha* xtab = (DecimateAlpha*)malloc(sizeof(DecimateAlpha) * 2 * (size_t)(sw + sh));
    DecimateAlpha* ytab = xtab + sw*2;
    int xtab_size = compute_area_tab(sw, dw, cn, scale_x, xtab);
    int ytab_size = compute_area_tab(sh, dh, 1, scale_y, ytab);
    const int dwidth = dw * cn;
    double* buf = (double*)malloc(sizeof(double) * 2 * (size_t)dwidth);
    double* sum = buf + dwidth;
    int prev_dy = ytab[0].di;
    for (int dx = 0; dx < dwidth; dx++) sum[dx] = 0;
    for (int j = 0; j < ytab_size; j++) {
        double beta = ytab[j].alpha; int dy = ytab[j].di, sy = ytab[j].si;
        const double* S = (const double*)(src + (size_t)sy * src_step);
        for (int dx = 0; dx < dwidth; dx++) buf[dx] = 0;
        for (int k = 0; k < xtab_size; k++) {
            int sxn = xtab[k].si, dxn = xtab[k].di; double alpha = xtab[k].alpha;
            for (int c = 0; c < cn; c++) buf[dxn + c] = buf[dxn + c] + S[sxn + c]*alpha;
        }
        if (dy != prev_dy) {
            memcpy(dst + (size_t)prev_dy * dst_step, sum, sizeof(double) * dwidth);
            for (int dx = 0; dx < dwidth; dx++) sum[dx] = beta * buf[dx];
            prev_dy = dy;
        } else
            for (int dx = 0; dx < dwidth; dx++) sum[dx] += beta * buf[dx];
    }
    memcpy(dst + (size_t)prev_dy * dst_step, sum, sizeof(double) * dwidth);
    free(buf); free(xtab);
    return ORC_OK;
}

/* INTER_CUBIC: resize.cpp:3884-3999 (coefficients), HResizeCubic :1993-2041, VResizeCubic :2044-2062 and its
 * SSE vector bodies VResizeCubicVec_32s8u :1410-1443, _32f16u :1445-1466, _32f16s :1468-1489, _32f :1491-1508.
 * The vector body covers the first (width & ~7) elements of a row (8U/16U/16S; width & ~3 for 32F) and the scalar
 * tail the rest; they order the additions differently (body: S0*b0 + (S1*b1 + (S2*b2 + S3*b3)), 8U in fp32 with
 * beta/2^22; tail: ((S0*b0 + S1*b1) + S2*b2) + S3*b3, 8U in int with (sum + 2^21) >> 22), so both are restated. */
static int resize_cubic(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                        uchar* dst, size_t dst_step, int dw, int dh, double inv_scale_x, double inv_scale_y)
{
    const double scale_x = 1./inv_scale_x, scale_y = 1./inv_scale_y;
    const int fixpt = depth == D_8U;
    const int dwidth = dw * cn, swidth = sw * cn;
    int* xofs = (int*)malloc(sizeof(int) * (size_t)dw);
    float* alpha = (float*)malloc(sizeof(float) * 4 * (size_t)dw);
    short* ialpha = (short*)malloc(sizeof(short) * 4 * (size_t)dw);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5)*scale_x - 0.5); int sx = floor_f(fx); fx -= sx;
        xofs[dx] = sx;
        coeffs_cubic(fx, alpha + dx*4);
        for (int k = 0; k < 4; k++) ialpha[dx*4 + k] = sat_s16(rne_f(alpha[dx*4 + k] * INTER_RESIZE_COEF_SCALE));
    }
    float* frows = (float*)malloc(sizeof(float) * 4 * (size_t)dwidth);
    int* irows = (int*)frows;
    const int lanes = depth == D_32F ? 4 : 8;
    const int nbody = dwidth - dwidth % lanes;
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5)*scale_y - 0.5); int sy = floor_f(fy); fy -= sy;
        float beta[4]; short ibeta[4];
        coeffs_cubic(fy, beta);
        for (int k = 0; k < 4; k++) ibeta[k] = sat_s16(rne_f(beta[k] * INTER_RESIZE_COEF_SCALE));
        for (int k = 0; k < 4; k++) {                                      /* :2205-2211 rows sy-1 .. sy+2, clipped */
            const uchar* S = src + (size_t)clipi(sy - 1 + k, 0, sh) * src_step;
            for (int o = 0; o < dwidth; o++) {
                int dx = o / cn, c = o % cn, sxe = xofs[dx]*cn + c - cn;
                int iv = 0; float fv = 0;
                for (int j = 0; j < 4; j++) {
                    int sxj = sxe + j*cn;
                    if ((unsigned)sxj >= (unsigned)swidth) { while (sxj < 0) sxj += cn; while (sxj >= swidth) sxj -= cn; }
                    if (fixpt) iv += S[sxj] * ialpha[dx*4 + j];
                    else fv += LD(depth, S, sxj) * alpha[dx*4 + j];
                }
                if (fixpt) irows[k*dwidth + o] = iv; else frows[k*dwidth + o] = fv;
            }
        }
        uchar* D = dst + (size_t)dy * dst_step;
        for (int o = 0; o < dwidth; o++) {
            if (fixpt) {
                int S0 = irows[o], S1 = irows[dwidth + o], S2 = irows[2*dwidth + o], S3 = irows[3*dwidth + o];
                if (o < nbody) {
                    const float scale = 1.f/(INTER_RESIZE_COEF_SCALE*INTER_RESIZE_COEF_SCALE);
                    float b0 = ibeta[0]*scale, b1 = ibeta[1]*scale, b2 = ibeta[2]*scale, b3 = ibeta[3]*scale;
                    float t = (float)S0*b0 + ((float)S1*b1 + ((float)S2*b2 + (float)S3*b3));
                    D[o] = sat_u8(sat_s16(rne_f(t)));                      /* v_round, v_pack, v_pack_u */
                } else
                    D[o] = sat_u8((S0*ibeta[0] + S1*ibeta[1] + S2*ibeta[2] + S3*ibeta[3] + (1 << 21)) >> 22);
            } else {
                float S0 = frows[o], S1 = frows[dwidth + o], S2 = frows[2*dwidth + o], S3 = frows[3*dwidth + o];
                float t = o < nbody ? S0*beta[0] + (S1*beta[1] + (S2*beta[2] + S3*beta[3]))
                                    : S0*beta[0] + S1*beta[1] + S2*beta[2] + S3*beta[3];
                if (depth == D_16U && o < nbody) ((ushort*)D)[o] = sat_u16(rne_f(t));          /* v_pack_u(int32) */
                else if (depth == D_16S && o < nbody) ((short*)D)[o] = sat_s16(rne_f(t));
                else store_sat_f(depth, D, o, t);
            }
        }
    }
    free(frows); free(ialpha); free(alpha); free(xofs);
    return ORC_OK;
}

/* INTER_LANCZOS4: resize.cpp:3884-3999 (coefficients: interpolateLanczos4 :974-1003, fp64 sin/cos of the host libm),
 * HResizeLanczos4 :2066-2116 (taps sx-3..sx+4, clamped), rows sy-3..sy+4 clamped (:2205-2211), VResizeLanczos4 :2120-2155.
 * Vertical pass: 8U int, (sum + 2^21) >> 22; 16U: the SSE4.1 body (resize.sse4_1.cpp:189-226) and both tails add left to
 * right; 16S (first width & ~7 elements) and 32F (first width & ~3) have a vector body that nests from the right
 * (:1562-1626): b0*S0 + (b1*S1 + (... + b7*S7)). */
static int resize_lanczos4(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                           uchar* dst, size_t dst_step, int dw, int dh, double inv_scale_x, double inv_scale_y)
{
    const double scale_x = 1./inv_scale_x, scale_y = 1./inv_scale_y;
    const int fixpt = depth == D_8U;
    const int dwidth = dw * cn, swidth = sw * cn;
    int* xofs = (int*)malloc(sizeof(int) * (size_t)dw);
    float* alpha = (float*)malloc(sizeof(float) * 8 * (size_t)dw);
    short* ialpha = (short*)malloc(sizeof(short) * 8 * (size_t)dw);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5)*scale_x - 0.5); int sx = floor_f(fx); fx -= sx;
        xofs[dx] = sx;
        coeffs_lanczos4(fx, alpha + dx*8);
        for (int k = 0; k < 8; k++) ialpha[dx*8 + k] = sat_s16(rne_f(alpha[dx*8 + k] * INTER_RESIZE_COEF_SCALE));
    }
    float* frows = (float*)malloc(sizeof(float) * 8 * (size_t)dwidth);
    int* irows = (int*)frows;
    const int nbody = depth == D_16S ? dwidth - dwidth % 8 : depth == D_32F ? dwidth - dwidth % 4 : 0;
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5)*scale_y - 0.5); int sy = floor_f(fy); fy -= sy;
        float beta[8]; short ibeta[8];
        coeffs_lanczos4(fy, beta);
        for (int k = 0; k < 8; k++) ibeta[k] = sat_s16(rne_f(beta[k] * INTER_RESIZE_COEF_SCALE));
        for (int k = 0; k < 8; k++) {
            const uchar* S = src + (size_t)clipi(sy - 3 + k, 0, sh) * src_step;
            for (int o = 0; o < dwidth; o++) {
                int dx = o / cn, c = o % cn, sxe = xofs[dx]*cn + c - 3*cn;
                int iv = 0; float fv = 0;
                for (int j = 0; j < 8; j++) {
                    int sxj = sxe + j*cn;
                    if ((unsigned)sxj >= (unsigned)swidth) { while (sxj < 0) sxj += cn; while (sxj >= swidth) sxj -= cn; }
                    if (fixpt) iv += S[sxj] * ialpha[dx*8 + j];
                    else fv += LD(depth, S, sxj) * alpha[dx*8 + j];
                }
                if (fixpt) irows[k*dwidth + o] = iv; else frows[k*dwidth + o] = fv;
            }
        }
        uchar* D = dst + (size_t)dy * dst_step;
        for (int o = 0; o < dwidth; o++) {
            if (fixpt) {
                int sum = 0;
                for (int k = 0; k < 8; k++) sum += irows[k*dwidth + o] * ibeta[k];
                D[o] = sat_u8((sum + (1 << 21)) >> 22);
            } else {
                float t;
                if (o < nbody) {
                    t = frows[7*dwidth + o] * beta[7];
                    for (int k = 6; k >= 0; k--) t = frows[k*dwidth + o] * beta[k] + t;
                } else {
                    t = frows[o] * beta[0];
                    for (int k = 1; k < 8; k++) t = t + frows[k*dwidth + o] * beta[k];
                }
                store_sat_f(depth, D, o, t);
            }
        }
    }
    free(frows); free(ialpha); free(alpha); free(xofs);
    return ORC_OK;
}

/* CV_64F INTER_CUBIC / INTER_LANCZOS4: resizeGeneric_<HResizeCubic<double,double,float>, VResizeCubic<double,double,float,
 * Cast<double,double>, VResizeNoVec>> and the Lanczos4 pair (resize.cpp cubic_tab / lanczos4_tab 64F rows): float coefficients, double
 * products and sums, taps and rows clamped as for the other depths, the vertical pass scalar (no vector body for 64F): left to right. */
static int resize_hq_f64(int ks, int cn, const uchar* src, size_t src_step, int sw, int sh,
                         uchar* dst, size_t dst_step, int dw, int dh, double inv_scale_x, double inv_scale_y)
{
    const double scale_x = 1./inv_scale_x, scale_y = 1./inv_scale_y;
    const int dwidth = dw * cn, swidth = sw * cn, org = ks/2 - 1;
    int* xofs = (int*)malloc(sizeof(int) * (size_t)dw);
    float* alpha = (float*)malloc(sizeof(float) * ks * (size_t)dw);
    double* rows = (double*)malloc(sizeof(double) * ks * (size_t)dwidth);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5)*scale_x - 0.5); int sx = floor_f(fx); fx -= sx;
        xofs[dx] = sx;
        if (ks == 4) coeffs_cubic(fx, alpha + dx*4); else coeffs_lanczos4(fx, alpha + dx*8);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5)*scale_y - 0.5); int sy = floor_f(fy); fy -= sy;
        float beta[8];
        if (ks == 4) coeffs_cubic(fy, beta); else coeffs_lanczos4(fy, beta);
        for (int k = 0; k < ks; k++) {
            const double* S = (const double*)(src + (size_t)clipi(sy - org + k, 0, sh) * src_step);
            for (int o = 0; o < dwidth; o++) {
                int dx = o / cn, c = o % cn, sxe = xofs[dx]*cn + c - org*cn;
                double v = 0;
                for (int j = 0; j < ks; j++) {
                    int sxj = sxe + j*cn;
                    if ((unsigned)sxj >= (unsigned)swidth) { while (sxj < 0) sxj += cn; while (sxj >= swidth) sxj -= cn; }
                    v += S[sxj] * alpha[dx*ks + j];
                }
                rows[k*dwidth + o] = v;
            }
        }
        double* D = (double*)(dst + (size_t)dy * dst_step);
        for (int o = 0; o < dwidth; o++) {
            double t = rows[o] * beta[0];
            for (int k = 1; k < ks; k++) t = t + rows[k*dwidth + o] * beta[k];
            D[o] = t;
        }
    }
    free(rows); free(alpha); free(xofs);
    return ORC_OK;
}

/* INTER_NEAREST_EXACT: resizeNN_bitexact resize.cpp:1267-1288, invoker :1174-1265 (16.16 fixed point, pixel centres) */
static void resize_nn_exact(int esz, const uchar* src, size_t src_step, int sw, int sh,
                            uchar* dst, size_t dst_step, int dw, int dh)
{
    int ifx = (int)((((unsigned)sw << 16) + dw / 2) / dw), ifx0 = ifx / 2 - sw % 2;
    int ify = (int)((((unsigned)sh << 16) + dh / 2) / dh), ify0 = ify / 2 - sh % 2;
    for (int y = 0; y < dh; y++) {
        int sy = (int)((unsigned)ify * (unsigned)y + (unsigned)ify0) >> 16; if (sy > sh - 1) sy = sh - 1;
        const uchar* S = src + (size_t)sy * src_step;
        uchar* D = dst + (size_t)y * dst_step;
        for (int x = 0; x < dw; x++) {
            int sx = (int)((unsigned)ifx * (unsigned)x + (unsigned)ifx0) >> 16; if (sx > sw - 1) sx = sw - 1;
            memcpy(D + (size_t)x*esz, S + (size_t)sx*esz, esz);
        }
    }
}

/* INTER_LINEAR_EXACT for 8U (ufixedpoint16, 8.8) and 16U (ufixedpoint32, 16.16): resize_bitExact resize.cpp:895-946,
   interpolationLinear::getCoeffs :789-828, hlineResize :74-109, vlineResize/vlineSet :708-786, invoker :842-886,
   fixed-point types fixedpoint.inl.hpp:224-374.  softdouble arithmetic == IEEE double arithmetic, op by op. */
typedef struct { int ofs; unsigned c0, c1; int where; } ExactCoef;   /* where: -1 left of min, 0 inside, +1 at/after max */
static ExactCoef exact_coef(int val, double scale, int maxsize, int shift)
{
    ExactCoef e; e.ofs = 0; e.c0 = e.c1 = 0;
    double fval = scale * ((double)val + 0.5) - 0.5;
    int ival = floor_d(fval);
    if (ival >= 0 && maxsize > 1) {
        if (ival < maxsize - 1) {
            e.where = 0; e.ofs = ival;
            double frac = fval - (double)ival;
            e.c1 = frac < 0 ? 0u : (unsigned)rne_d(frac * (double)(1 << shift));
            unsigned one = 1u << shift;
            e.c0 = one > e.c1 ? one - e.c1 : 0u;
        } else { e.where = 1; e.ofs = maxsize - 1; }
    } else e.where = -1;
    return e;
}

/* INTER_LINEAR_EXACT on signed data: CV_8S and CV_16S both run in fixedpoint32 (16.16, int32_t; fixedtype<int8_t / int16_t, .>,
 * resize.cpp:66-71), vertical products in fixedpoint64 (fixedpoint.inl.hpp:169-221).  Coefficients as for the unsigned types; sample *
 * coefficient and the sums saturate to the fixed-point type (operator*, operator+), the cast back adds half an LSB in the UNSIGNED
 * type of the same width and shifts arithmetically (the saturate_cast member). */
static long long sat_ll(long long v, long long lo, long long hi) { return v < lo ? lo : v > hi ? hi : v; }
static int resize_linear_exact_signed(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                                      uchar* dst, size_t dst_step, int dw, int dh, double inv_scale_x, double inv_scale_y)
{
    const int shift = 16;
    const long long ft_lo = -2147483648LL, ft_hi = 2147483647LL;
    const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
    ExactCoef* cx = (ExactCoef*)malloc(sizeof(ExactCoef) * (size_t)dw);
    for (int x = 0; x < dw; x++) cx[x] = exact_coef(x, scale_x, sw, shift);
    for (int y = 0; y < dh; y++) {
        ExactCoef cy = exact_coef(y, scale_y, sh, shift);
        uchar* D = dst + (size_t)y * dst_step;
        const uchar* R0 = src + (size_t)(cy.where < 0 ? 0 : cy.where > 0 ? sh - 1 : cy.ofs) * src_step;
        const uchar* R1 = cy.where == 0 ? R0 + src_step : R0;
        for (int x = 0; x < dw; x++) for (int c = 0; c < cn; c++) {
            long long hh[2];
            for (int k = 0; k < 2; k++) {
                const uchar* R = k ? R1 : R0;
                #define LDS_(i) ((long long)(depth == D_8S ? ((const signed char*)R)[i] : ((const short*)R)[i]))
                long long h;
                if (cx[x].where < 0) h = LDS_(c) * (1LL << shift);
                else if (cx[x].where > 0) h = LDS_((sw - 1)*cn + c) * (1LL << shift);
                else {
                    long long a = sat_ll((long long)cx[x].c0 * LDS_(cx[x].ofs*cn + c), ft_lo, ft_hi);
                    long long b = sat_ll((long long)cx[x].c1 * LDS_((cx[x].ofs + 1)*cn + c), ft_lo, ft_hi);
                    h = sat_ll(a + b, ft_lo, ft_hi);
                }
                #undef LDS_
                hh[k] = h;
            }
            long long v;
            if (cy.where != 0) v = (int)((unsigned)hh[0] + 32768u) >> 16;          /* vlineSet: (ET)FT */
            else {                                                                  /* vlineResize: (ET)(m0*s0 + m1*s1), products in WT */
                __int128 acc = (__int128)hh[0] * cy.c0 + (__int128)hh[1] * cy.c1;
                const __int128 lo = -((__int128)1 << 63), hi = ((__int128)1 << 63) - 1;
                if (acc < lo) acc = lo;
                if (acc > hi) acc = hi;
                v = (long long)((unsigned long long)(long long)acc + 0x80000000ull) >> 32;
            }
            if (depth == D_8S) ((signed char*)D)[x*cn + c] = (signed char)sat_ll(v, -128, 127);
            else ((short*)D)[x*cn + c] = (short)sat_ll(v, -32768, 32767);
        }
    }
    free(cx);
    return ORC_OK;
}

/* INTER_LINEAR_EXACT on CV_32S: fixedpoint64 (32.32, int64_t) for the horizontal AND the vertical pass (fixedpoint64::WT is
 * fixedpoint64 itself, fixedpoint.inl.hpp:14-99).  Its product works on magnitudes: (|a| * |b| + 2^31) >> 32 as a 128-bit quantity,
 * saturated at 2^63 - 1 (negative: -2^63), the sign applied afterwards -- i.e. it rounds half AWAY from zero; sums saturate; the cast
 * back is (int64)((uint64)v + 2^31) >> 32 saturated to int.  Coefficients: cvRound64(frac * 2^32), one = 2^32. */
static long long fp64_mul(long long a, long long b)
{
    const int neg = (a < 0) != (b < 0);
    const unsigned long long ua = a < 0 ? 0ull - (unsigned long long)a : (unsigned long long)a;
    const unsigned long long ub = b < 0 ? 0ull - (unsigned long long)b : (unsigned long long)b;
    const unsigned __int128 full = (unsigned __int128)ua * ub;
    const unsigned __int128 r = (full + ((unsigned __int128)1 << 31)) >> 32;
    /* the reference tests sh2 > 0x7FFFFFFF || val0_h > 0x7FFFFFFF, i.e. r >= 2^63 */
    if (r >> 63) return neg ? (long long)0x8000000000000000ull : 0x7FFFFFFFFFFFFFFFll;
    return neg ? -(long long)(unsigned long long)r : (long long)(unsigned long long)r;
}
static long long fp64_add(long long a, long long b)
{
    const long long res = (long long)((unsigned long long)a + (unsigned long long)b);
    return (((a ^ res) & (b ^ res)) < 0) ? (long long)~((unsigned long long)res & 0x8000000000000000ull) : res;
}
typedef struct { int ofs; long long c0, c1; int where; } ExactCoef64;
static ExactCoef64 exact_coef64(int val, double scale, int maxsize)
{
    ExactCoef64 e; e.ofs = 0; e.c0 = e.c1 = 0;
    double fval = scale * ((double)val + 0.5) - 0.5;
    int ival = floor_d(fval);
    if (ival >= 0 && maxsize > 1) {
        if (ival < maxsize - 1) {
            e.where = 0; e.ofs = ival;
            e.c1 = llrint((fval - (double)ival) * 4294967296.0);               /* cvRound64 */
            /* one() - c1 with fixedpoint64's saturating operator- : no overflow possible here */
            e.c0 = 4294967296LL - e.c1;
        } else { e.where = 1; e.ofs = maxsize - 1; }
    } else e.where = -1;
    return e;
}
static int resize_linear_exact_s32(int cn, const uchar* src, size_t src_step, int sw, int sh,
                                   uchar* dst, size_t dst_step, int dw, int dh, double inv_scale_x, double inv_scale_y)
{
    const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
    ExactCoef64* cx = (ExactCoef64*)malloc(sizeof(ExactCoef64) * (size_t)dw);
    for (int x = 0; x < dw; x++) cx[x] = exact_coef64(x, scale_x, sw);
    for (int y = 0; y < dh; y++) {
        ExactCoef64 cy = exact_coef64(y, scale_y, sh);
        int* D = (int*)(dst + (size_t)y * dst_step);
        const int* R0 = (const int*)(src + (size_t)(cy.where < 0 ? 0 : cy.where > 0 ? sh - 1 : cy.ofs) * src_step);
        const int* R1 = cy.where == 0 ? (const int*)((const uchar*)R0 + src_step) : R0;
        for (int x = 0; x < dw; x++) for (int c = 0; c < cn; c++) {
            long long hh[2];
            for (int k = 0; k < 2; k++) {
                const int* R = k ? R1 : R0;
                if (cx[x].where < 0) hh[k] = (long long)((unsigned long long)(long long)R[c] << 32);
                else if (cx[x].where > 0) hh[k] = (long long)((unsigned long long)(long long)R[(sw - 1)*cn + c] << 32);
                else hh[k] = fp64_add(fp64_mul(cx[x].c0, (long long)((unsigned long long)(long long)R[cx[x].ofs*cn + c] << 32)),
                                      fp64_mul(cx[x].c1, (long long)((unsigned long long)(long long)R[(cx[x].ofs + 1)*cn + c] << 32)));
            }
            long long acc = cy.where != 0 ? hh[0] : fp64_add(fp64_mul(hh[0], cy.c0), fp64_mul(hh[1], cy.c1));
            long long v = (long long)((unsigned long long)acc + 0x80000000ull) >> 32;
            D[x*cn + c] = (int)sat_ll(v, -2147483648LL, 2147483647LL);
        }
    }
    free(cx);
    return ORC_OK;
}

static int resize_linear_exact(int depth, int cn, const uchar* src, size_t src_step, int sw, int sh,
                               uchar* dst, size_t dst_step, int dw, int dh, double inv_scale_x, double inv_scale_y)
{
    if (depth == D_32S)
        return resize_linear_exact_s32(cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
    if (depth == D_8S || depth == D_16S)
        return resize_linear_exact_signed(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
    if (depth != D_8U && depth != D_16U) return ORC_NOT_IMPLEMENTED;
    const int shift = depth == D_8U ? 8 : 16;
    const double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
    ExactCoef* cx = (ExactCoef*)malloc(sizeof(ExactCoef) * (size_t)dw);
    for (int x = 0; x < dw; x++) cx[x] = exact_coef(x, scale_x, sw, shift);
    /* the reference's min/max are monotone thresholds (:823-826); a column is "left" iff its own ival < 0 etc. */
    for (int y = 0; y < dh; y++) {
        ExactCoef cy = exact_coef(y, scale_y, sh, shift);
        uchar* D = dst + (size_t)y * dst_step;
        const uchar* R0 = src + (size_t)(cy.where < 0 ? 0 : cy.where > 0 ? sh - 1 : cy.ofs) * src_step;
        const uchar* R1 = cy.where == 0 ? R0 + src_step : R0;
        for (int x = 0; x < dw; x++) for (int c = 0; c < cn; c++) {
            unsigned long long h0, h1;      /* horizontally interpolated values of the two rows, FT raw */
            for (int k = 0; k < 2; k++) {
                const uchar* R = k ? R1 : R0;
                unsigned long long h;
                if (cx[x].where < 0) h = (unsigned long long)(depth == D_8U ? R[c] : ((const ushort*)R)[c]) << shift;
                else if (cx[x].where > 0) h = (unsigned long long)(depth == D_8U ? R[(sw - 1)*cn + c] : ((const ushort*)R)[(sw - 1)*cn + c]) << shift;
                else {
                    unsigned p0 = depth == D_8U ? R[cx[x].ofs*cn + c] : ((const ushort*)R)[cx[x].ofs*cn + c];
                    unsigned p1 = depth == D_8U ? R[(cx[x].ofs + 1)*cn + c] : ((const ushort*)R)[(cx[x].ofs + 1)*cn + c];
                    unsigned long long a = (unsigned long long)cx[x].c0 * p0, b = (unsigned long long)cx[x].c1 * p1;
                    unsigned long long cap = depth == D_8U ? 0xFFFFull : 0xFFFFFFFFull;
                    if (a > cap) a = cap;                             /* saturate_cast in operator* */
                    if (b > cap) b = cap;
                    h = a + b; if (h > cap) h = cap;                 /* saturating operator+ */
                }
                if (k) h1 = h; else h0 = h;
            }
            unsigned long long v;
            if (cy.where != 0) {                                     /* vlineSet: FT -> ET with rounding */
                v = (h0 + (1ull << (shift - 1))) >> shift;           /* no FT wrap possible: h0 <= max<<shift */
            } else {                                                 /* vlineResize: WT = FT*FT, 2*shift fractional bits */
                unsigned long long acc = h0 * cy.c0 + h1 * cy.c1;
                v = (acc + (1ull << (2*shift - 1))) >> (2*shift);
            }
            unsigned maxv = depth == D_8U ? 255u : 65535u;
            if (v > maxv) v = maxv;
            if (depth == D_8U) D[x*cn + c] = (uchar)v; else ((ushort*)D)[x*cn + c] = (ushort)v;
        }
    }
    free(cx);
    return ORC_OK;
}

int orc_resize(int src_type, const uchar* src, size_t src_step, int sw, int sh,
               uchar* dst, size_t dst_step, int dw, int dh,
               double inv_scale_x, double inv_scale_y, int interpolation)
{
    int depth = MAT_DEPTH(src_type), cn = MAT_CN(src_type);
    if (inv_scale_x < DBL_EPSILON || inv_scale_y < DBL_EPSILON) {    /* :3642-3646 */
        inv_scale_x = (double)dw / sw; inv_scale_y = (double)dh / sh;
    }
    /* the reference re-derives dsize (:3652-3653) and writes that many pixels */
    if (rne_d(sw*inv_scale_x) != dw || rne_d(sh*inv_scale_y) != dh) return ORC_NOT_IMPLEMENTED;
    if (interpolation == I_NEAREST) {
        int esz = (depth <= D_8S ? 1 : depth <= D_16S ? 2 : depth <= D_32F ? 4 : 8) * cn;
        resize_nn(esz, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
        return ORC_OK;
    }
    if (interpolation == I_NEAREST_EXACT) {                          /* :3805-3809 */
        int esz = (depth <= D_8S ? 1 : depth <= D_16S ? 2 : depth <= D_32F ? 4 : 8) * cn;
        if (sw >= 32768 || sh >= 32768) return ORC_NOT_IMPLEMENTED;
        resize_nn_exact(esz, src, src_step, sw, sh, dst, dst_step, dw, dh);
        return ORC_OK;
    }
    double scale_x = 1./inv_scale_x, scale_y = 1./inv_scale_y;
    int iscale_x = rne_d(scale_x), iscale_y = rne_d(scale_y);       /* :3769-3770 */
    int is_area_fast = fabs(scale_x - iscale_x) < DBL_EPSILON && fabs(scale_y - iscale_y) < DBL_EPSILON;
    if (interpolation == I_LINEAR_EXACT) {                           /* :3782-3798 */
        if (is_area_fast && iscale_x == 2 && iscale_y == 2 && cn != 2) interpolation = I_AREA;
        else return resize_linear_exact(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
    }
    if (depth == D_64F) {                                            /* linear, cubic and Lanczos4 are restated for 64F (no area) */
        if (interpolation == I_CUBIC || interpolation == I_LANCZOS4)
            return resize_hq_f64(interpolation == I_CUBIC ? 4 : 8, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
        if (interpolation == I_LINEAR && is_area_fast && iscale_x == 2 && iscale_y == 2) interpolation = I_AREA;      /* :3817-3818 */
        if (interpolation == I_AREA && scale_x >= 1 && scale_y >= 1) {
            if (is_area_fast) { resize_area_fast_f64(cn, src, src_step, sw, sh, dst, dst_step, dw, dh, iscale_x, iscale_y); return ORC_OK; }
            if (cn > 4) return ORC_ERR;
            return resize_area_f64(cn, src, src_step, sw, sh, dst, dst_step, dw, dh, scale_x, scale_y);
        }
        if (interpolation != I_LINEAR && interpolation != I_AREA) return ORC_NOT_IMPLEMENTED;
        return resize_linear(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y, interpolation == I_AREA);
    }
    if (!(depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F)) return ORC_NOT_IMPLEMENTED;
    if (interpolation == I_LANCZOS4) return resize_lanczos4(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
    if (interpolation == I_CUBIC) return resize_cubic(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y);
    if (interpolation != I_LINEAR && interpolation != I_AREA) return ORC_NOT_IMPLEMENTED;
    if (interpolation == I_LINEAR && is_area_fast && iscale_x == 2 && iscale_y == 2)   /* :3817-3818 */
        interpolation = I_AREA;
    if (interpolation == I_AREA && scale_x >= 1 && scale_y >= 1) {   /* :3822 */
        if (is_area_fast) {
            resize_area_fast(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, iscale_x, iscale_y);
            return ORC_OK;
        }
        if (cn > 4) return ORC_ERR;                                  /* CV_Assert :3851 */
        return resize_area(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh, scale_x, scale_y);
    }
    return resize_linear(depth, cn, src, src_step, sw, sh, dst, dst_step, dw, dh,
                         inv_scale_x, inv_scale_y, interpolation == I_AREA);
}

/* ------------------------------------------------------------------------------------------
 * cv::dnn::blobFromImages, CV_8U frames -> fp32 NCHW blob: dnn/src/dnn_utils.cpp:275-353 (resize INTER_LINEAR when the
 * sizes differ), :130-241 (planar copy with swapRB, then (p - mean[ch]) * scale[ch] per blob channel).
 * ------------------------------------------------------------------------------------------ */
int orc_blobFromImages(int src_type, const uchar* src, size_t sfs, size_t src_step, int sw, int sh,
                       float* blob, int bw, int bh, int n_frames, const double scalefactor[4], const double mean[4], int swapRB)
{
    const int depth = MAT_DEPTH(src_type), cn = MAT_CN(src_type);
    if (depth != D_8U || !(cn == 1 || cn == 3 || cn == 4)) return ORC_NOT_IMPLEMENTED;
    const size_t wh = (size_t)bw * bh;
    uchar* tmp = (uchar*)malloc(wh * cn);
    for (int f = 0; f < n_frames; f++) {
        const uchar* img = src + (size_t)f * sfs; size_t step = src_step;
        if (bw != sw || bh != sh) {
            int rc = orc_resize(src_type, img, src_step, sw, sh, tmp, (size_t)bw * cn, bw, bh, 0.0, 0.0, I_LINEAR);
            if (rc != ORC_OK) { free(tmp); return rc; }
            img = tmp; step = (size_t)bw * cn;
        }
        float* pb = blob + (size_t)f * cn * wh;
        for (int k = 0; k < cn; k++) {
            const int ch = (swapRB && cn >= 3 && k < 3) ? 2 - k : k;
            const float cur_mean = (float)(mean ? mean[ch] : 0.0), cur_scale = (float)(scalefactor ? scalefactor[ch] : 1.0);
            for (int y = 0; y < bh; y++)
                for (int x = 0; x < bw; x++) {
                    float v = (float)img[y * step + (size_t)x * cn + k];
                    pb[ch * wh + (size_t)y * bw + x] = (v - cur_mean) * cur_scale;
                }
        }
    }
    free(tmp);
    return ORC_OK;
}

/* blobFromImagesWithParams with a padding mode (dnn_utils.cpp:303-337): 0 = DNN_PMODE_NULL (resize to the blob size, as above),
 * 1 = DNN_PMODE_CROP_CENTER (resize by max of the two float ratios through (fx, fy), then the centred crop, :309-317),
 * 2 = DNN_PMODE_LETTERBOX (resize by the min ratio to (int(w*f), int(h*f)), constant border of borderValue around it, :318-331). */
int orc_blobFromImagesWithParams(int src_type, const uchar* src, size_t sfs, size_t src_step, int sw, int sh,
                                 float* blob, int bw, int bh, int n_frames, const double scalefactor[4], const double mean[4], int swapRB,
                                 int paddingmode, const double borderValue[4])
{
    const int depth = MAT_DEPTH(src_type), cn = MAT_CN(src_type);
    if (depth != D_8U || !(cn == 1 || cn == 3 || cn == 4)) return ORC_NOT_IMPLEMENTED;
    if (paddingmode == 0 || (bw == sw && bh == sh))
        return orc_blobFromImages(src_type, src, sfs, src_step, sw, sh, blob, bw, bh, n_frames, scalefactor, mean, swapRB);
    const size_t wh = (size_t)bw * bh;
    int rw, rh, ox = 0, oy = 0; double inv = 0.0;
    if (paddingmode == 1) {
        float a = bw / (float)sw, b = bh / (float)sh, rf = a > b ? a : b;
        inv = rf;
        rw = rne_d(sw * inv); rh = rne_d(sh * inv);                   /* resize.cpp:4015-4016 saturate_cast<int>(ssize * inv_scale) */
        ox = (int)(0.5 * (rw - bw)); oy = (int)(0.5 * (rh - bh));     /* Point(double, double): truncation */
        if (rw <= 0 || rh <= 0 || ox < 0 || oy < 0 || ox + bw > rw || oy + bh > rh) return ORC_NOT_IMPLEMENTED;   /* Mat::operator()(Rect) asserts */
    } else if (paddingmode == 2) {
        float a = bw / (float)sw, b = bh / (float)sh, rf = a < b ? a : b;
        rh = (int)(sh * rf); rw = (int)(sw * rf);
        if (rw <= 0 || rh <= 0) return ORC_NOT_IMPLEMENTED;
        oy = -((bh - rh) / 2); ox = -((bw - rw) / 2);
    } else return ORC_NOT_IMPLEMENTED;
    uchar* tmp = (uchar*)malloc((size_t)rw * rh * cn);
    for (int f = 0; f < n_frames; f++) {
        const uchar* img = src + (size_t)f * sfs; size_t step = src_step;
        if (rw != sw || rh != sh) {
            int rc = orc_resize(src_type, img, src_step, sw, sh, tmp, (size_t)rw * cn, rw, rh, inv, inv, I_LINEAR);
            if (rc != ORC_OK) { free(tmp); return rc; }
            img = tmp; step = (size_t)rw * cn;
        }
        float* pb = blob + (size_t)f * cn * wh;
        for (int k = 0; k < cn; k++) {
            const int ch = (swapRB && cn >= 3 && k < 3) ? 2 - k : k;
            const float cur_mean = (float)(mean ? mean[ch] : 0.0), cur_scale = (float)(scalefactor ? scalefactor[ch] : 1.0);
            int pad = rne_d(borderValue ? borderValue[k] : 0.0); pad = pad < 0 ? 0 : pad > 255 ? 255 : pad;   /* saturate_cast<uchar>(double) */
            for (int y = 0; y < bh; y++)
                for (int x = 0; x < bw; x++) {
                    const int rx = x + ox, ry = y + oy;
                    const int inside = rx >= 0 && rx < rw && ry >= 0 && ry < rh;
                    float v = (float)(inside ? img[ry * step + (size_t)rx * cn + k] : pad);
                    pb[ch * wh + (size_t)y * bw + x] = (v - cur_mean) * cur_scale;
                }
        }
    }
    free(tmp);
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------
 * Host-side helpers of the public API, above the hooks (used to pin the host mirror).
 * ------------------------------------------------------------------------------------------ */
/* cv::warpAffine's closed-form inversion, imgwarp.cpp:2401-2411 (this exact operation order) */
void orc_invert_affine(const double Min[6], double M[6])
{
    memcpy(M, Min, 6*sizeof(double));
    double D = M[0]*M[4] - M[1]*M[3];
    D = D != 0 ? 1./D : 0;
    double A11 = M[4]*D, A22 = M[0]*D;
    M[0] = A11; M[1] *= -D;
    M[3] *= -D; M[4] = A22;
    double b1 = -M[0]*M[2] - M[1]*M[5];
    double b2 = -M[3]*M[2] - M[4]*M[5];
    M[2] = b1; M[5] = b2;
}

/* source-footprint bitmap for the roofline's algorithmic bytes: counts distinct in-bounds
   source pixels touched by any tap of a LINEAR warpAffine.  Not reference code: a measurement helper. */
long long orc_affine_footprint(int sw, int sh, int dw, int dh, const double M[6])
{
    uchar* bm = (uchar*)calloc((size_t)sw * sh, 1);
    const int AB_BITS = 10, AB_SCALE = 1 << AB_BITS, rd = AB_SCALE/INTER_TAB_SIZE/2;
    for (int y = 0; y < dh; y++) {
        int X0 = rne_d((M[1]*y + M[2])*AB_SCALE) + rd, Y0 = rne_d((M[4]*y + M[5])*AB_SCALE) + rd;
        for (int x = 0; x < dw; x++) {
            int X = (X0 + rne_d(M[0]*x*AB_SCALE)) >> 5, Y = (Y0 + rne_d(M[3]*x*AB_SCALE)) >> 5;
            int sx = sat_s16(X >> 5), sy = sat_s16(Y >> 5);
            for (int j = 0; j < 2; j++) for (int i = 0; i < 2; i++) {
                int xx = sx + i, yy = sy + j;
                if ((unsigned)xx < (unsigned)sw && (unsigned)yy < (unsigned)sh) bm[(size_t)yy*sw + xx] = 1;
            }
        }
    }
    long long n = 0;
    for (size_t i = 0; i < (size_t)sw*sh; i++) n += bm[i];
    free(bm);
    return n;
}
