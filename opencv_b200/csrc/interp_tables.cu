// INTER_CUBIC / INTER_LANCZOS4 coefficient tables (reference: imgwarp.cpp:89-206).
//
// The reference builds them once on the host: 1-D coefficients for the 32 sub-pixel positions (fp32 polynomial for
// bicubic, fp64 sin/cos for Lanczos4), their 2-D float products, and for 8U an int16 copy of the products scaled by
// 2^15 whose sum is forced to 2^15 by adjusting one of the four centre taps.  Lanczos4 depends on the host libm's
// sin/cos, so the tables are built on the host here as well (same libm as a reference running on the same box) and
// copied once per device into fixed __device__ arrays -- no allocation.  The kernels read the 1-D table (staged in
// shared memory; w2d = vy*vx is the float table's own definition) and the int16 2-D table.
#include "interp_tables.cuh"
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <mutex>

namespace b200 {
namespace {

constexpr int TAB = 32;
__device__ float g_cubic_1d[TAB * 4];
__device__ float g_lanczos_1d[TAB * 8];
__device__ short g_cubic_2d_i[TAB * TAB * 16];
__device__ short g_lanczos_2d_i[TAB * TAB * 64];

struct HostTables {
    float cubic1[TAB * 4], lanczos1[TAB * 8];
    short cubic2[TAB * TAB * 16], lanczos2[TAB * TAB * 64];
};

void cubic_coeffs(float x, float* c)                     // interpolateCubic, imgwarp.cpp:89-97 (all fp32)
{
    const float A = -0.75f;
    const float x1 = x + 1, xr = 1 - x;
    c[0] = ((A * x1 - 5 * A) * x1 + 8 * A) * x1 - 4 * A;
    c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
    c[2] = ((A + 2) * xr - (A + 3)) * xr * xr + 1;
    c[3] = 1.f - c[0] - c[1] - c[2];
}

void lanczos4_coeffs(float x, float* c)                  // interpolateLanczos4, imgwarp.cpp:99-127
{
    const double s45 = 0.70710678118654752440084436210485, pi = 3.1415926535897932384626433832795;
    const double cs[8][2] = {{1, 0}, {-s45, -s45}, {0, 1}, {s45, -s45}, {-1, 0}, {s45, s45}, {0, -1}, {-s45, s45}};
    if (x < FLT_EPSILON) {
        for (int i = 0; i < 8; i++) c[i] = 0;
        c[3] = 1;
        return;
    }
    float sum = 0;
    const float x3 = x + 3;                              // float add first, as in the reference's -(x+3)*CV_PI*0.25
    const double y0 = -x3 * pi * 0.25, s0 = std::sin(y0), c0 = std::cos(y0);
    for (int i = 0; i < 8; i++) {
        const float xi = x3 - i;
        const double y = -xi * pi * 0.25;
        c[i] = (float)((cs[i][0] * s0 + cs[i][1] * c0) / (y * y));
        sum += c[i];
    }
    sum = 1.f / sum;
    for (int i = 0; i < 8; i++) c[i] *= sum;
}

short sat_short_rne(float v)                             // saturate_cast<short>(float): cvRound then clamp
{
    const float r = std::nearbyintf(v);
    const int i = (r >= -2147483648.0f && r < 2147483648.0f) ? (int)r : INT_MIN;
    return (short)(i < -32768 ? -32768 : i > 32767 ? 32767 : i);
}

void build_2d(int ks, const float* t1, short* itab)      // initInterTab2D, imgwarp.cpp:154-206
{
    for (int i = 0; i < TAB; i++)
        for (int j = 0; j < TAB; j++, itab += ks * ks) {
            int isum = 0;
            for (int k1 = 0; k1 < ks; k1++)
                for (int k2 = 0; k2 < ks; k2++) {
                    const float v = t1[i * ks + k1] * t1[j * ks + k2];
                    isum += itab[k1 * ks + k2] = sat_short_rne(v * 32768);
                }
            if (isum == 32768) continue;
            // the reference moves the rounding excess onto the largest (deficit) or smallest (surplus) of the four centre taps
            const int diff = isum - 32768, h = ks / 2;
            int M1 = h, M2 = h, m1 = h, m2 = h;
            for (int k1 = h; k1 < h + 2; k1++)
                for (int k2 = h; k2 < h + 2; k2++) {
                    if (itab[k1 * ks + k2] < itab[m1 * ks + m2]) m1 = k1, m2 = k2;
                    else if (itab[k1 * ks + k2] > itab[M1 * ks + M2]) M1 = k1, M2 = k2;
                }
            if (diff < 0) itab[M1 * ks + M2] = (short)(itab[M1 * ks + M2] - diff);
            else itab[m1 * ks + m2] = (short)(itab[m1 * ks + m2] - diff);
        }
}

const HostTables& host_tables()
{
    static HostTables t;
    static std::once_flag once;
    std::call_once(once, [] {
        const float scale = 1.f / TAB;
        for (int i = 0; i < TAB; i++) { cubic_coeffs(i * scale, t.cubic1 + i * 4); lanczos4_coeffs(i * scale, t.lanczos1 + i * 8); }
        build_2d(4, t.cubic1, t.cubic2);
        build_2d(8, t.lanczos1, t.lanczos2);
    });
    return t;
}

} // namespace

cudaError_t interp_tables(int interp, const float** tab1d, const short** tab2d_i)
{
    static std::mutex mu;
    static unsigned long long uploaded = 0;              // bit per device (64, the limit noalloc_pipeline.cu uses too)
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev >= 64) return cudaErrorInvalidDevice;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!(uploaded & (1ull << dev))) {
            const HostTables& t = host_tables();
            if ((e = cudaMemcpyToSymbol(g_cubic_1d, t.cubic1, sizeof(t.cubic1))) != cudaSuccess) return e;
            if ((e = cudaMemcpyToSymbol(g_lanczos_1d, t.lanczos1, sizeof(t.lanczos1))) != cudaSuccess) return e;
            if ((e = cudaMemcpyToSymbol(g_cubic_2d_i, t.cubic2, sizeof(t.cubic2))) != cudaSuccess) return e;
            if ((e = cudaMemcpyToSymbol(g_lanczos_2d_i, t.lanczos2, sizeof(t.lanczos2))) != cudaSuccess) return e;
            // The copies above come from pageable memory on the legacy default stream and may return before the DMA has
            // landed; the kernels that read the tables run on non-blocking streams, which do not order against it.
            if ((e = cudaDeviceSynchronize()) != cudaSuccess) return e;
            uploaded |= 1ull << dev;
        }
    }
    void *p1 = nullptr, *p2 = nullptr;
    if (interp == 2) {
        if ((e = cudaGetSymbolAddress(&p1, g_cubic_1d)) != cudaSuccess) return e;
        if ((e = cudaGetSymbolAddress(&p2, g_cubic_2d_i)) != cudaSuccess) return e;
    } else {
        if ((e = cudaGetSymbolAddress(&p1, g_lanczos_1d)) != cudaSuccess) return e;
        if ((e = cudaGetSymbolAddress(&p2, g_lanczos_2d_i)) != cudaSuccess) return e;
    }
    *tab1d = (const float*)p1; *tab2d_i = (const short*)p2;
    return cudaSuccess;
}

void lanczos4_coeffs_host(float x, float* c) { lanczos4_coeffs(x, c); }

void interp_tables_host(int interp, const float** tab1d, const short** tab2d_i)
{
    const HostTables& t = host_tables();
    *tab1d = interp == 2 ? t.cubic1 : t.lanczos1;
    *tab2d_i = interp == 2 ? t.cubic2 : t.lanczos2;
}

} // namespace b200
