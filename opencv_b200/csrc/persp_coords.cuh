// warpPerspective coordinate derivation shared by the generic gather kernel (warp_kernels.cu) and the TMA-staged
// kernel (fast_kernels.cu).  Reference: imgwarp.cpp:2689-2731, :2758-2824; imgwarp.sse4_1.cpp:313-458.
#pragma once
#include "warp_kernels.cuh"

namespace b200 {

// Per tile row: X0/Y0/W0 at the origin xb of the reference's coordinate block (imgwarp.cpp:2712-2714), evaluated left to
// right in fp64.  A 32-wide tile lies inside one block whenever bw0 is a multiple of 32 (bw0 = 64 for every image of at
// least 64x16), so 8 threads do this once per tile instead of every pixel doing it.
struct PerspRow { double X0, Y0, W0; };

template <class Args>
__device__ __forceinline__ PerspRow persp_row_terms(const Args& a, int xb, int y)
{
    const double dxb = (double)xb, dy = (double)y;
    PerspRow r;
    r.X0 = __dadd_rn(__dadd_rn(__dmul_rn(a.M[0], dxb), __dmul_rn(a.M[1], dy)), a.M[2]);
    r.Y0 = __dadd_rn(__dadd_rn(__dmul_rn(a.M[3], dxb), __dmul_rn(a.M[4], dy)), a.M[5]);
    r.W0 = __dadd_rn(__dadd_rn(__dmul_rn(a.M[6], dxb), __dmul_rn(a.M[7], dy)), a.M[8]);
    return r;
}

// exact path: IEEE division, the reference's clamps and its NaN split (imgwarp.cpp:2803-2821, imgwarp.sse4_1.cpp:313-458)
template <bool LINEAR>
__device__ __noinline__ int2 persp_exact(double Nx, double Ny, double W, bool vec)      // by value: no stack traffic on the fast path
{
    int X, Y;
    W = (W != 0.0) ? __ddiv_rn(LINEAR ? 32.0 : 1.0, W) : 0.0;
    double fX = __dmul_rn(Nx, W), fY = __dmul_rn(Ny, W);
    // clamp to [INT_MIN, INT_MAX]; NaN: vector body (first bw&~15 px of a block) keeps NaN -> INT_MIN,
    // scalar tail maps NaN -> INT_MAX (imgwarp.sse4_1.cpp:326-330 vs :451-452)
    if (fX != fX) fX = vec ? fX : 2147483647.0; else fX = fmax(-2147483648.0, fmin(2147483647.0, fX));
    if (fY != fY) fY = vec ? fY : 2147483647.0; else fY = fmax(-2147483648.0, fmin(2147483647.0, fY));
    X = rne_d(fX); Y = rne_d(fY);
    return make_int2(X, Y);
}

// The reference computes X = cvRound((X0 + M0*x1) * (32/W)) with every step rounded to double.  The numerators and W are
// evaluated exactly as the reference does; the quotient is first approximated (MUFU.RCP seed + two fp64 Newton steps,
// relative error < 2^-49) and the product rounded to the nearest integer.  If that product is farther than its error bound
// from every half-integer the rounded result provably equals the reference's; otherwise (|frac| within 2^-24 of .5,
// huge or non-finite values: ~1 pixel in 10^7) the pixel takes the exact IEEE-division path.  Bit-exactness is kept, the
// ~20-instruction fp64 division leaves the common path (first capture: 86 % issue-slot utilisation, fp64-bound).
// returns the reference's rounded fixed-point coordinate pair (X, Y) (LINEAR: in 1/32 pixel units) for pixel x1 of a block
template <bool LINEAR, class Args>
__device__ __forceinline__ int2 persp_xy(const Args& a, const PerspRow& r, int x1, int bw)
{
    const double dx1 = (double)x1;
    const double W = __dadd_rn(r.W0, __dmul_rn(a.M[6], dx1));
    const double Nx = __dadd_rn(r.X0, __dmul_rn(a.M[0], dx1));
    const double Ny = __dadd_rn(r.Y0, __dmul_rn(a.M[3], dx1));
    int X, Y;
    bool slow = true;
    const double aw = fabs(W);
    if (aw > 1e-30 && aw < 1e30) {
        float rf;
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rf) : "f"(__double2float_rn(W)));
        double q = (double)rf;
        q = __fma_rn(q, __fma_rn(-W, q, 1.0), q);
        q = __fma_rn(q, __fma_rn(-W, q, 1.0), q);
        if (LINEAR) q = __dmul_rn(q, 32.0);                              // exact (power of two)
        const double pX = __dmul_rn(Nx, q), pY = __dmul_rn(Ny, q);
        X = __double2int_rn(pX); Y = __double2int_rn(pY);
        const double dX = fabs(__dsub_rn(pX, (double)X)), dY = fabs(__dsub_rn(pY, (double)Y));   // exact differences
        // |p| < 2^30 here, so the approximation error is < 2^30 * 2^-49 = 2^-19 ... keep a wide 2^-16 guard band
        const double guard = 0.5 - 1.52587890625e-05;
        slow = !(fabs(pX) < 1073741824.0 && fabs(pY) < 1073741824.0 && dX < guard && dY < guard);
    }
    if (slow) {
        const int2 e = LINEAR ? persp_exact<true>(Nx, Ny, W, x1 < (bw & ~15)) : persp_exact<false>(Nx, Ny, W, x1 < (bw & ~15));
        X = e.x; Y = e.y;
    }
    return make_int2(X, Y);
}

template <bool LINEAR, class Args>
__device__ __forceinline__ void persp_coord(const Args& a, const PerspRow& r, int x1, int bw,
                                            int& sx, int& sy, int& fx, int& fy)
{
    const int2 p = persp_xy<LINEAR>(a, r, x1, bw);
    const int X = p.x, Y = p.y;
    if (LINEAR) { sx = sat_s16(X >> INTER_BITS); sy = sat_s16(Y >> INTER_BITS); fx = X & 31; fy = Y & 31; }
    else { sx = sat_s16(X); sy = sat_s16(Y); fx = fy = 0; }
}

} // namespace b200
