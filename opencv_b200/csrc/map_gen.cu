// Remap map generators on the device: cv::fisheye::initUndistortRectifyMap (calib3d/src/fisheye.cpp:530-631), the producer of
// BASELINE config 4's fisheye-undistort map.  The map is written straight into device memory in either format cv::remap takes
// (CV_16SC2 + CV_16UC1 fixed point, or two CV_32FC1 planes), so a frame loop never builds it on the host and uploads it.
//
// The reference walks a row with three fp64 accumulators advanced by REPEATED addition (_x += iR00, fisheye.cpp:627-629), which
// is not _x0 + j*iR00 once rounding is counted.  The kernel reproduces that: one CTA per row; thread 0 runs the recurrence for
// the next 256 columns into shared memory (three independent add chains, ~8 cycles a step), then 256 threads evaluate their
// pixel (two divisions, sqrt, atan, the theta polynomial in the reference's order, each operation rounded separately) while
// thread 0 already advances.  atan is the CUDA math library's (<= 1 ulp), the reference uses the host libm's: a map entry can
// differ in its last float bit / last 1/32 pixel where the fp64 value sits within an ulp of a rounding boundary (measured:
// see tests/test_gpu_mapgen.py), everything else is bit-identical.
#include "../../include/b200_geom.h"
#include "geom_common.cuh"
#include "trace.cuh"
#include <cmath>
#include <vector>

namespace b200 { void count_launch(); int report_cuda_error(const char* what, cudaError_t e); }   // capi.cu

namespace b200 {
namespace {

struct FisheyeArgs {
    double iR[9], f0, f1, c0, c1, k[4];
    uint8_t* map1; int step1; uint8_t* map2; int step2;
    int w, h, fixed;           // fixed: CV_16SC2 + CV_16UC1, else 2 x CV_32FC1
};

constexpr int MG_T = 256;

__global__ void __launch_bounds__(MG_T) fisheye_map_kernel(const FisheyeArgs a)
{
    __shared__ double sx[2][MG_T], sy[2][MG_T], sw[2][MG_T];
    const int i = blockIdx.x, tid = threadIdx.x;
    // row start: i*iR(0,1) + iR(0,2) etc. (:584-586), each product and sum rounded
    double _x = __dadd_rn(__dmul_rn((double)i, a.iR[1]), a.iR[2]);
    double _y = __dadd_rn(__dmul_rn((double)i, a.iR[4]), a.iR[5]);
    double _w = __dadd_rn(__dmul_rn((double)i, a.iR[7]), a.iR[8]);
    uint8_t* r1 = a.map1 + (long long)i * a.step1;
    uint8_t* r2 = a.map2 + (long long)i * a.step2;
    const int chunks = (a.w + MG_T - 1) / MG_T;
    auto advance = [&](int buf, int n) {          // thread 0: the next n columns of the recurrence
        for (int j = 0; j < n; j++) {
            sx[buf][j] = _x; sy[buf][j] = _y; sw[buf][j] = _w;
            _x = __dadd_rn(_x, a.iR[0]); _y = __dadd_rn(_y, a.iR[3]); _w = __dadd_rn(_w, a.iR[6]);
        }
    };
    if (tid == 0) advance(0, min(MG_T, a.w));
    __syncthreads();
    for (int c = 0; c < chunks; c++) {
        const int buf = c & 1, j = c * MG_T + tid;
        if (tid == 0 && c + 1 < chunks) advance(buf ^ 1, min(MG_T, a.w - (c + 1) * MG_T));
        if (j < a.w) {
            const double X = sx[buf][tid], Y = sy[buf][tid], Wv = sw[buf][tid];
            double u, v;
            if (Wv <= 0) {
                u = (X > 0) ? -__longlong_as_double(0x7ff0000000000000LL) : __longlong_as_double(0x7ff0000000000000LL);
                v = (Y > 0) ? -__longlong_as_double(0x7ff0000000000000LL) : __longlong_as_double(0x7ff0000000000000LL);
            } else {
                const double x = __ddiv_rn(X, Wv), y = __ddiv_rn(Y, Wv);
                const double r = __dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)));
                const double theta = atan(r);
                const double t2 = __dmul_rn(theta, theta), t4 = __dmul_rn(t2, t2), t6 = __dmul_rn(t4, t2), t8 = __dmul_rn(t4, t4);
                // theta * (1 + k0*t2 + k1*t4 + k2*t6 + k3*t8), left to right (:603)
                const double poly = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(1.0, __dmul_rn(a.k[0], t2)), __dmul_rn(a.k[1], t4)),
                                                        __dmul_rn(a.k[2], t6)), __dmul_rn(a.k[3], t8));
                const double theta_d = __dmul_rn(theta, poly);
                const double scale = (r == 0) ? 1.0 : __ddiv_rn(theta_d, r);
                u = __dadd_rn(__dmul_rn(__dmul_rn(a.f0, x), scale), a.c0);
                v = __dadd_rn(__dmul_rn(__dmul_rn(a.f1, y), scale), a.c1);
            }
            if (a.fixed) {                                                   // :611-617
                const int iu = rne_d(__dmul_rn(u, 32.0)), iv = rne_d(__dmul_rn(v, 32.0));
                reinterpret_cast<short2*>(r1)[j] = make_short2((short)(iu >> 5), (short)(iv >> 5));
                reinterpret_cast<unsigned short*>(r2)[j] = (unsigned short)((iv & 31) * 32 + (iu & 31));
            } else {                                                         // :618-622
                reinterpret_cast<float*>(r1)[j] = __double2float_rn(u);
                reinterpret_cast<float*>(r2)[j] = __double2float_rn(v);
            }
        }
        __syncthreads();
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------------
// cv::initUndistortRectifyMap (calib3d/src/undistort.dispatch.cpp:82-161, undistort.simd.hpp:102-277) as a stock x86 build runs
// it: the AVX2 dispatch variant.  Its vector body takes a row in chunks of 8 pixels with fused multiply-adds (v_muladd = FMA3),
// advances the row accumulators by 8*ir per chunk and adds the table ir*i per lane; the scalar tail (last width % 8 pixels)
// rounds every operation, advances by ir per pixel and packs CV_16SC2 with a wrapping cast where the body saturates.  All of it
// is IEEE arithmetic (no transcendental function), so the kernel is bit-identical: __fma_rn where the reference fuses,
// __dmul_rn / __dadd_rn / __ddiv_rn where it does not.  One CTA per row; thread 0 runs the chunk recurrence for the next 256
// chunks into shared memory while the other threads evaluate theirs (8 pixels each).
struct UndistArgs {
    double ir[9], tilt[9], fx, fy, u0, v0, k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4;
    uint8_t* map1; int step1; uint8_t* map2; int step2;
    int w, h, m1type;
    double Ar[9]; int stripe_rows;      // cv::undistort: stripes of stripe_rows rows, each with the inverse of its own Ar (0: one map, ir as given)
};

namespace {

// cv::Mat::inv(DECOMP_LU) of a 3x3 CV_64F matrix is cv::invert's closed form (core/src/lapack.cpp:711-713 det3, :946-968), every
// operation rounded
__device__ void inv3_rounded(const double* m, double* t)
{
    const double c0 = __dsub_rn(__dmul_rn(m[4], m[8]), __dmul_rn(m[5], m[7]));
    const double c1 = __dsub_rn(__dmul_rn(m[3], m[8]), __dmul_rn(m[5], m[6]));
    const double c2 = __dsub_rn(__dmul_rn(m[3], m[7]), __dmul_rn(m[4], m[6]));
    double d = __dadd_rn(__dsub_rn(__dmul_rn(m[0], c0), __dmul_rn(m[1], c1)), __dmul_rn(m[2], c2));
    if (d == 0.0) { for (int i = 0; i < 9; i++) t[i] = 0.0; return; }
    d = __ddiv_rn(1.0, d);
    t[0] = __dmul_rn(c0, d);
    t[1] = __dmul_rn(__dsub_rn(__dmul_rn(m[2], m[7]), __dmul_rn(m[1], m[8])), d);
    t[2] = __dmul_rn(__dsub_rn(__dmul_rn(m[1], m[5]), __dmul_rn(m[2], m[4])), d);
    t[3] = __dmul_rn(__dsub_rn(__dmul_rn(m[5], m[6]), __dmul_rn(m[3], m[8])), d);
    t[4] = __dmul_rn(__dsub_rn(__dmul_rn(m[0], m[8]), __dmul_rn(m[2], m[6])), d);
    t[5] = __dmul_rn(__dsub_rn(__dmul_rn(m[2], m[3]), __dmul_rn(m[0], m[5])), d);
    t[6] = __dmul_rn(c2, d);
    t[7] = __dmul_rn(__dsub_rn(__dmul_rn(m[1], m[6]), __dmul_rn(m[0], m[7])), d);
    t[8] = __dmul_rn(__dsub_rn(__dmul_rn(m[0], m[4]), __dmul_rn(m[1], m[3])), d);
}

__device__ __forceinline__ void undist_store(const UndistArgs& a, uint8_t* r1, uint8_t* r2, bool body, int j, double u, double v)
{
    if (a.m1type == 11) {
        const int iu = rne_d(__dmul_rn(u, 32.0)), iv = rne_d(__dmul_rn(v, 32.0));
        const int xs = iu >> 5, ys = iv >> 5;
        reinterpret_cast<short2*>(r1)[j] = body ? make_short2((short)sat_s16(xs), (short)sat_s16(ys)) : make_short2((short)xs, (short)ys);
        reinterpret_cast<unsigned short*>(r2)[j] = (unsigned short)((iv & 31) * 32 + (iu & 31));
    } else if (a.m1type == 5) {
        reinterpret_cast<float*>(r1)[j] = __double2float_rn(u);
        reinterpret_cast<float*>(r2)[j] = __double2float_rn(v);
    } else {
        reinterpret_cast<float2*>(r1)[j] = make_float2(__double2float_rn(u), __double2float_rn(v));
    }
}

__global__ void __launch_bounds__(MG_T) undistort_map_kernel(const __grid_constant__ UndistArgs a)
{
    __shared__ double sx[2][MG_T], sy[2][MG_T], sw[2][MG_T], tail0[3], sir[9];
    const int tid = threadIdx.x;
    int i = blockIdx.x;                                        // row inside its map (stripe)
    uint8_t* r1 = a.map1 + (long long)blockIdx.x * a.step1;
    uint8_t* r2 = a.map2 ? a.map2 + (long long)blockIdx.x * a.step2 : nullptr;
    if (a.stripe_rows > 0) {                                   // undistort.dispatch.cpp:311-325: Ar(1,2) = v0 - y, a new inverse per stripe
        const int y0 = (i / a.stripe_rows) * a.stripe_rows;
        i -= y0;
        if (tid == 0) {
            double m[9];
            for (int k = 0; k < 9; k++) m[k] = a.Ar[k];
            m[5] = __dsub_rn(m[5], (double)y0);
            for (int k = 0; k < 9; k++) m[k] = __dadd_rn(m[k], 0.0);      // Ar * I by cv::gemm: x*1 + 0 + 0
            inv3_rounded(m, sir);
        }
    } else if (tid < 9) sir[tid] = a.ir[tid];
    __syncthreads();
    double ir[9];
#pragma unroll
    for (int k = 0; k < 9; k++) ir[k] = sir[k];
    const int nb = a.w / 8, nseg = (nb + MG_T - 1) / MG_T;
    double _x = __dadd_rn(__dmul_rn((double)i, ir[1]), ir[2]);                 // :115
    double _y = __dadd_rn(__dmul_rn((double)i, ir[4]), ir[5]);
    double _w = __dadd_rn(__dmul_rn((double)i, ir[7]), ir[8]);
    const double dx8 = __dmul_rn(8.0, ir[0]), dy8 = __dmul_rn(8.0, ir[3]), dw8 = __dmul_rn(8.0, ir[6]);
    auto advance = [&](int buf, int n) {          // thread 0: accumulator values at the start of the next n chunks (:128)
        for (int c = 0; c < n; c++) {
            sx[buf][c] = _x; sy[buf][c] = _y; sw[buf][c] = _w;
            _x = __dadd_rn(_x, dx8); _y = __dadd_rn(_y, dy8); _w = __dadd_rn(_w, dw8);
        }
    };
    if (tid == 0) {
        if (nseg > 0) advance(0, min(MG_T, nb));
        if (nseg <= 1) { tail0[0] = _x; tail0[1] = _y; tail0[2] = _w; }
    }
    __syncthreads();
    for (int seg = 0; seg < nseg; seg++) {
        const int buf = seg & 1, c = seg * MG_T + tid;
        if (tid == 0 && seg + 1 < nseg) {
            advance(buf ^ 1, min(MG_T, nb - (seg + 1) * MG_T));
            if (seg + 2 >= nseg) { tail0[0] = _x; tail0[1] = _y; tail0[2] = _w; }
        }
        if (c < nb) {
            const double X0 = sx[buf][tid], Y0 = sy[buf][tid], W0 = sw[buf][tid];
#pragma unroll 2
            for (int l = 0; l < 8; l++) {
                const double dl = (double)l;
                const double m = __ddiv_rn(1.0, __dadd_rn(W0, __dmul_rn(ir[6], dl)));                 // :131-132 (s_w[l] = ir[6]*l)
                double x = __dmul_rn(__dadd_rn(X0, __dmul_rn(ir[0], dl)), m);
                double y = __dmul_rn(__dadd_rn(Y0, __dmul_rn(ir[3], dl)), m);
                double xd = __dmul_rn(x, x), yd = __dmul_rn(y, y);
                const double r2v = __dadd_rn(xd, yd);
                const double kr = __ddiv_rn(__fma_rn(__fma_rn(__fma_rn(a.k3, r2v, a.k2), r2v, a.k1), r2v, 1.0),
                                            __fma_rn(__fma_rn(__fma_rn(a.k6, r2v, a.k5), r2v, a.k4), r2v, 1.0));   // :150-155
                xd = __fma_rn(2.0, xd, r2v); yd = __fma_rn(2.0, yd, r2v);                                 // :158-161
                const double xy2 = __dmul_rn(__dmul_rn(x, y), 2.0);                                        // :162
                x = __dmul_rn(x, kr); y = __dmul_rn(y, kr);
                xd = __fma_rn(xd, a.p2, x); yd = __fma_rn(yd, a.p1, y);                                   // :169-172
                xd = __fma_rn(a.p1, xy2, xd); yd = __fma_rn(a.p2, xy2, yd);                               // :174-177
                const double r4 = __dmul_rn(r2v, r2v);
                xd = __fma_rn(a.s1, r2v, __fma_rn(a.s2, r4, xd));                                         // :183-184
                yd = __fma_rn(a.s3, r2v, __fma_rn(a.s4, r4, yd));
                const double X = __fma_rn(a.tilt[0], xd, __fma_rn(a.tilt[1], yd, a.tilt[2]));
                const double Y = __fma_rn(a.tilt[3], xd, __fma_rn(a.tilt[4], yd, a.tilt[5]));
                const double Z = __fma_rn(a.tilt[6], xd, __fma_rn(a.tilt[7], yd, a.tilt[8]));
                const double inv = (Z == 0.0) ? 1.0 : __ddiv_rn(1.0, Z);                                   // :207-209
                const double u = __fma_rn(__dmul_rn(a.fx, inv), X, a.u0), v = __fma_rn(__dmul_rn(a.fy, inv), Y, a.v0);
                undist_store(a, r1, r2, true, c * 8 + l, u, v);
            }
        }
        __syncthreads();
    }
    // scalar tail: pixel nb*8 + tid, accumulators advanced by tid repeated additions (:242)
    const int j = nb * 8 + tid;
    if (j < a.w) {
        double tx = tail0[0], ty = tail0[1], tw = tail0[2];
        for (int t = 0; t < tid; t++) { tx = __dadd_rn(tx, ir[0]); ty = __dadd_rn(ty, ir[3]); tw = __dadd_rn(tw, ir[6]); }
        const double ww = __ddiv_rn(1.0, tw), x = __dmul_rn(tx, ww), y = __dmul_rn(ty, ww);
        const double x2 = __dmul_rn(x, x), y2 = __dmul_rn(y, y);
        const double r2v = __dadd_rn(x2, y2), xy2 = __dmul_rn(__dmul_rn(2.0, x), y);
        const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.k3, r2v), a.k2), r2v), a.k1), r2v));
        const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(a.k6, r2v), a.k5), r2v), a.k4), r2v));
        const double kr = __ddiv_rn(num, den);
        // x*kr + p1*_2xy + p2*(r2 + 2*x2) + s1*r2 + s2*r2*r2, left to right (:247)
        double xd = __dadd_rn(__dmul_rn(x, kr), __dmul_rn(a.p1, xy2));
        xd = __dadd_rn(xd, __dmul_rn(a.p2, __dadd_rn(r2v, __dmul_rn(2.0, x2))));
        xd = __dadd_rn(xd, __dmul_rn(a.s1, r2v));
        xd = __dadd_rn(xd, __dmul_rn(__dmul_rn(a.s2, r2v), r2v));
        double yd = __dadd_rn(__dmul_rn(y, kr), __dmul_rn(a.p1, __dadd_rn(r2v, __dmul_rn(2.0, y2))));
        yd = __dadd_rn(yd, __dmul_rn(a.p2, xy2));
        yd = __dadd_rn(yd, __dmul_rn(a.s3, r2v));
        yd = __dadd_rn(yd, __dmul_rn(__dmul_rn(a.s4, r2v), r2v));
        // Matx33d * Vec3d(xd, yd, 1): s = 0 + a0*xd, += a1*yd, += a2*1
        const double X = __dadd_rn(__dadd_rn(__dmul_rn(a.tilt[0], xd), __dmul_rn(a.tilt[1], yd)), a.tilt[2]);
        const double Y = __dadd_rn(__dadd_rn(__dmul_rn(a.tilt[3], xd), __dmul_rn(a.tilt[4], yd)), a.tilt[5]);
        const double Z = __dadd_rn(__dadd_rn(__dmul_rn(a.tilt[6], xd), __dmul_rn(a.tilt[7], yd)), a.tilt[8]);
        const double inv = (Z != 0.0) ? __ddiv_rn(1.0, Z) : 1.0;
        const double u = __dadd_rn(__dmul_rn(__dmul_rn(a.fx, inv), X), a.u0), v = __dadd_rn(__dmul_rn(__dmul_rn(a.fy, inv), Y), a.v0);
        undist_store(a, r1, r2, false, j, u, v);
    }
}

} // namespace
} // namespace b200

extern "C" __attribute__((visibility("default")))
int b200_fisheye_initUndistortRectifyMap_dev(const double K[9], const double D[4], const double iR[9], int width, int height,
                                             int m1type, void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream)
{
    B200_TRACE();
    using namespace b200;
    if (m1type <= 0) m1type = 11;                                            // CV_16SC2 (:536)
    if ((m1type != 11 && m1type != 5) || !K || !iR || !map1 || !map2 || width <= 0 || height <= 0) return B200_NOT_IMPLEMENTED;
    if (map1_step > 0x7fffffffu || map2_step > 0x7fffffffu) return B200_NOT_IMPLEMENTED;
    const int a1 = 4, a2 = m1type == 11 ? 2 : 4;
    if (((uintptr_t)map1 % a1) || (map1_step % a1) || ((uintptr_t)map2 % a2) || (map2_step % a2)) return B200_NOT_IMPLEMENTED;
    FisheyeArgs a;
    for (int i = 0; i < 9; i++) a.iR[i] = iR[i];
    a.f0 = K[0]; a.f1 = K[4]; a.c0 = K[2]; a.c1 = K[5];
    for (int i = 0; i < 4; i++) a.k[i] = D ? D[i] : 0.0;
    a.map1 = (uint8_t*)map1; a.step1 = (int)map1_step; a.map2 = (uint8_t*)map2; a.step2 = (int)map2_step;
    a.w = width; a.h = height; a.fixed = m1type == 11;
    fisheye_map_kernel<<<height, MG_T, 0, (cudaStream_t)stream>>>(a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return report_cuda_error("fisheye map kernel launch", e);
    count_launch();
    return B200_OK;
}

// cv::fisheye::undistortImage (calib3d/src/fisheye.cpp:636-646): initUndistortRectifyMap(K, D, I, Knew, size, CV_16SC2) followed by
// cv::remap(INTER_LINEAR, BORDER_CONSTANT), both on the device and on the same stream.  The caller lends the map storage
// (map1: 4 bytes per destination pixel, map2: 2 bytes), so a frame loop can keep the maps and call the remap entry alone.
extern "C" __attribute__((visibility("default")))
int b200_fisheye_undistortImage_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                                    void* dst, size_t dst_step, int dst_width, int dst_height,
                                    const double K[9], const double D[4], const double iKnew[9],
                                    void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream)
{
    B200_TRACE();
    int rc = b200_fisheye_initUndistortRectifyMap_dev(K, D, iKnew, dst_width, dst_height, 11, map1, map1_step, map2, map2_step, stream);
    if (rc != B200_OK) return rc;
    const double bv[4] = {0, 0, 0, 0};
    return b200_remap16s_dev(src_type, src, src_step, src_width, src_height, dst, dst_step, dst_width, dst_height,
                             (const short*)map1, map1_step, (const unsigned short*)map2, map2_step, 1 /* INTER_LINEAR */, 0 /* BORDER_CONSTANT */,
                             bv, stream);
}

namespace b200 {
int launch_undistort_map(const double* K, const double* dist12, const double* iR, const double* Ar, int stripe_rows, const double* tilt,
                         int width, int height, int m1type, void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream)
{
    if (m1type <= 0) m1type = 11;                                            // undistort.dispatch.cpp:90-91
    if ((m1type != 11 && m1type != 5 && m1type != 13) || !K || (!iR && !Ar) || !map1 || width <= 0 || height <= 0) return B200_NOT_IMPLEMENTED;
    if (m1type != 13 && !map2) return B200_NOT_IMPLEMENTED;
    if (map1_step > 0x7fffffffu || map2_step > 0x7fffffffu) return B200_NOT_IMPLEMENTED;
    const int a1 = m1type == 13 ? 8 : 4, a2 = m1type == 11 ? 2 : 4;
    if (((uintptr_t)map1 % a1) || (map1_step % a1)) return B200_NOT_IMPLEMENTED;
    if (m1type != 13 && (((uintptr_t)map2 % a2) || (map2_step % a2))) return B200_NOT_IMPLEMENTED;
    UndistArgs a;
    for (int i = 0; i < 9; i++) {
        a.ir[i] = iR ? iR[i] : 0.0; a.Ar[i] = Ar ? Ar[i] : 0.0;
        a.tilt[i] = tilt ? tilt[i] : (i % 4 == 0 ? 1.0 : 0.0);
    }
    a.stripe_rows = Ar ? stripe_rows : 0;
    a.fx = K[0]; a.fy = K[4]; a.u0 = K[2]; a.v0 = K[5];
    const double z12[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const double* d = dist12 ? dist12 : z12;
    a.k1 = d[0]; a.k2 = d[1]; a.p1 = d[2]; a.p2 = d[3]; a.k3 = d[4]; a.k4 = d[5]; a.k5 = d[6]; a.k6 = d[7];
    a.s1 = d[8]; a.s2 = d[9]; a.s3 = d[10]; a.s4 = d[11];
    a.map1 = (uint8_t*)map1; a.step1 = (int)map1_step; a.map2 = m1type == 13 ? nullptr : (uint8_t*)map2; a.step2 = (int)map2_step;
    a.w = width; a.h = height; a.m1type = m1type;
    undistort_map_kernel<<<height, MG_T, 0, (cudaStream_t)stream>>>(a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return report_cuda_error("undistort map kernel launch", e);
    count_launch();
    return B200_OK;
}
}

extern "C" __attribute__((visibility("default")))
int b200_initUndistortRectifyMap_dev(const double K[9], const double dist12[12], const double iR[9], const double tilt[9],
                                     int width, int height, int m1type, void* map1, size_t map1_step, void* map2, size_t map2_step,
                                     void* stream)
{
    B200_TRACE();
    if (!iR) return B200_NOT_IMPLEMENTED;
    return b200::launch_undistort_map(K, dist12, iR, nullptr, 0, tilt, width, height, m1type, map1, map1_step, map2, map2_step, stream);
}

// cv::undistort (undistort.dispatch.cpp:284-328).  The reference takes the image in stripes of max(1, 4096 / cols) rows and gives each
// stripe a map of its own, built from a new camera matrix whose v0 is lowered by the stripe's first row: every stripe inverts its own
// matrix and counts its rows from 0, so the entries are NOT those of one map over the whole image.  One launch builds all stripes
// (each CTA inverts its stripe's matrix), then cv::remap INTER_LINEAR BORDER_CONSTANT.  newK = NULL: the camera matrix.
extern "C" __attribute__((visibility("default")))
int b200_undistort_dev(int src_type, const void* src, size_t src_step, int width, int height, void* dst, size_t dst_step,
                       const double K[9], const double dist12[12], const double newK[9], const double tilt[9],
                       void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream)
{
    B200_TRACE();
    if (!K || width <= 0 || height <= 0) return B200_NOT_IMPLEMENTED;
    int stripe = 4096 / width; if (stripe < 1) stripe = 1; if (stripe > height) stripe = height;           // :297
    int rc = b200::launch_undistort_map(K, dist12, nullptr, newK ? newK : K, stripe, tilt, width, height, 11, map1, map1_step, map2, map2_step, stream);
    if (rc != B200_OK) return rc;
    const double bv[4] = {0, 0, 0, 0};
    return b200_remap16s_dev(src_type, src, src_step, width, height, dst, dst_step, width, height,
                             (const short*)map1, map1_step, (const unsigned short*)map2, map2_step, 1, 0, bv, stream);
}

// The two matrices cv::initUndistortRectifyMap derives before its pixel loop (undistort.dispatch.cpp:123, :155-157), computed the way
// the reference does (cv::gemm's 3x3 form, cv::invert's closed form, Matx products, host libm sin / cos): host code, no device work.
namespace {
__attribute__((optimize("fp-contract=off"))) void matx33_mul(const double* a, const double* b, double* c)
{
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            volatile double s = 0;
            for (int k = 0; k < 3; k++) { volatile double p = a[i * 3 + k] * b[k * 3 + j]; s = s + p; }
            c[i * 3 + j] = s;
        }
}
}
extern "C" __attribute__((visibility("default"))) __attribute__((optimize("fp-contract=off")))
int b200_undistort_params(const double newK[9], const double R[9], double tauX, double tauY, double iR_out[9], double tilt_out[9])
{
    B200_TRACE();
    static const double I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (!newK || !iR_out) return B200_NOT_IMPLEMENTED;
    const double* b = R ? R : I3;
    volatile double m[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            volatile double p0 = newK[i * 3] * b[j], p1 = newK[i * 3 + 1] * b[3 + j], p2 = newK[i * 3 + 2] * b[6 + j];
            volatile double t = p0 + p1; t = t + p2;
            m[i * 3 + j] = t + 0.0;
        }
    volatile double a0 = m[4] * m[8], a1 = m[5] * m[7], b0 = m[3] * m[8], b1 = m[5] * m[6], e0 = m[3] * m[7], e1 = m[4] * m[6];
    volatile double c0 = a0 - a1, c1 = b0 - b1, c2 = e0 - e1;
    volatile double q0 = m[0] * c0, q1 = m[1] * c1, q2 = m[2] * c2;
    volatile double d = q0 - q1; d = d + q2;
    if (d == 0.0) { for (int i = 0; i < 9; i++) iR_out[i] = 0.0; }
    else {
        d = 1.0 / d;
        auto cof = [&](int p, int q, int r, int s2) { volatile double x = m[p] * m[q], y = m[r] * m[s2]; volatile double z = x - y; return (double)(z * d); };
        iR_out[0] = c0 * d;            iR_out[1] = cof(2, 7, 1, 8); iR_out[2] = cof(1, 5, 2, 4);
        iR_out[3] = cof(5, 6, 3, 8);   iR_out[4] = cof(0, 8, 2, 6); iR_out[5] = cof(2, 3, 0, 5);
        iR_out[6] = c2 * d;            iR_out[7] = cof(1, 6, 0, 7); iR_out[8] = cof(0, 4, 1, 3);
    }
    if (tilt_out) {
        for (int i = 0; i < 9; i++) tilt_out[i] = I3[i];
        if (tauX != 0 || tauY != 0) {                                                // distortion_model.hpp:75-94
            const double cx = cos(tauX), sx = sin(tauX), cy = cos(tauY), sy = sin(tauY);
            const double rx[9] = {1, 0, 0, 0, cx, sx, 0, -sx, cx}, ry[9] = {cy, 0, -sy, 0, 1, 0, sy, 0, cy};
            double rxy[9];
            matx33_mul(ry, rx, rxy);
            const double pz[9] = {rxy[8], 0, -rxy[2], 0, rxy[8], -rxy[5], 0, 0, 1};
            matx33_mul(pz, rxy, tilt_out);
        }
    }
    return B200_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// The map loops of the stitching module's rotation warpers (stitching/include/opencv2/stitching/detail/warpers_inl.hpp:74-99,
// stitching/src/warpers.cpp:192-243): P::mapBackward for every pixel of the result ROI, plane (:235-248), spherical (:263-281) and
// cylindrical (:296-314) projectors, float arithmetic in the reference's order with every operation rounded.  The plane projector is
// pure IEEE arithmetic -> bit-identical; the other two take sinf / cosf from the CUDA math library where the reference takes the
// host libm's (both within 2 ulp of the true value, not of each other), so their maps agree to a few ulp.
namespace b200 {
struct StitchArgs { float k[9], t[3], scale; int kind, tl_x, tl_y, w, h; uint8_t* xmap; int xstep; uint8_t* ymap; int ystep; };
namespace {
__global__ void __launch_bounds__(256) stitch_map_kernel(const StitchArgs a)
{
    const int i = blockIdx.x * 256 + threadIdx.x, j = blockIdx.y;
    if (i >= a.w) return;
    float u = (float)(a.tl_x + i), v = (float)(a.tl_y + j), x, y, z;
    const float* k = a.k;
    auto dot3 = [](float c0, float p, float c1, float q, float c2, float r) {       // c0*p + c1*q + c2*r, left to right
        return __fadd_rn(__fadd_rn(__fmul_rn(c0, p), __fmul_rn(c1, q)), __fmul_rn(c2, r));
    };
    if (a.kind == 0) {
        u = __fsub_rn(__fdiv_rn(u, a.scale), a.t[0]); v = __fsub_rn(__fdiv_rn(v, a.scale), a.t[1]);
        const float w1 = __fsub_rn(1.f, a.t[2]);
        x = dot3(k[0], u, k[1], v, k[2], w1); y = dot3(k[3], u, k[4], v, k[5], w1); z = dot3(k[6], u, k[7], v, k[8], w1);
        x = __fdiv_rn(x, z); y = __fdiv_rn(y, z);
    } else {
        const float pi = 3.14159274101257324f;                              // static_cast<float>(CV_PI)
        u = __fdiv_rn(u, a.scale); v = __fdiv_rn(v, a.scale);
        float x_, y_, z_;
        if (a.kind == 1) { const float sinv = sinf(__fsub_rn(pi, v)); x_ = __fmul_rn(sinv, sinf(u)); y_ = cosf(__fsub_rn(pi, v)); z_ = __fmul_rn(sinv, cosf(u)); }
        else { x_ = sinf(u); y_ = v; z_ = cosf(u); }
        x = dot3(k[0], x_, k[1], y_, k[2], z_); y = dot3(k[3], x_, k[4], y_, k[5], z_); z = dot3(k[6], x_, k[7], y_, k[8], z_);
        if (z > 0) { x = __fdiv_rn(x, z); y = __fdiv_rn(y, z); } else x = y = -1.f;
    }
    reinterpret_cast<float*>(a.xmap + (long long)j * a.xstep)[i] = x;
    reinterpret_cast<float*>(a.ymap + (long long)j * a.ystep)[i] = y;
}
}
}

extern "C" __attribute__((visibility("default")))
int b200_stitch_buildMaps_dev(int kind, const float k_rinv[9], const float t[3], float scale, int tl_x, int tl_y, int width, int height,
                              float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, void* stream)
{
    B200_TRACE();
    using namespace b200;
    if (kind < 0 || kind > 2 || !k_rinv || !xmap || !ymap || width <= 0 || height <= 0 || height > 65535) return B200_NOT_IMPLEMENTED;
    if (xmap_step > 0x7fffffffu || ymap_step > 0x7fffffffu || ((uintptr_t)xmap & 3) || ((uintptr_t)ymap & 3) || (xmap_step & 3) || (ymap_step & 3))
        return B200_NOT_IMPLEMENTED;
    StitchArgs a;
    for (int i = 0; i < 9; i++) a.k[i] = k_rinv[i];
    for (int i = 0; i < 3; i++) a.t[i] = t ? t[i] : 0.f;
    a.scale = scale; a.kind = kind; a.tl_x = tl_x; a.tl_y = tl_y; a.w = width; a.h = height;
    a.xmap = (uint8_t*)xmap; a.xstep = (int)xmap_step; a.ymap = (uint8_t*)ymap; a.ystep = (int)ymap_step;
    stitch_map_kernel<<<dim3((width + 255) / 256, height), 256, 0, (cudaStream_t)stream>>>(a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return report_cuda_error("stitching map kernel launch", e);
    count_launch();
    return B200_OK;
}

// RotationWarperBase<P>::warp / PlaneWarper::warp (warpers_inl.hpp:102-113, warpers.cpp:246-254): the maps above + cv::remap on one stream
extern "C" __attribute__((visibility("default")))
int b200_stitch_warp_dev(int kind, const float k_rinv[9], const float t[3], float scale, int tl_x, int tl_y,
                         int src_type, const void* src, size_t src_step, int src_width, int src_height,
                         void* dst, size_t dst_step, int dst_width, int dst_height,
                         float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, int interpolation, int border, void* stream)
{
    B200_TRACE();
    int rc = b200_stitch_buildMaps_dev(kind, k_rinv, t, scale, tl_x, tl_y, dst_width, dst_height, xmap, xmap_step, ymap, ymap_step, stream);
    if (rc != B200_OK) return rc;
    const double bv[4] = {0, 0, 0, 0};
    return b200_remap32f_dev(src_type, src, src_step, src_width, src_height, dst, dst_step, dst_width, dst_height,
                             xmap, xmap_step, ymap, ymap_step, interpolation, border, bv, stream);
}

// ------------------------------------------------------------------------------------------------------------------------------
// cv::warpPolar, forward direction (reference imgproc/src/imgwarp.cpp:3163-3224): CV_32FC1 maps of the polar image (column = radius,
// row = angle), then cv::remap.  The reference takes its transcendental values per column (the rho table, :3190-3204) and per row
// (cos / sin, :3208-3210) only, so they are computed on the host with the same libm, as the reference does, and the device does the
// one multiply-add per entry (:3216-3220, rounded separately in double, then to float): the maps are bit-identical by construction.
namespace b200 {
struct PolarArgs { const float* rho; const double* cs; double cx, cy; int w, h; uint8_t* xmap; int xstep; uint8_t* ymap; int ystep; };
namespace {
__global__ void __launch_bounds__(256) polar_map_kernel(const PolarArgs a)
{
    const int i = blockIdx.x * 256 + threadIdx.x, j = blockIdx.y;
    if (i >= a.w) return;
    const double r = (double)__ldg(a.rho + i), cp = __ldg(a.cs + 2 * j), sp = __ldg(a.cs + 2 * j + 1);
    reinterpret_cast<float*>(a.xmap + (long long)j * a.xstep)[i] = __double2float_rn(__dadd_rn(__dmul_rn(r, cp), a.cx));
    reinterpret_cast<float*>(a.ymap + (long long)j * a.ystep)[i] = __double2float_rn(__dadd_rn(__dmul_rn(r, sp), a.cy));
}

// the tables of the last geometry, per host thread (a video loop unwraps every frame the same way); uploads are ordered on the caller's stream
struct PolarCache {
    int device = -1, w = 0, h = 0, semilog = 0; double max_radius = 0;
    void* p = nullptr; size_t cap = 0;
};
thread_local PolarCache g_polar;

int polar_tables(int w, int h, double max_radius, int semilog, cudaStream_t s, const float** rho, const double** cs)
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return B200_ERROR;
    PolarCache& c = g_polar;
    const size_t cs_off = ((size_t)w * 4 + 15) & ~(size_t)15, bytes = cs_off + (size_t)h * 16;
    if (!(c.device == dev && c.w == w && c.h == h && c.semilog == semilog && c.max_radius == max_radius && c.p)) {
        if (c.device != dev) { c.p = nullptr; c.cap = 0; }             // another device's allocation is not ours to reuse here
        if (bytes > c.cap) {
            if (c.p) cudaFree(c.p);
            c.p = nullptr; c.cap = 0;
            const cudaError_t e = cudaMalloc(&c.p, bytes);
            if (e != cudaSuccess) return report_cuda_error("warpPolar tables", e);
            c.cap = bytes;
        }
        std::vector<unsigned char> host(bytes);
        float* hr = reinterpret_cast<float*>(host.data()); double* hc = reinterpret_cast<double*>(host.data() + cs_off);
        if (semilog) { const double Kmag = std::log(max_radius) / w; for (int r = 0; r < w; r++) hr[r] = (float)(std::exp(r * Kmag) - 1.0); }
        else { const double Kmag = max_radius / w; for (int r = 0; r < w; r++) hr[r] = (float)(r * Kmag); }
        const double Kangle = 6.283185307179586476925286766559 / h;    // CV_2PI
        for (int phi = 0; phi < h; phi++) { const double KKy = Kangle * phi; hc[2 * phi] = std::cos(KKy); hc[2 * phi + 1] = std::sin(KKy); }
        // pageable source: the runtime stages it before returning, so `host` may go out of scope
        const cudaError_t e = cudaMemcpyAsync(c.p, host.data(), bytes, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) { c.device = -1; return report_cuda_error("warpPolar tables upload", e); }
        c.device = dev; c.w = w; c.h = h; c.semilog = semilog; c.max_radius = max_radius;
    }
    *rho = reinterpret_cast<const float*>(c.p);
    *cs = reinterpret_cast<const double*>(reinterpret_cast<const unsigned char*>(c.p) + cs_off);
    return B200_OK;
}
}
}

extern "C" __attribute__((visibility("default")))
int b200_warpPolar_buildMaps_dev(int width, int height, float center_x, float center_y, double max_radius, int flags,
                                 float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, void* stream)
{
    B200_TRACE();
    using namespace b200;
    if (flags & 16) return B200_NOT_IMPLEMENTED;                       // WARP_INVERSE_MAP: left to the reference (cartToPolar + cv::log on the host)
    if (!xmap || !ymap || width <= 0 || height <= 0 || height > 65535) return B200_NOT_IMPLEMENTED;
    if (xmap_step > 0x7fffffffu || ymap_step > 0x7fffffffu || ((uintptr_t)xmap & 3) || ((uintptr_t)ymap & 3) || (xmap_step & 3) || (ymap_step & 3))
        return B200_NOT_IMPLEMENTED;
    PolarArgs a;
    int rc = polar_tables(width, height, max_radius, (flags & 256) != 0, (cudaStream_t)stream, &a.rho, &a.cs);
    if (rc != B200_OK) return rc;
    a.cx = (double)center_x; a.cy = (double)center_y; a.w = width; a.h = height;
    a.xmap = (uint8_t*)xmap; a.xstep = (int)xmap_step; a.ymap = (uint8_t*)ymap; a.ystep = (int)ymap_step;
    polar_map_kernel<<<dim3((width + 255) / 256, height), 256, 0, (cudaStream_t)stream>>>(a);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return report_cuda_error("warpPolar map kernel launch", e);
    count_launch();
    return B200_OK;
}

// cv::warpPolar without WARP_INVERSE_MAP: the maps above + cv::remap(src, dst, mapx, mapy, flags & INTER_MAX, WARP_FILL_OUTLIERS ?
// BORDER_CONSTANT : BORDER_TRANSPARENT) (imgwarp.cpp:3223) on one stream
extern "C" __attribute__((visibility("default")))
int b200_warpPolar_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                       void* dst, size_t dst_step, int dst_width, int dst_height, float center_x, float center_y, double max_radius, int flags,
                       float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, void* stream)
{
    B200_TRACE();
    int rc = b200_warpPolar_buildMaps_dev(dst_width, dst_height, center_x, center_y, max_radius, flags, xmap, xmap_step, ymap, ymap_step, stream);
    if (rc != B200_OK) return rc;
    const double bv[4] = {0, 0, 0, 0};
    return b200_remap32f_dev(src_type, src, src_step, src_width, src_height, dst, dst_step, dst_width, dst_height,
                             xmap, xmap_step, ymap, ymap_step, flags & 7, (flags & 8) ? 0 : 5, bv, stream);
}
