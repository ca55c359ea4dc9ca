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
} // namespace b200

extern "C" __attribute__((visibility("default")))
int b200_fisheye_initUndistortRectifyMap_dev(const double K[9], const double D[4], const double iR[9], int width, int height,
                                             int m1type, void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream)
{
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
    int rc = b200_fisheye_initUndistortRectifyMap_dev(K, D, iKnew, dst_width, dst_height, 11, map1, map1_step, map2, map2_step, stream);
    if (rc != B200_OK) return rc;
    const double bv[4] = {0, 0, 0, 0};
    return b200_remap16s_dev(src_type, src, src_step, src_width, src_height, dst, dst_step, dst_width, dst_height,
                             (const short*)map1, map1_step, (const unsigned short*)map2, map2_step, 1 /* INTER_LINEAR */, 0 /* BORDER_CONSTANT */,
                             bv, stream);
}
