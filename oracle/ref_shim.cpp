// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Thin extern "C" shim over the UNMODIFIED reference implementation (compiled from
// /root/reference by oracle/build_ref.sh into oracle/_ref/libcvref.so).  It lets the
// tests and bench.py's cpu_baseline / `--impl reference` legs call the reference's
// own CPU code on raw buffers through ctypes, with the same argument lists as the
// HAL replacement hooks (modules/imgproc/src/hal_replacement.hpp:369,387,428,483,506,530).
#include "opencv2/core.hpp"
#include "opencv2/core/utility.hpp"
#include "opencv2/imgproc.hpp"
#include "opencv2/imgproc/hal/hal.hpp"
#include "opencv2/dnn.hpp"
#include "opencv2/calib3d.hpp"
#include "distortion_model.hpp"
#include "opencv2/stitching/detail/warpers.hpp"          // calib3d/src (on the include path): cv::detail::computeTiltProjectionMatrix
#include <vector>
#include <cstring>

// dnn_utils.cpp references cv::cvtColor only on its NHWC + swapRB branch, which this library never takes; color.cpp is
// not part of the build, so the symbol is satisfied here.
namespace cv {
void cvtColor(InputArray, OutputArray, int, int, AlgorithmHint) { CV_Error(Error::StsNotImplemented, "cvtColor is not built into libcvref"); }
// fisheye.cpp's calibration entry points (never called through this library) reference three calib3d functions that are
// not part of the build
void Rodrigues(InputArray, OutputArray, OutputArray) { CV_Error(Error::StsNotImplemented, "Rodrigues is not built into libcvref"); }
bool solvePnP(InputArray, InputArray, InputArray, InputArray, OutputArray, OutputArray, bool, int) { CV_Error(Error::StsNotImplemented, "solvePnP is not built into libcvref"); }
bool solvePnPRansac(InputArray, InputArray, InputArray, InputArray, OutputArray, OutputArray, bool, int, float, double, OutputArray, int)
{ CV_Error(Error::StsNotImplemented, "solvePnPRansac is not built into libcvref"); }
}

using namespace cv;

extern "C" {

// -- HAL-level entries: exactly what sits below the cv_hal_* hooks ------------------
int cvref_hal_warpAffine(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                         uchar* dst, size_t dst_step, int dw, int dh, const double M[6],
                         int interpolation, int borderType, const double borderValue[4])
{
    try {
        cv::hal::warpAffine(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, M,
                            interpolation, borderType, borderValue);
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_hal_warpPerspective(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                              uchar* dst, size_t dst_step, int dw, int dh, const double M[9],
                              int interpolation, int borderType, const double borderValue[4])
{
    try {
        cv::hal::warpPerspective(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, M,
                                 interpolation, borderType, borderValue);
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_hal_resize(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                     uchar* dst, size_t dst_step, int dw, int dh,
                     double inv_scale_x, double inv_scale_y, int interpolation)
{
    try {
        cv::hal::resize(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh,
                        inv_scale_x, inv_scale_y, interpolation);
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

static int remap_common(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                        uchar* dst, size_t dst_step, int dw, int dh,
                        const Mat& m1, const Mat& m2, int interpolation, int borderType,
                        const double bv[4])
{
    try {
        Mat s(sh, sw, src_type, const_cast<uchar*>(src), src_step);
        Mat d(dh, dw, src_type, dst, dst_step);
        uchar* before = d.data;
        cv::remap(s, d, m1, m2, interpolation, borderType, Scalar(bv[0], bv[1], bv[2], bv[3]));
        if (d.data != before) return -2;
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_remap32f(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                   uchar* dst, size_t dst_step, int dw, int dh,
                   float* mapx, size_t mapx_step, float* mapy, size_t mapy_step,
                   int interpolation, int borderType, const double bv[4])
{
    Mat m1(dh, dw, CV_32FC1, mapx, mapx_step), m2(dh, dw, CV_32FC1, mapy, mapy_step);
    return remap_common(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, m1, m2, interpolation, borderType, bv);
}

int cvref_remap32fc2(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                     uchar* dst, size_t dst_step, int dw, int dh,
                     float* map, size_t map_step, int interpolation, int borderType, const double bv[4])
{
    Mat m1(dh, dw, CV_32FC2, map, map_step);
    return remap_common(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, m1, Mat(), interpolation, borderType, bv);
}

int cvref_remap16s(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                   uchar* dst, size_t dst_step, int dw, int dh,
                   short* mapx, size_t mapx_step, ushort* mapy, size_t mapy_step,
                   int interpolation, int borderType, const double bv[4])
{
    Mat m1(dh, dw, CV_16SC2, mapx, mapx_step), m2;
    if (mapy) m2 = Mat(dh, dw, CV_16UC1, mapy, mapy_step);
    return remap_common(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, m1, m2, interpolation, borderType, bv);
}

// -- public-API entries (for the host-side mirror: matrix inversion, dsize/fx/fy rules) --
int cvref_cv_warpAffine(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                        uchar* dst, size_t dst_step, int dw, int dh, const double M[6],
                        int flags, int borderType, const double bv[4])
{
    try {
        Mat s(sh, sw, src_type, const_cast<uchar*>(src), src_step);
        Mat d(dh, dw, src_type, dst, dst_step);
        Mat m(2, 3, CV_64F, const_cast<double*>(M));
        cv::warpAffine(s, d, m, Size(dw, dh), flags, borderType, Scalar(bv[0], bv[1], bv[2], bv[3]));
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_cv_warpPerspective(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                             uchar* dst, size_t dst_step, int dw, int dh, const double M[9],
                             int flags, int borderType, const double bv[4])
{
    try {
        Mat s(sh, sw, src_type, const_cast<uchar*>(src), src_step);
        Mat d(dh, dw, src_type, dst, dst_step);
        Mat m(3, 3, CV_64F, const_cast<double*>(M));
        cv::warpPerspective(s, d, m, Size(dw, dh), flags, borderType, Scalar(bv[0], bv[1], bv[2], bv[3]));
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

// cv::resize with (dsize | fx,fy); writes the resulting size to out_wh and the data to dst
// (dst must be large enough: caller computes the size with the same rule first via dst==NULL).
int cvref_cv_resize(int src_type, const uchar* src, size_t src_step, int sw, int sh,
                    uchar* dst, size_t dst_step, int dw, int dh, double fx, double fy,
                    int interpolation, int* out_wh)
{
    try {
        Mat s(sh, sw, src_type, const_cast<uchar*>(src), src_step);
        Mat d;
        cv::resize(s, d, Size(dw, dh), fx, fy, interpolation);
        if (out_wh) { out_wh[0] = d.cols; out_wh[1] = d.rows; }
        if (dst) {
            Mat o(d.rows, d.cols, src_type, dst, dst_step);
            d.copyTo(o);
        }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_convertMaps(float* mapx, size_t mapx_step, float* mapy, size_t mapy_step, int w, int h,
                      short* out1, size_t out1_step, ushort* out2, size_t out2_step, int nninterpolation)
{
    try {
        Mat m1(h, w, CV_32FC1, mapx, mapx_step), m2(h, w, CV_32FC1, mapy, mapy_step);
        Mat o1, o2;
        cv::convertMaps(m1, m2, o1, o2, CV_16SC2, nninterpolation != 0);
        Mat d1(h, w, CV_16SC2, out1, out1_step);
        o1.copyTo(d1);
        if (!nninterpolation && out2) { Mat d2(h, w, CV_16UC1, out2, out2_step); o2.copyTo(d2); }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_convertMaps_any(int m1type, void* map1, size_t s1, int m2type, void* map2, size_t s2,
                          int dtype, void* dst1, size_t ds1, void* dst2, size_t ds2, int w, int h, int nn)
{
    try {
        Mat m1(h, w, m1type, map1, s1), m2;
        if (map2) m2 = Mat(h, w, m2type, map2, s2);
        Mat o1, o2;
        cv::convertMaps(m1, m2, o1, o2, dtype, nn != 0);
        Mat d1(h, w, o1.type(), dst1, ds1);
        o1.copyTo(d1);
        if (!o2.empty() && dst2) { Mat d2(h, w, o2.type(), dst2, ds2); o2.copyTo(d2); }
        return o1.type() | (o2.empty() ? 0 : 0x10000);
    } catch (const cv::Exception&) { return -1; }
}

// cv::dnn::blobFromImagesWithParams (dnn/src/dnn_utils.cpp, compiled into this library on its own)
int cvref_blobFromImages(int src_type, unsigned char* src, size_t sfs, size_t src_step, int sw, int sh,
                         float* blob, int bw, int bh, int n_frames, const double* scalefactor, const double* mean, int swapRB)
{
    try {
        std::vector<Mat> imgs;
        for (int f = 0; f < n_frames; f++) imgs.push_back(Mat(sh, sw, src_type, src + (size_t)f * sfs, src_step));
        cv::dnn::Image2BlobParams prm(Scalar(scalefactor[0], scalefactor[1], scalefactor[2], scalefactor[3]), Size(bw, bh),
                                      Scalar(mean[0], mean[1], mean[2], mean[3]), swapRB != 0, CV_32F);
        Mat out;
        cv::dnn::blobFromImagesWithParams(imgs, out, prm);
        if (out.dims != 4 || out.size[0] != n_frames || out.size[2] != bh || out.size[3] != bw || out.type() != CV_32F) return -2;
        std::memcpy(blob, out.ptr<float>(), out.total() * sizeof(float));
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

// the same with a padding mode (0 DNN_PMODE_NULL, 1 DNN_PMODE_CROP_CENTER, 2 DNN_PMODE_LETTERBOX) and the letterbox border value
int cvref_blobFromImagesWithParams(int src_type, unsigned char* src, size_t sfs, size_t src_step, int sw, int sh,
                                   float* blob, int bw, int bh, int n_frames, const double* scalefactor, const double* mean, int swapRB,
                                   int paddingmode, const double* borderValue)
{
    try {
        std::vector<Mat> imgs;
        for (int f = 0; f < n_frames; f++) imgs.push_back(Mat(sh, sw, src_type, src + (size_t)f * sfs, src_step));
        cv::dnn::Image2BlobParams prm(Scalar(scalefactor[0], scalefactor[1], scalefactor[2], scalefactor[3]), Size(bw, bh),
                                      Scalar(mean[0], mean[1], mean[2], mean[3]), swapRB != 0, CV_32F, cv::dnn::DNN_LAYOUT_NCHW,
                                      (cv::dnn::ImagePaddingMode)paddingmode,
                                      borderValue ? Scalar(borderValue[0], borderValue[1], borderValue[2], borderValue[3]) : Scalar());
        Mat out;
        cv::dnn::blobFromImagesWithParams(imgs, out, prm);
        if (out.dims != 4 || out.size[0] != n_frames || out.size[2] != bh || out.size[3] != bw || out.type() != CV_32F) return -2;
        std::memcpy(blob, out.ptr<float>(), out.total() * sizeof(float));
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

int cvref_getRotationMatrix2D(double cx, double cy, double angle, double scale, double* M6)
{
    Mat m = cv::getRotationMatrix2D(Point2f((float)cx, (float)cy), angle, scale);
    std::memcpy(M6, m.ptr<double>(), 6 * sizeof(double));
    return 0;
}

int cvref_getPerspectiveTransform(const float* src8, const float* dst8, double* M9)
{
    try {
        Point2f s[4], d[4];
        for (int i = 0; i < 4; i++) { s[i] = Point2f(src8[2*i], src8[2*i+1]); d[i] = Point2f(dst8[2*i], dst8[2*i+1]); }
        Mat m = cv::getPerspectiveTransform(s, d);
        std::memcpy(M9, m.ptr<double>(), 9 * sizeof(double));
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

// cv::RNG fill, so tests can regenerate the reference tests' inputs (e.g. Resize.Area_half RNG(17))
int cvref_rng_fill(unsigned long long seed, int type, uchar* data, size_t step, int w, int h,
                   double lo, double hi, int saturate)
{
    try {
        RNG rng(seed);
        Mat m(h, w, type, data, step);
        rng.fill(m, RNG::UNIFORM, lo, hi, saturate != 0);
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

// cv::fisheye::initUndistortRectifyMap (calib3d/src/fisheye.cpp:530-631).  R / P may be null (identity).  m1type: CV_16SC2 (11) or
// CV_32FC1 (5).  iR_out (optional) receives the reference's own (P*R)^-1 (Matx inv(DECOMP_SVD), :575), the one input the
// restatements take from outside.
int cvref_fisheye_initUndistortRectifyMap(const double* K9, const double* D4, const double* R9, const double* P9, int w, int h,
                                          int m1type, void* map1, size_t step1, void* map2, size_t step2, double* iR_out)
{
    try {
        Mat K(3, 3, CV_64F, (void*)K9), D, R, P;
        if (D4) D = Mat(1, 4, CV_64F, (void*)D4);
        if (R9) R = Mat(3, 3, CV_64F, (void*)R9);
        if (P9) P = Mat(3, 3, CV_64F, (void*)P9);
        Mat m1, m2;
        cv::fisheye::initUndistortRectifyMap(K, D, R, P, Size(w, h), m1type, m1, m2);
        const size_t row1 = (size_t)w * m1.elemSize(), row2 = (size_t)w * m2.elemSize();
        for (int y = 0; y < h; y++) {
            std::memcpy((uchar*)map1 + y * step1, m1.ptr(y), row1);
            std::memcpy((uchar*)map2 + y * step2, m2.ptr(y), row2);
        }
        if (iR_out) {
            Matx33d RR = Matx33d::eye(), PP = Matx33d::eye();
            if (R9) RR = Matx33d(R9);
            if (P9) PP = Matx33d(P9);
            Matx33d iR = (PP * RR).inv(DECOMP_SVD);
            std::memcpy(iR_out, iR.val, 9 * sizeof(double));
        }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}

// cv::initUndistortRectifyMap (calib3d/src/undistort.dispatch.cpp:82-161).  dist: 4, 5, 8, 12 or 14 coefficients; R / newK may be null.
// m1type: CV_16SC2 (11), CV_32FC1 (5) or CV_32FC2 (13).  iR_out / tilt_out (optional) receive the reference's own
// (newK(:,0:3) * R)^-1 (LU inverse, :123) and tilt matrix (:155-157), the inputs the restatements take from outside.
int cvref_initUndistortRectifyMap(const double* K9, const double* dist, int ndist, const double* R9, const double* newK9, int w, int h,
                                  int m1type, void* map1, size_t step1, void* map2, size_t step2, double* iR_out, double* tilt_out)
{
    try {
        Mat K(3, 3, CV_64F, (void*)K9), D, R, P;
        if (dist && ndist > 0) D = Mat(1, ndist, CV_64F, (void*)dist);
        if (R9) R = Mat(3, 3, CV_64F, (void*)R9);
        if (newK9) P = Mat(3, 3, CV_64F, (void*)newK9);
        Mat m1, m2;
        cv::initUndistortRectifyMap(K, D, R, P, Size(w, h), m1type, m1, m2);
        for (int y = 0; y < h; y++) {
            std::memcpy((uchar*)map1 + y * step1, m1.ptr(y), (size_t)w * m1.elemSize());
            if (!m2.empty() && map2) std::memcpy((uchar*)map2 + y * step2, m2.ptr(y), (size_t)w * m2.elemSize());
        }
        if (iR_out) {
            Mat_<double> Rm = Mat_<double>::eye(3, 3), A(K), Ar;
            Ar = newK9 ? Mat_<double>(P) : getDefaultNewCameraMatrix(A, Size(w, h), true);
            if (R9) Rm = Mat_<double>(R);
            Mat_<double> iR = (Ar.colRange(0, 3) * Rm).inv(DECOMP_LU);
            std::memcpy(iR_out, iR.ptr<double>(), 9 * sizeof(double));
        }
        if (tilt_out) {
            Matx33d t = Matx33d::eye();
            if (dist && ndist >= 14) cv::detail::computeTiltProjectionMatrix(dist[12], dist[13], &t);
            std::memcpy(tilt_out, t.val, 9 * sizeof(double));
        }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
// cv::undistort (calib3d/src/undistort.dispatch.cpp:284-328): dist = 4..14 coefficients or null, newK9 may be null
int cvref_undistort(int type, const void* src, size_t src_step, int w, int h, void* dst, size_t dst_step,
                    const double* K9, const double* dist, int ndist, const double* newK9)
{
    try {
        Mat s(h, w, type, (void*)src, src_step), d(h, w, type, dst, dst_step), d0 = d;
        Mat K(3, 3, CV_64F, (void*)K9), D, P;
        if (dist && ndist > 0) D = Mat(1, ndist, CV_64F, (void*)dist);
        if (newK9) P = Mat(3, 3, CV_64F, (void*)newK9);
        cv::undistort(s, d, K, D, P);
        if (d.data != d0.data) return -2;
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
// cv::pyrDown (imgproc/src/pyramids.cpp:1266-1310); dw / dh = 0: the default size ((w+1)/2, (h+1)/2); out_wh receives the size
int cvref_pyrDown(int type, const uchar* src, size_t src_step, int sw, int sh, uchar* dst, size_t dst_step, int dw, int dh,
                  int borderType, int* out_wh)
{
    try {
        Mat s(sh, sw, type, const_cast<uchar*>(src), src_step), d;
        cv::pyrDown(s, d, Size(dw, dh), borderType);
        if (out_wh) { out_wh[0] = d.cols; out_wh[1] = d.rows; }
        if (dst) { Mat o(d.rows, d.cols, type, dst, dst_step); d.copyTo(o); }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
// cv::pyrUp (imgproc/src/pyramids.cpp:1374-1416); dw / dh = 0: the default size (2w, 2h)
int cvref_pyrUp(int type, const uchar* src, size_t src_step, int sw, int sh, uchar* dst, size_t dst_step, int dw, int dh, int borderType)
{
    try {
        Mat s(sh, sw, type, const_cast<uchar*>(src), src_step), d;
        cv::pyrUp(s, d, Size(dw, dh), borderType);
        if (dw && (d.cols != dw || d.rows != dh)) return -2;
        Mat o(d.rows, d.cols, type, dst, dst_step); d.copyTo(o);
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
// cv::warpPolar (imgproc/src/imgwarp.cpp:3163-3281); dst is pre-filled by the caller (BORDER_TRANSPARENT keeps it where the map leaves)
int cvref_warpPolar(int type, const uchar* src, size_t src_step, int sw, int sh, uchar* dst, size_t dst_step, int dw, int dh,
                    float cx, float cy, double maxRadius, int flags)
{
    try {
        Mat s(sh, sw, type, const_cast<uchar*>(src), src_step), d(dh, dw, type, dst, dst_step);
        const uchar* d0 = d.data;
        cv::warpPolar(s, d, Size(dw, dh), Point2f(cx, cy), maxRadius, flags);
        if (d.data != d0) return -2;
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
// cv::detail::{Plane,Spherical,Cylindrical}Warper (stitching/src/warpers.cpp, detail/warpers_inl.hpp): kind 0 / 1 / 2.
// K9 / R9 / T3: CV_32F.  roi_out = {tl.x, tl.y, width, height} of the maps (buildMaps' Rect + 1), kr_out / t_out = the projector's
// k_rinv and t after setCameraParams (what the map loop reads).  xmap / ymap may be null (size query).
static cv::Ptr<cv::detail::RotationWarper> make_warper(int kind, float scale)
{
    if (kind == 0) return cv::makePtr<cv::detail::PlaneWarper>(scale);
    if (kind == 1) return cv::makePtr<cv::detail::SphericalWarper>(scale);
    return cv::makePtr<cv::detail::CylindricalWarper>(scale);
}
int cvref_stitch_buildMaps(int kind, int w, int h, const float* K9, const float* R9, const float* T3, float scale,
                           float* xmap, size_t xstep, float* ymap, size_t ystep, int* roi_out, float* kr_out, float* t_out)
{
    try {
        Mat K(3, 3, CV_32F, (void*)K9), R(3, 3, CV_32F, (void*)R9), T(3, 1, CV_32F, (void*)T3), xm, ym;
        Rect r;
        if (kind == 0) { cv::detail::PlaneWarper wp(scale); r = wp.buildMaps(Size(w, h), K, R, T, xm, ym); }
        else r = make_warper(kind, scale)->buildMaps(Size(w, h), K, R, xm, ym);
        if (roi_out) { roi_out[0] = r.x; roi_out[1] = r.y; roi_out[2] = xm.cols; roi_out[3] = xm.rows; }
        cv::detail::PlaneProjector pr; pr.scale = scale;
        pr.setCameraParams(K, R, kind == 0 ? T : Mat::zeros(3, 1, CV_32F));
        if (kr_out) std::memcpy(kr_out, pr.k_rinv, 9 * sizeof(float));
        if (t_out) std::memcpy(t_out, pr.t, 3 * sizeof(float));
        if (xmap && ymap)
            for (int y = 0; y < xm.rows; y++) {
                std::memcpy((uchar*)xmap + y * xstep, xm.ptr(y), xm.cols * sizeof(float));
                std::memcpy((uchar*)ymap + y * ystep, ym.ptr(y), ym.cols * sizeof(float));
            }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
// RotationWarper::warp: maps + cv::remap; dst must hold roi_out[2] x roi_out[3] pixels (query with cvref_stitch_buildMaps first)
int cvref_stitch_warp(int kind, int type, const uchar* src, size_t src_step, int w, int h, const float* K9, const float* R9, const float* T3,
                      float scale, int interp, int border, uchar* dst, size_t dst_step, int* tl_out)
{
    try {
        Mat K(3, 3, CV_32F, (void*)K9), R(3, 3, CV_32F, (void*)R9), T(3, 1, CV_32F, (void*)T3), s(h, w, type, (void*)src, src_step), d;
        Point tl;
        if (kind == 0) { cv::detail::PlaneWarper wp(scale); tl = wp.warp(s, K, R, T, interp, border, d); }
        else tl = make_warper(kind, scale)->warp(s, K, R, interp, border, d);
        if (tl_out) { tl_out[0] = tl.x; tl_out[1] = tl.y; tl_out[2] = d.cols; tl_out[3] = d.rows; }
        if (dst) { Mat o(d.rows, d.cols, type, dst, dst_step); d.copyTo(o); }
    } catch (const cv::Exception&) { return -1; }
    return 0;
}
int cvref_have_avx2(void) { return cv::checkHardwareSupport(CV_CPU_AVX2) ? 1 : 0; }

void cvref_setNumThreads(int n) { cv::setNumThreads(n); }
int  cvref_getNumThreads(void) { return cv::getNumThreads(); }
const char* cvref_version(void) { return CV_VERSION; }
const char* cvref_parallel_backend(void)
{
#if defined HAVE_PTHREADS_PF
    return "pthreads";
#else
    return "other";
#endif
}

} // extern "C"
