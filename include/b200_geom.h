/*
 * b200_geom.h -- C ABI of libb200geom.so: OpenCV's imgproc geometric-transform hot path
 * (cv::warpAffine, cv::warpPerspective, cv::remap, cv::resize) as hand-written sm_100a kernels.
 *
 * Plain pointers and sizes only.  Three layers, all with the reference's argument meaning:
 *
 *   b200_hal_*      HOST pointers.  These are drop-in bodies for OpenCV's HAL replacement hooks
 *                   (hal/cuda_b200/cuda_b200_hal.hpp maps cv_hal_* onto them).  H2D copy, kernel,
 *                   D2H copy and a stream sync happen inside the call.
 *   b200_*_dev      DEVICE pointers + a CUDA stream; asynchronous; identical arithmetic.
 *   b200_*_batch    device-resident batches of independent frames (one launch for many frames) and
 *                   the multi-GPU frame scheduler (contiguous frame blocks per GPU, no collective).
 *
 * Return codes follow modules/core/include/opencv2/core/hal/interface.h:9-11:
 *   0 = CV_HAL_ERROR_OK, 1 = CV_HAL_ERROR_NOT_IMPLEMENTED (caller falls back to its CPU path;
 *   dst untouched), -1 = CV_HAL_ERROR_UNKNOWN (CUDA failure; b200_last_error() has the text).
 * There is NO CPU fallback inside this library.
 *
 * Supported: depth CV_8U / CV_16U / CV_16S / CV_32F (INTER_NEAREST[_EXACT]: any depth), 1..4 channels (the warpAffine /
 * warpPerspective / remap HOOKS also more: they run groups of 4 channels; not with INTER_CUBIC / LANCZOS4, as in the reference),
 * INTER_NEAREST / INTER_LINEAR / INTER_AREA / INTER_CUBIC / INTER_LANCZOS4 in warpAffine, warpPerspective, remap
 * (imgwarp.cpp:268-1059) and resize (resize.cpp:1993-2155; LANCZOS4 coefficients are built on the host with the host libm and cached
 * on the device per geometry), plus resize's bit-exact modes INTER_LINEAR_EXACT (CV_8U, CV_16U, CV_8S, CV_16S, CV_32S) and
 * INTER_NEAREST_EXACT; CV_64F with every interpolation (not with BORDER_TRANSPARENT);
 * BORDER_CONSTANT / REPLICATE / REFLECT / WRAP / REFLECT_101 / TRANSPARENT
 * (the host-pointer hooks then read dst as well: pixels mapped outside the source keep its contents).
 * Everything else returns NOT_IMPLEMENTED.
 */
#ifndef B200_GEOM_H
#define B200_GEOM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_NOT_IMPLEMENTED 1
#define B200_ERROR (-1)

/* ------------------------------------------------------------------ HAL hooks (host pointers) --
 * Each replaces the hook of the same name in modules/imgproc/src/hal_replacement.hpp. */

/* replaces cv_hal_resize (hal_replacement.hpp:369; called from resize.cpp:3648).  Receives the raw
 * `interpolation`; applies the reference's own mode rules (INTER_LINEAR at exact 2x -> 2x2 box,
 * resize.cpp:3817; INTER_AREA with a scale < 1 -> area-mode bilinear, resize.cpp:3822,3905-3916). */
int b200_hal_resize(int src_type, const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                    unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height,
                    double inv_scale_x, double inv_scale_y, int interpolation);

/* replaces cv_hal_warpAffine (hal_replacement.hpp:387; called from imgwarp.cpp:2255).
 * M is the inverse (dst->src) 2x3 matrix, exactly as cv::warpAffine hands it to the hook. */
int b200_hal_warpAffine(int src_type, const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                        unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height,
                        const double M[6], int interpolation, int borderType, const double borderValue[4]);

/* replaces cv_hal_warpPerspective (hal_replacement.hpp:428; called from imgwarp.cpp:2749). */
int b200_hal_warpPerspective(int src_type, const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                             unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height,
                             const double M[9], int interpolation, int borderType, const double borderValue[4]);

/* replaces cv_hal_remap32f (hal_replacement.hpp:483-487; called from imgwarp.cpp:1608-1612). */
int b200_hal_remap32f(int src_type, const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                      unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height,
                      float* mapx, size_t mapx_step, float* mapy, size_t mapy_step,
                      int interpolation, int border_type, const double border_value[4]);

/* replaces cv_hal_remap32fc2 (hal_replacement.hpp:506-509; called from imgwarp.cpp:1613-1617). */
int b200_hal_remap32fc2(int src_type, const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                        unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height,
                        float* map, size_t map_step, int interpolation, int border_type, const double border_value[4]);

/* replaces cv_hal_remap16s (hal_replacement.hpp:530-534; called from imgwarp.cpp:1618-1622).
 * mapy may be NULL (CV_16SC2 only). */
int b200_hal_remap16s(int src_type, const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                      unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height,
                      short* mapx, size_t mapx_step, unsigned short* mapy, size_t mapy_step,
                      int interpolation, int border_type, const double border_value[4]);

/* Hooks return NOT_IMPLEMENTED for dst_width*dst_height below this many pixels (default 1<<16, env
 * OPENCV_B200_MIN_PIXELS): the reference's CPU fallback of warpAffine/warpPerspective re-enters
 * cv_hal_remap16s once per <=4096-px tile (imgwarp.cpp:2040-2044,2723-2727).  Returns the old value. */
long long b200_hal_set_min_pixels(long long min_pixels);

/* Which cells the per-image hooks take.  0 = "profitable" (default; env OPENCV_B200_POLICY=profitable): the hooks return
 * NOT_IMPLEMENTED for the cells in which one image's upload + kernel + download lose to the reference's CPU code on a many-core host
 * whatever the kernel does (measured with the reference's own perf grids, profiles/r2_perf_view.md): every INTER_NEAREST cell,
 * cv::resize with INTER_LINEAR / INTER_LINEAR_EXACT / the 2×2 INTER_AREA box, and INTER_LINEAR remaps of more than 2^20 pixels,
 * whose maps cross PCIe with every call.  1 = "all" (env OPENCV_B200_POLICY=all): every supported cell -- for hosts with few cores,
 * for pinned buffers, and for tests.  The b200_*_dev and *_batch entries are not affected.  Returns the old value. */
int b200_hal_set_policy(int all);

/* ------------------------------------------------------ device-resident entries (async, stream) --
 * Same arguments as the hooks above, but every image / map pointer is DEVICE memory on the current
 * device, and `stream` is a cudaStream_t (NULL = default stream).  No host synchronisation. */
int b200_resize_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                    void* dst, size_t dst_step, int dst_width, int dst_height,
                    double inv_scale_x, double inv_scale_y, int interpolation, void* stream);
int b200_warpAffine_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                        void* dst, size_t dst_step, int dst_width, int dst_height,
                        const double M[6], int interpolation, int borderType, const double borderValue[4], void* stream);
int b200_warpPerspective_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                             void* dst, size_t dst_step, int dst_width, int dst_height,
                             const double M[9], int interpolation, int borderType, const double borderValue[4], void* stream);
int b200_remap32f_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                      void* dst, size_t dst_step, int dst_width, int dst_height,
                      const float* mapx, size_t mapx_step, const float* mapy, size_t mapy_step,
                      int interpolation, int border_type, const double border_value[4], void* stream);
int b200_remap32fc2_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                        void* dst, size_t dst_step, int dst_width, int dst_height,
                        const float* map, size_t map_step, int interpolation, int border_type,
                        const double border_value[4], void* stream);
int b200_remap16s_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                      void* dst, size_t dst_step, int dst_width, int dst_height,
                      const short* mapx, size_t mapx_step, const unsigned short* mapy, size_t mapy_step,
                      int interpolation, int border_type, const double border_value[4], void* stream);

/* cv::convertMaps (imgwarp.cpp:1682-1986) on device-resident maps: float maps (CV_32FC1 pair or CV_32FC2) to the
 * fixed-point pair CV_16SC2 + CV_16UC1 that remap's fastest path reads, and back; merge / split / copy for the
 * remaining pairs.  Types are OpenCV type codes (CV_16SC2 = 11, CV_16UC1 = 2, CV_32FC1 = 5, CV_32FC2 = 13);
 * map2 may be NULL (then map2_type is ignored), dstmap1_type <= 0 picks the reference's default, dstmap2 is
 * written only where the reference creates it (!nninterpolation and dstmap1_type != CV_32FC2).
 * The reference has no HAL hook for convertMaps; a maintainer would call this from cv::convertMaps itself. */
int b200_convertMaps_dev(int map1_type, const void* map1, size_t map1_step, int map2_type, const void* map2, size_t map2_step,
                         int dstmap1_type, void* dstmap1, size_t dstmap1_step, void* dstmap2, size_t dstmap2_step,
                         int width, int height, int nninterpolation, void* stream);

/* cv::dnn::blobFromImages for CV_8U frames on the device (dnn/src/dnn_utils.cpp:130-241, :275-353; a caller of
 * cv::resize worth fusing): n_frames frames of src_width x src_height are resized to
 * blob_width x blob_height with INTER_LINEAR exactly as cv::resize does (no resize when the sizes agree) and written
 * as a planar fp32 NCHW blob [n_frames][channels][blob_height][blob_width] with (v - mean[c]) * scalefactor[c];
 * swapRB exchanges channels 0 and 2, and mean / scalefactor index the BLOB channel as in the reference.
 * One kernel, no intermediate image.  1, 3 or 4 channels. */
int b200_blobFromImages_dev(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                            float* blob, int blob_width, int blob_height, int n_frames,
                            const double scalefactor[4], const double mean[4], int swapRB, void* stream);

/* cv::dnn::blobFromImagesWithParams with a padding mode (dnn_utils.cpp:303-337), same fused kernel: paddingmode 0 = DNN_PMODE_NULL (the
 * entry above), 1 = DNN_PMODE_CROP_CENTER (resize by the larger of the two float ratios through (fx, fy), then the centred crop, :309-317),
 * 2 = DNN_PMODE_LETTERBOX (resize by the smaller ratio to (int(w*f), int(h*f)), then borderValue[channel] around it, :318-331;
 * borderValue NULL = 0).  Geometries the reference would reject (empty resized image, crop outside it) return NOT_IMPLEMENTED. */
int b200_blobFromImagesWithParams_dev(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                      float* blob, int blob_width, int blob_height, int n_frames,
                                      const double scalefactor[4], const double mean[4], int swapRB,
                                      int paddingmode, const double borderValue[4], void* stream);

/* --------------------------------------------------------------- batched, device-resident -------
 * `n_frames` independent frames of identical geometry; frame i of a batch starts at
 * base + i*frame_stride bytes.  One launch covers the whole batch (blockIdx.z = frame).
 * New API: the reference has no batched entry; arithmetic per frame is that of the entries above. */
int b200_resize_batch_dev(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                          void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                          int n_frames, double inv_scale_x, double inv_scale_y, int interpolation, void* stream);
int b200_warpAffine_batch_dev(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                              void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                              int n_frames, const double M[6], int interpolation, int borderType,
                              const double borderValue[4], void* stream);

int b200_warpPerspective_batch_dev(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                   void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                                   int n_frames, const double M[9], int interpolation, int borderType,
                                   const double borderValue[4], void* stream);
/* n_frames frames through ONE map (a video stream through a fixed undistortion / rectification map): the map is read
 * once from HBM and then served by L2.  map1_type: CV_32FC1 = 5 (map2 = the y plane), CV_32FC2 = 13 (map2 ignored),
 * CV_16SC2 = 11 (map2 = the CV_16UC1 fraction plane or NULL); `interpolation` may carry WARP_RELATIVE_MAP. */
int b200_remap_batch_dev(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                         void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames,
                         int map1_type, const void* map1, size_t map1_step, const void* map2, size_t map2_step,
                         int interpolation, int borderType, const double borderValue[4], void* stream);

/* The BASELINE "frame stream" pipeline: per frame resize(src -> mid) then warpAffine(mid -> dst).
 * `mid` is caller-provided device scratch of n_frames * mid_frame_stride bytes (mid_frame_stride >=
 * mid_step*mid_height); its contents after the call are unspecified.  Equivalent to cv::resize followed by
 * cv::warpAffine per frame.  For CV_8U, an exact 2x downscale and a bilinear warp, batches of more than ~20 frames
 * run as ONE persistent kernel whose half-size frames stay in L2 (a ring in the first part of `mid`, dependency
 * counters behind it); everything else runs as two batched launches.  Same results either way. */
int b200_resize_warpAffine_batch_dev(int src_type,
                                     const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                     void* mid, size_t mid_frame_stride, size_t mid_step, int mid_width, int mid_height,
                                     void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                                     int n_frames, int resize_interpolation,
                                     const double M[6], int warp_interpolation, int borderType,
                                     const double borderValue[4], void* stream);

/* The same pipeline for a stream of frames in HOST memory (new API, no counterpart in the reference: the hooks
 * above are one synchronous call per image, so each pays its own upload, kernel and download back to back).
 * Frames f = 0..n_frames-1 at src + f*src_frame_stride are resized to mid_width x mid_height and warped into
 * dst + f*dst_frame_stride; equivalent to cv::resize + cv::warpAffine per frame (imgproc.hpp:2517,2540).
 * Work is cut into chunks of `chunk_frames` frames (0 = choose: about 24 MB of source per chunk) and the upload of
 * chunk k+1, the kernels of chunk k and the download of chunk k-1 run concurrently on three streams with three
 * staging slots, so the call is bound by the slower PCIe direction rather than by the sum of the stages.
 * Returns when every dst frame is complete in host memory.  Pinned (page-locked) buffers are copied by DMA;
 * pageable ones work but are staged by the driver.  Staging is per host thread, grow-only, released at thread exit. */
int b200_hal_resize_warpAffine_batch(int src_type,
                                     const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                     int mid_width, int mid_height,
                                     void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                                     int n_frames, int resize_interpolation,
                                     const double M[6], int warp_interpolation, int borderType,
                                     const double borderValue[4], int chunk_frames);

/* The same chunked host pipeline (three streams, three staging slots, PCIe-bound instead of the sum of the stages) around the single
 * batched operations, so that frame streams of BASELINE configs 2-4 need not pay one synchronous hook call per frame: n_frames frames
 * in host memory through cv::resize / cv::warpAffine / cv::warpPerspective with one size / matrix, or through cv::remap with ONE map
 * (map1_type as in b200_remap_batch_dev; the map is uploaded once).  Arguments as in the b200_*_batch_dev entries with host pointers;
 * chunk_frames = 0 chooses.  BORDER_TRANSPARENT returns NOT_IMPLEMENTED (the staging never sees the caller's dst). */
int b200_hal_resize_batch(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                          void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames,
                          double inv_scale_x, double inv_scale_y, int interpolation, int chunk_frames);
int b200_hal_warpAffine_batch(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                              void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames,
                              const double M[6], int interpolation, int borderType, const double borderValue[4], int chunk_frames);
int b200_hal_warpPerspective_batch(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                   void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames,
                                   const double M[9], int interpolation, int borderType, const double borderValue[4], int chunk_frames);
int b200_hal_remap_batch(int src_type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                         void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames,
                         int map1_type, const void* map1, size_t map1_step, const void* map2, size_t map2_step,
                         int interpolation, int borderType, const double borderValue[4], int chunk_frames);

/* Same pipeline for callers that cannot allocate device memory at all (a capture loop that must stay free of
 * allocator calls once it is running, a plugin with no memory budget of its own, code on a path where
 * cudaMalloc would stall other work on the GPU): the intermediate frame lives in a fixed array inside the
 * library instead of a caller buffer.  Has no counterpart in the reference.
 *   - Limit: the intermediate frame, rows padded to 256 bytes, must fit B200_NOALLOC_SCRATCH_BYTES (8 MiB, e.g.
 *     up to 1920x1080 CV_8UC3); larger requests return B200_NOALLOC_TOO_LARGE and explain on stderr.  Every GPU
 *     that loads the library holds those 8 MiB until the process ends.
 *   - It calls no cudaMalloc / cudaFree / allocating routine.  The first call on each device looks up the
 *     array's device address (cudaGetSymbolAddress); make that first call before starting any stream capture.  The
 *     address is looked up again on every call (it is a table lookup in the runtime, no allocation), so it stays right
 *     across a cudaDeviceReset() of that device.
 *   - Frames are processed one after another through the one shared array, which keeps whatever the last
 *     call left in it; nothing in it needs clearing between calls.  Two calls whose GPU work can overlap
 *     (different streams, different threads) race on the array: serialise them yourself. */
#define B200_NOALLOC_SCRATCH_BYTES (8u << 20)
#define B200_NOALLOC_TOO_LARGE (-2)
int b200_resize_warpAffine_noalloc_dev(int src_type,
                                       const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                       int mid_width, int mid_height,
                                       void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height,
                                       int n_frames, int resize_interpolation,
                                       const double M[6], int warp_interpolation, int borderType,
                                       const double borderValue[4], void* stream);

/* cv::fisheye::initUndistortRectifyMap (calib3d/src/fisheye.cpp:530-631) written straight into device memory, in either map
 * format cv::remap takes: m1type CV_16SC2 (11, the default for <= 0; map2 = CV_16UC1) or CV_32FC1 (5; map2 = CV_32FC1).
 * K = camera matrix (row-major 3x3; fx, fy, cx, cy are read), D = the four fisheye coefficients (NULL: zeros),
 * iR = (P(:, 0:3) * R)^-1, row-major, computed by the caller (the reference uses its SVD-based Matx inverse, :575).
 * Every step of the reference's fp64 sequence is reproduced, the row accumulators included; atan comes from the CUDA math
 * library instead of the host libm, so an entry whose fp64 value sits within an ulp of a rounding boundary can differ in its
 * last bit.  Returns B200_OK / B200_NOT_IMPLEMENTED (other map types, misaligned maps) / B200_ERROR. */
int b200_fisheye_initUndistortRectifyMap_dev(const double K[9], const double D[4], const double iR[9], int width, int height,
                                             int m1type, void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream);

/* replaces cv_hal_pyrdown (imgproc/src/hal_replacement.hpp:1461-1464; called from cv::pyrDown, pyramids.cpp:1292): 5x5 binomial
 * filter + decimation by 2.  CV_8U / CV_16U / CV_16S / CV_32F, 1..4 channels, BORDER_REPLICATE / REFLECT / WRAP / REFLECT_101
 * (| BORDER_ISOLATED); dst size within +-2 of half the source as the reference requires (pyramids.cpp:892-894).  CV_32F follows the
 * two evaluation orders of the reference's default x86 build (vector body / scalar tail, pyramids.cpp:322-404, 494-514, 955-1034).
 * Host pointers.  CV_64F and more than 4 channels return NOT_IMPLEMENTED (the CPU path runs). */
int b200_hal_pyrdown(const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                     unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height, int depth, int cn, int border_type);

/* the same on device pointers (asynchronous on `stream`), and over n_frames frames in one launch */
int b200_pyrDown_dev(int type, const void* src, size_t src_step, int src_width, int src_height,
                     void* dst, size_t dst_step, int dst_width, int dst_height, int border, void* stream);
int b200_pyrDown_batch_dev(int type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                           void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames, int border,
                           void* stream);

/* replaces cv_hal_pyrup (imgproc/src/hal_replacement.hpp:1508-1511; called from cv::pyrUp, pyramids.cpp:1391): each source pixel
 * becomes a 2x2 block filtered with the even (1 6 1) and odd (4 4) phases of the binomial kernel, / 64 (pyrUp_, pyramids.cpp:1036-1155).
 * CV_8U / CV_16U / CV_16S / CV_32F / CV_64F, 1..4 channels, BORDER_REFLECT_101 only (cv::pyrUp asserts BORDER_DEFAULT, :1378); dst
 * size 2n, 2n - 1 or 2n + 1 per axis (:1055-1056).  Host pointers.  Other depths, more than 4 channels, and a one-pixel-wide source
 * with a 3-column dst (the reference reads an unwritten cell there) return NOT_IMPLEMENTED. */
int b200_hal_pyrup(const unsigned char* src_data, size_t src_step, int src_width, int src_height,
                   unsigned char* dst_data, size_t dst_step, int dst_width, int dst_height, int depth, int cn, int border_type);

/* the same on device pointers (asynchronous on `stream`), and over n_frames frames in one launch */
int b200_pyrUp_dev(int type, const void* src, size_t src_step, int src_width, int src_height,
                   void* dst, size_t dst_step, int dst_width, int dst_height, int border, void* stream);
int b200_pyrUp_batch_dev(int type, const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                         void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames, int border,
                         void* stream);

/* cv::initUndistortRectifyMap (calib3d/src/undistort.dispatch.cpp:82-161) written straight into device memory: m1type CV_16SC2 (11,
 * default for <= 0; map2 = CV_16UC1), CV_32FC1 (5; map2 = CV_32FC1) or CV_32FC2 (13; map2 unused).  K = camera matrix (fx, fy, u0, v0
 * are read); dist12 = k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4 (NULL: zeros; shorter models: trailing zeros); iR = (newK(:,0:3) * R)^-1
 * and tilt = the 3x3 tilt projection matrix of (tauX, tauY) (NULL: identity), both row-major and computed by the caller as
 * undistort.dispatch.cpp:123, :155-157 do.  Bit-identical to the AVX2 dispatch variant a stock x86 build runs (vector body in
 * chunks of 8 pixels with fused multiply-adds, scalar tail for the last width % 8: undistort.simd.hpp:128-277). */
int b200_initUndistortRectifyMap_dev(const double K[9], const double dist12[12], const double iR[9], const double tilt[9],
                                     int width, int height, int m1type, void* map1, size_t map1_step, void* map2, size_t map2_step,
                                     void* stream);

/* The two matrices cv::initUndistortRectifyMap derives before its pixel loop, computed on the host exactly as the reference does:
 * iR_out = (newK(:,0:3) * R)^-1 (undistort.dispatch.cpp:123: cv::gemm's 3x3 form, core/src/matmul.simd.hpp:1470-1481, then
 * cv::invert's closed form, core/src/lapack.cpp:946-968) and tilt_out = computeTiltProjectionMatrix(tauX, tauY)
 * (calib3d/src/distortion_model.hpp:75-94; identity for (0, 0); may be NULL).  R = NULL: identity.  No device work. */
int b200_undistort_params(const double newK[9], const double R[9], double tauX, double tauY, double iR_out[9], double tilt_out[9]);

/* cv::undistort (undistort.dispatch.cpp:284-328), CV_16SC2 maps + cv::remap INTER_LINEAR BORDER_CONSTANT(0) on one stream.  The
 * reference works in stripes of max(1, 4096 / width) rows, each with a map built from a new camera matrix whose v0 is lowered by the
 * stripe's first row (:311-325) -- its entries are not those of one map over the whole image, and this entry reproduces the stripes
 * (one launch; every row block inverts its stripe's matrix).  newK = NULL: the camera matrix.  map1 (4 bytes per pixel) / map2
 * (2 bytes per pixel) are the caller's and can be re-used with b200_remap16s_dev for the following frames. */
int b200_undistort_dev(int src_type, const void* src, size_t src_step, int width, int height, void* dst, size_t dst_step,
                       const double K[9], const double dist12[12], const double newK[9], const double tilt[9],
                       void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream);

/* The stitching module's rotation warpers on the device.  b200_stitch_buildMaps_dev = the map loop of
 * RotationWarperBase<P>::buildMaps / PlaneWarper::buildMaps (stitching/include/opencv2/stitching/detail/warpers_inl.hpp:74-99,
 * stitching/src/warpers.cpp:192-243): kind 0 = PlaneProjector, 1 = SphericalProjector, 2 = CylindricalProjector (mapBackward,
 * warpers_inl.hpp:235-248, :263-281, :296-314).  k_rinv / t / scale are the projector's members after setCameraParams
 * (warpers.cpp:128-161; t = NULL: zeros), (tl_x, tl_y, width, height) the result ROI the warper's detectResultRoi found -- both are
 * host work of the warper object and stay with the caller.  CV_32FC1 maps.  Plane: bit-identical; spherical / cylindrical: a few ulp
 * (sinf / cosf of the CUDA math library against the host libm's).  b200_stitch_warp_dev = RotationWarper::warp: the maps, then
 * cv::remap(src, dst, xmap, ymap, interpolation, border) with border value 0 on the same stream; xmap / ymap are the caller's storage
 * (4 bytes per destination pixel each) and can serve b200_remap32f_dev for the following frames of the same camera. */
int b200_stitch_buildMaps_dev(int kind, const float k_rinv[9], const float t[3], float scale, int tl_x, int tl_y, int width, int height,
                              float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, void* stream);
int b200_stitch_warp_dev(int kind, const float k_rinv[9], const float t[3], float scale, int tl_x, int tl_y,
                         int src_type, const void* src, size_t src_step, int src_width, int src_height,
                         void* dst, size_t dst_step, int dst_width, int dst_height,
                         float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, int interpolation, int border, void* stream);

/* cv::warpPolar without WARP_INVERSE_MAP (imgproc/src/imgwarp.cpp:3163-3224; cv::linearPolar / cv::logPolar call it, :3284-3296).
 * b200_warpPolar_buildMaps_dev = its CV_32FC1 maps for a width x height polar image (column = radius, row = angle): flags & 256
 * (WARP_POLAR_LOG) selects the semilog radius table.  The reference evaluates exp / log per column and cos / sin per row only; the
 * library does the same on the host (same libm) and keeps the tables of the last geometry per host thread, the device does the
 * multiply-add per entry in the reference's rounding order: bit-identical maps.  b200_warpPolar_dev = the maps, then cv::remap with
 * interpolation flags & 7 and BORDER_CONSTANT (flags & 8, WARP_FILL_OUTLIERS; value 0) or BORDER_TRANSPARENT (dst keeps its contents
 * where the map leaves the source) on the same stream.  xmap / ymap are the caller's (4 bytes per destination pixel each) and can
 * serve b200_remap32f_dev for the following frames.  WARP_INVERSE_MAP (flags & 16) returns NOT_IMPLEMENTED: that direction runs
 * cartToPolar and cv::log on the host in the reference and is left to it. */
int b200_warpPolar_buildMaps_dev(int width, int height, float center_x, float center_y, double max_radius, int flags,
                                 float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, void* stream);
int b200_warpPolar_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                       void* dst, size_t dst_step, int dst_width, int dst_height, float center_x, float center_y, double max_radius, int flags,
                       float* xmap, size_t xmap_step, float* ymap, size_t ymap_step, void* stream);

/* cv::fisheye::undistortImage (calib3d/src/fisheye.cpp:636-646) = the map above (CV_16SC2, R = I, P = Knew) + cv::remap INTER_LINEAR
 * BORDER_CONSTANT, on one stream.  iKnew = Knew(:, 0:3)^-1.  map1 (4 bytes per destination pixel, rows of map1_step bytes) and map2
 * (2 bytes per pixel) are the caller's: keep them and call b200_remap16s_dev alone for the following frames of a stream. */
int b200_fisheye_undistortImage_dev(int src_type, const void* src, size_t src_step, int src_width, int src_height,
                                    void* dst, size_t dst_step, int dst_width, int dst_height,
                                    const double K[9], const double D[4], const double iKnew[9],
                                    void* map1, size_t map1_step, void* map2, size_t map2_step, void* stream);

/* Synthetic frame generator: byte (frame,y,x,c) = hash(seed, frame, y, x, c); 8U only. */
int b200_synth_frames_dev(void* dst, size_t frame_stride, size_t step, int width, int height, int cn,
                          int first_frame, int n_frames, unsigned int seed, void* stream);

/* ----------------------------------------------------------- multi-GPU frame scheduler ----------
 * Shards frame range [0, n_frames) into contiguous blocks [g*n/G, (g+1)*n/G) over the visible devices
 * listed in `devices` (n_devices entries), one host thread + stream per device, no collective.
 * Frames are synthetic (b200_synth_frames_dev, generated on the owning device, outside the timing);
 * each device runs the resize->warpAffine pipeline over its block in chunks of `chunk` frames.
 * Writes per-device elapsed milliseconds (CUDA events) to elapsed_ms[n_devices] and a per-device
 * 64-bit checksum of all outputs to checksums[n_devices].  A device that fails (lost, out of memory, bad ordinal) does not fail the
 * job: its block is run again on a device whose own block went through and b200_last_error() names it; only when no device is left
 * the first failure is returned (the same holds for b200_sched_run_pipeline_host). */
int b200_sched_run_pipeline(const int* devices, int n_devices, int n_frames, int chunk,
                            int src_width, int src_height, int cn, int mid_width, int mid_height,
                            int dst_width, int dst_height, const double M[6], unsigned int seed,
                            float* elapsed_ms, unsigned long long* checksums);
/* The same sharding for frames in HOST memory: device g streams its block through b200_hal_resize_warpAffine_batch on its
 * own host thread (one PCIe link per GPU, no collective).  elapsed_ms[n_devices] (optional) = wall time per device.
 * Returns when every destination frame is complete in host memory. */
int b200_sched_run_pipeline_host(const int* devices, int n_devices, int src_type,
                                 const void* src, size_t src_frame_stride, size_t src_step, int src_width, int src_height,
                                 int mid_width, int mid_height,
                                 void* dst, size_t dst_frame_stride, size_t dst_step, int dst_width, int dst_height, int n_frames,
                                 int resize_interpolation, const double M[6], int warp_interpolation, int borderType,
                                 const double borderValue[4], float* elapsed_ms);
/* frame block of rank `rank` out of `world`: writes first and count (contiguous-block rule above) */
void b200_sched_partition(int n_frames, int world, int rank, int* first, int* count);

/* ------------------------------------------------------------------------------ utilities ------ */
const char* b200_last_error(void);          /* thread-local text of the last failure */
int b200_device_count(void);
const char* b200_version(void);
/* Per-kernel timing for measurement harnesses: between begin and end, every kernel this library launches from
 * the calling thread is bracketed by CUDA events recorded on its launch stream.  b200_profile_end waits for them
 * and returns, per kind (0 = resize kernels, 1 = warp/remap kernels, 2 = other), the summed milliseconds and
 * the number of launches.  No counterpart in the reference. */
void b200_profile_begin(void);
int b200_profile_end(double* ms_by_kind, unsigned long long* launches_by_kind, int n_kinds);
/* diagnostics: non-zero routes every call to the generic kernels (skips the TMA / vectorised hot-cell
 * kernels); both families produce identical bytes, tests compare them.  Returns the old value. */
int b200_set_force_generic(int on);
/* number of kernel launches issued by this library in this process (all threads) */
unsigned long long b200_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_GEOM_H */
