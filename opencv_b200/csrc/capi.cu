// extern "C" surface of libb200geom.so (declared in include/b200_geom.h): argument checking,
// reference-faithful dispatch rules, host<->device staging for the HAL hooks, batched entries,
// synthetic frames and the multi-GPU frame scheduler.  No CPU fallback lives here: unsupported
// cases return B200_NOT_IMPLEMENTED so the caller (OpenCV's CALL_HAL) uses its own path.
#include "../../include/b200_geom.h"
#include "warp_kernels.cuh"
#include "interp_tables.cuh"
#include "fast_kernels.cuh"
#include "resize_kernels.cuh"

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <chrono>
#include <thread>
#include <vector>
#include <new>
#include <algorithm>
#include <cfloat>
#include <climits>
#include "trace.cuh"

using namespace b200;

namespace {

// (B200_TRACE, trace.cuh: an NVTX range around every compute entry of the C ABI -- SURVEY section 5: the reference wraps its
// entries in CV_INSTRUMENT_REGION / OPENCV_TRACE; this is what nsys / ncu show instead.  Free when no tool is attached.)

thread_local std::string g_err;
std::atomic<unsigned long long> g_launches{0};
std::atomic<long long> g_min_pixels{-1};
std::atomic<int> g_force_generic{0};
std::atomic<int> g_policy{-1};                   // which cells the HAL hooks take: see hal_policy_all()

// ---- optional per-launch CUDA-event timing (b200_profile_begin/end): events are recorded on the launch
// stream directly around each kernel of this library, so bench.py can report each kernel's mean duration
// from inside its timed region.
enum { PK_RESIZE = 0, PK_WARP = 1, PK_OTHER = 2, PK_KINDS = 3 };
struct ProfRec { int kind; cudaEvent_t e0, e1; };
struct Profiler {
    bool on = false;
    std::vector<ProfRec> recs;
    std::vector<cudaEvent_t> pool;
    cudaEvent_t get() {
        if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
        cudaEvent_t e; cudaEventCreate(&e); return e;
    }
};
thread_local Profiler g_prof;
struct ProfScope {
    ProfRec r; bool active; cudaStream_t s;
    ProfScope(int kind, cudaStream_t stream) : active(g_prof.on), s(stream) {
        if (active) { r.kind = kind; r.e0 = g_prof.get(); r.e1 = g_prof.get(); cudaEventRecord(r.e0, s); }
    }
    ~ProfScope() { if (active) { cudaEventRecord(r.e1, s); g_prof.recs.push_back(r); } }
    void cancel() { if (active) { g_prof.pool.push_back(r.e0); g_prof.pool.push_back(r.e1); active = false; } }   // nothing was launched
};

int fail(const char* what, cudaError_t e)
{
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return B200_ERROR;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(#call, e_); } while (0)

// HAL hooks only.  OpenCV's CALL_HAL (imgproc/src/hal_replacement.hpp:1608-1614) turns every return code other than 0 / 1
// into CV_Error(StsInternal), so a hook may only report an error once it has written to the caller's dst.  Everything that
// goes wrong before that -- no driver, no device, no stream, no memory, a failed upload or launch -- declines instead
// (the caller's own CPU path then answers, as hal/ipp/src/warp_ipp.cpp:81-115 does for its failures), says so once on
// stderr, and leaves the CUDA error state clean.
std::atomic<int> g_hal_unusable{0};        // 1: no usable CUDA device in this process -- later calls decline without asking again
std::atomic<int> g_hal_logged{0};
int hal_decline(const char* what, cudaError_t e)
{
    g_err = std::string(what) + ": " + cudaGetErrorString(e) + " (declined; the caller's CPU path runs)";
    cudaGetLastError();                    // do not leave a non-sticky error behind for the caller's own CUDA code
    if (!g_hal_logged.exchange(1)) fprintf(stderr, "b200 HAL: %s -- falling back to the CPU path\n", g_err.c_str());
    return B200_NOT_IMPLEMENTED;
}
#define HCK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return hal_decline(#call, e_); } while (0)

// x86 cvRound(double) (core/fast_math.hpp:199-211) for host-side saturate_cast of the border value
int host_rne(double v)
{
    double r = std::nearbyint(v);
    if (!(r >= -2147483648.0 && r <= 2147483647.0)) return INT_MIN;
    return (int)r;
}

int elem_size1(int depth) { return depth <= D_8S ? 1 : depth <= D_16S ? 2 : depth <= D_32F ? 4 : 8; }

bool depth_ok_linear(int depth) { return depth == D_8U || depth == D_16U || depth == D_16S || depth == D_32F; }

// cval[k] = saturate_cast<T>(borderValue[k & 3])  (imgwarp.cpp:278-279, :630-631)
void fill_cval(uint8_t* cval, int depth, int cn, const double bv[4])
{
    memset(cval, 0, 32);
    for (int k = 0; k < cn; k++) {
        double v = bv ? bv[k & 3] : 0.0;
        int i = host_rne(v);
        switch (depth) {
        case D_8U:  cval[k] = (uint8_t)std::max(0, std::min(255, i)); break;
        case D_8S:  ((int8_t*)cval)[k] = (int8_t)std::max(-128, std::min(127, i)); break;
        case D_16U: ((uint16_t*)cval)[k] = (uint16_t)std::max(0, std::min(65535, i)); break;
        case D_16S: ((int16_t*)cval)[k] = (int16_t)std::max(-32768, std::min(32767, i)); break;
        case D_32S: ((int32_t*)cval)[k] = i; break;
        case D_32F: ((float*)cval)[k] = (float)v; break;
        case D_64F: ((double*)cval)[k] = v; break;
        default: break;
        }
    }
}

bool aligned_to(const void* p, size_t step, int a) { return ((uintptr_t)p % a) == 0 && (step % a) == 0; }

// ---------------------------------------------------------------- warp-family common path ----
struct WarpCall {
    int src_type; const void* src; size_t src_step; int sw, sh;
    void* dst; size_t dst_step; int dw, dh;
    int coord; const double* M;
    const void* map1; size_t map1_step; const void* map2; size_t map2_step;
    int interpolation, border; const double* bv;
    size_t src_frame_stride = 0, dst_frame_stride = 0; int n_frames = 1;
    int keep_in_l2 = 0;
};

// validates + fills the kernel argument block; returns B200_OK / B200_NOT_IMPLEMENTED
int prepare_warp(const WarpCall& c, WarpArgs& a, int& depth, int& cn, bool resolve_tables = true)
{
    depth = B200_DEPTH(c.src_type); cn = B200_CN(c.src_type);
    int interp = c.interpolation;
    a.relative = 0;
    if (c.coord >= COORD_MAP32F) {                      // remap hooks get the raw flags (imgwarp.cpp:1608-1626)
        a.relative = (interp & WARP_RELATIVE_MAP) != 0;
        interp &= ~WARP_RELATIVE_MAP;
    }
    if (interp == I_AREA) interp = I_LINEAR;            // imgwarp.cpp:1625-1626, :2395, :2855
    if (interp != I_NEAREST && interp != I_LINEAR && interp != I_CUBIC && interp != I_LANCZOS4) return B200_NOT_IMPLEMENTED;
    if (cn < 1 || cn > 4) return B200_NOT_IMPLEMENTED;
    // 64F: float table weights, double arithmetic (imgwarp.cpp:1546, :1564, :1588); not with BORDER_TRANSPARENT
    if (interp != I_NEAREST && !depth_ok_linear(depth) && !(depth == D_64F && c.border != B_TRANSPARENT)) return B200_NOT_IMPLEMENTED;
    if (depth > D_64F) return B200_NOT_IMPLEMENTED;
    if (c.border < B_CONSTANT || c.border > B_TRANSPARENT) return B200_NOT_IMPLEMENTED;
    if (c.sw <= 0 || c.sh <= 0 || c.dw <= 0 || c.dh <= 0) return B200_NOT_IMPLEMENTED;
    if (c.sw >= 32767 || c.sh >= 32767 || c.dw >= 32767 || c.dh >= 32767) return B200_NOT_IMPLEMENTED;  // imgwarp.cpp:1603
    // LINEAR from a bare CV_16SC2 map is undefined in the reference (reads the map as floats): decline
    if (c.coord == COORD_MAP16S && interp != I_NEAREST && !c.map2) return B200_NOT_IMPLEMENTED;
    const int e1 = elem_size1(depth);
    if (!aligned_to(c.src, c.src_step, e1) || !aligned_to(c.dst, c.dst_step, e1)) return B200_NOT_IMPLEMENTED;
    if (c.coord == COORD_MAP32F && (!aligned_to(c.map1, c.map1_step, 4) || !aligned_to(c.map2, c.map2_step, 4))) return B200_NOT_IMPLEMENTED;
    if (c.coord == COORD_MAP32FC2 && !aligned_to(c.map1, c.map1_step, 8)) return B200_NOT_IMPLEMENTED;
    if (c.coord == COORD_MAP16S && (!aligned_to(c.map1, c.map1_step, 4) || (c.map2 && !aligned_to(c.map2, c.map2_step, 2)))) return B200_NOT_IMPLEMENTED;

    if (c.src_step > 0x7fffffffu || c.dst_step > 0x7fffffffu || c.map1_step > 0x7fffffffu || c.map2_step > 0x7fffffffu) return B200_NOT_IMPLEMENTED;
    a.src = Image{(const uint8_t*)c.src, (int)c.src_step, c.sw, c.sh};
    a.dst = ImageOut{(uint8_t*)c.dst, (int)c.dst_step, c.dw, c.dh};
    a.src_frame_stride = (long long)c.src_frame_stride; a.dst_frame_stride = (long long)c.dst_frame_stride;
    for (int i = 0; i < 9; i++) a.M[i] = 0;
    if (c.coord == COORD_AFFINE) memcpy(a.M, c.M, 6 * sizeof(double));
    else if (c.coord == COORD_PERSP) memcpy(a.M, c.M, 9 * sizeof(double));
    a.border = c.border; a.linear = interp != I_NEAREST; a.interp = interp;
    a.tab1d = nullptr; a.tab2d_i = nullptr;
    if (resolve_tables && (interp == I_CUBIC || interp == I_LANCZOS4)) {
        cudaError_t e = interp_tables(interp, &a.tab1d, &a.tab2d_i);
        if (e != cudaSuccess) return fail("interpolation tables", e);
    }
    a.force_generic = g_force_generic.load();
    a.keep_in_l2 = c.keep_in_l2;
    // word-granular access: 4-byte pixels (src and dst words) or 8UC3 (aligned source words, byte stores)
    a.vec4 = ((e1 * cn == 4) && aligned_to(c.src, c.src_step, 4) && aligned_to(c.dst, c.dst_step, 4) &&
              (c.src_frame_stride % 4 == 0) && (c.dst_frame_stride % 4 == 0)) ||
             (depth == D_8U && cn == 3 && aligned_to(c.src, c.src_step, 4) && (c.src_frame_stride % 4 == 0));
    // warpPerspective block width: bh0 = min(16,H); bw0 = min(1024/bh0, W)  (imgwarp.cpp:2695-2697)
    int bh0 = std::min(16, c.dh);
    a.bw0 = std::min(1024 / bh0, c.dw);
    a.bw0_shift = -1;
    for (int sft = 0; sft < 16; sft++) if ((1 << sft) == a.bw0) a.bw0_shift = sft;
    a.map1 = (const uint8_t*)c.map1; a.map1_step = (int)c.map1_step;
    a.map2 = (const uint8_t*)c.map2; a.map2_step = (int)c.map2_step;
    fill_cval(a.cval, depth, cn, c.bv);
    return B200_OK;
}

int run_warp_dev(const WarpCall& c, cudaStream_t stream)
{
    WarpArgs a; int depth, cn;
    int rc = prepare_warp(c, a, depth, cn);
    if (rc != B200_OK) return rc;
    ProfScope prof(PK_WARP, stream);
    cudaError_t e = (cudaError_t)launch_warp(a, depth, cn, c.coord, c.n_frames, stream);
    if (e != cudaSuccess) return fail("warp kernel launch", e);
    g_launches++;
    return B200_OK;
}

// ---------------------------------------------------------------- resize common path ---------
struct ResizeCall {
    int src_type; const void* src; size_t src_step; int sw, sh;
    void* dst; size_t dst_step; int dw, dh;
    double inv_scale_x, inv_scale_y; int interpolation;
    size_t src_frame_stride = 0, dst_frame_stride = 0; int n_frames = 1;
    int keep_in_l2 = 0;
    const float* ltab_x = nullptr; const float* ltab_y = nullptr;   // INTER_LANCZOS4: host-built coefficient tables on the device
};

int prepare_resize(const ResizeCall& c, ResizeArgs& a, int& mode, int& depth, int& cn)
{
    depth = B200_DEPTH(c.src_type); cn = B200_CN(c.src_type);
    if (cn < 1 || cn > 4 || depth > D_64F) return B200_NOT_IMPLEMENTED;
    if (c.sw <= 0 || c.sh <= 0 || c.dw <= 0 || c.dh <= 0) return B200_NOT_IMPLEMENTED;
    a.inv_scale_x = c.inv_scale_x; a.inv_scale_y = c.inv_scale_y;
    if (a.inv_scale_x < DBL_EPSILON || a.inv_scale_y < DBL_EPSILON) {        // resize.cpp:3642-3646
        a.inv_scale_x = (double)c.dw / c.sw; a.inv_scale_y = (double)c.dh / c.sh;
    }
    // the reference re-derives dsize from the scales (resize.cpp:3652-3653); decline if it disagrees
    if (host_rne(c.sw * a.inv_scale_x) != c.dw || host_rne(c.sh * a.inv_scale_y) != c.dh) return B200_NOT_IMPLEMENTED;
    a.force_generic = g_force_generic.load();
    a.keep_in_l2 = c.keep_in_l2;
    a.cn = cn;
    a.src.w = c.sw; a.src.h = c.sh; a.dst.w = c.dw; a.dst.h = c.dh;       // mode selection reads the sizes
    mode = resize_select_mode(a, c.interpolation);
    if (mode < 0) return B200_NOT_IMPLEMENTED;
    if (mode == RS_LINEAR_EXACT && !(depth == D_8U || depth == D_16U || depth == D_8S || depth == D_16S || depth == D_32S)) return B200_NOT_IMPLEMENTED;
    if (mode != RS_NEAREST && mode != RS_NEAREST_EXACT && !depth_ok_linear(depth) && !((depth == D_8S || depth == D_32S) && mode == RS_LINEAR_EXACT) && !(depth == D_64F && (mode == RS_LINEAR || mode == RS_CUBIC || mode == RS_LANCZOS4 || mode == RS_AREA || mode == RS_AREA_FAST))) return B200_NOT_IMPLEMENTED;
    const int e1 = elem_size1(depth);
    if (!aligned_to(c.src, c.src_step, e1) || !aligned_to(c.dst, c.dst_step, e1)) return B200_NOT_IMPLEMENTED;
    if (mode == RS_AREA && (size_t)(128 * a.kmax_x + 16 * a.kmax_y) * 4 > 40 * 1024) return B200_NOT_IMPLEMENTED;   // shared-memory tables of resize_area_kernel
    if (c.src_step > 0x7fffffffu || c.dst_step > 0x7fffffffu) return B200_NOT_IMPLEMENTED;
    a.src = Image{(const uint8_t*)c.src, (int)c.src_step, c.sw, c.sh};
    a.dst = ImageOut{(uint8_t*)c.dst, (int)c.dst_step, c.dw, c.dh};
    a.src_frame_stride = (long long)c.src_frame_stride; a.dst_frame_stride = (long long)c.dst_frame_stride;
    return B200_OK;
}

int lanczos4_device_tables(const ResizeArgs& pre, int sw, int sh, int dw, int dh, cudaStream_t s, const float** tx, const float** ty);

int run_resize_dev(const ResizeCall& c, cudaStream_t stream)
{
    ResizeArgs a; int mode, depth, cn;
    int rc = prepare_resize(c, a, mode, depth, cn);
    if (rc != B200_OK) return rc;
    if (mode == RS_LANCZOS4) {
        if (c.ltab_x && c.ltab_y) { a.ltab_x = c.ltab_x; a.ltab_y = c.ltab_y; }     // the hook uploaded them with the image
        else {
            if (c.n_frames > 1) return B200_NOT_IMPLEMENTED;
            rc = lanczos4_device_tables(a, c.sw, c.sh, c.dw, c.dh, stream, &a.ltab_x, &a.ltab_y);
            if (rc != B200_OK) return rc;
        }
    }
    ProfScope prof(PK_RESIZE, stream);
    cudaError_t e = (cudaError_t)launch_resize(a, mode, depth, cn, c.n_frames, stream);
    if (e != cudaSuccess) return fail("resize kernel launch", e);
    g_launches++;
    return B200_OK;
}

// ---------------------------------------------------------------- HAL (host pointer) staging --
// Per host thread: one stream and grow-only device buffers (hooks are called concurrently from any
// thread, incl. OpenCV pool workers).  Copies use cudaMemcpy2DAsync straight from/to the
// caller's memory (honours `step`; DMA when the caller's buffer is pinned, staged by the driver otherwise).
struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = (n + (1u << 20)) & ~((size_t)(1u << 20) - 1);
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    ~DevBuf() { /* process teardown: leave to the driver */ }
};
struct HalCtx {
    int device = -1;
    cudaStream_t stream = nullptr;
    DevBuf src, dst, map1, map2, grp_src, grp_dst;      // grp_*: one group of <= 4 channels of an image with more (hal_warp)
    // a host thread that ends gives its staging buffers and stream back (errors ignored: at process exit the CUDA
    // context may already be gone, and then there is nothing left to free)
    ~HalCtx() {
        if (!stream) return;
        int cur = -1;
        if (cudaGetDevice(&cur) != cudaSuccess) return;
        if (cur != device && cudaSetDevice(device) != cudaSuccess) return;
        for (DevBuf* b : {&src, &dst, &map1, &map2, &grp_src, &grp_dst}) if (b->p) cudaFree(b->p);
        cudaStreamDestroy(stream);
        if (cur != device) cudaSetDevice(cur);
    }
};
thread_local HalCtx g_ctx;

int hal_ctx_init()
{
    if (g_ctx.stream) {
        int cur = -1;
        if (cudaGetDevice(&cur) == cudaSuccess && cur == g_ctx.device) return B200_OK;
        if (getenv("OPENCV_B200_DEVICES") && cudaSetDevice(g_ctx.device) == cudaSuccess) return B200_OK;   // this thread's assigned device
        // the thread switched devices since its last call: release the old device's staging and start over
        g_ctx.~HalCtx();
        new (&g_ctx) HalCtx();
    }
    if (g_hal_unusable.load()) { g_err = "no CUDA device (declined)"; return B200_NOT_IMPLEMENTED; }
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_hal_unusable.store(1);
        return hal_decline("no CUDA device", e != cudaSuccess ? e : cudaErrorNoDevice);
    }
    // OPENCV_B200_DEVICES="0,2,5": the hooks spread the host threads that call them (OpenCV pool workers, application
    // threads) over these devices, thread k of first use -> entry k mod count; unset: every thread uses its current device
    static const std::vector<int> devices = [n]() {
        std::vector<int> v;
        if (const char* s = getenv("OPENCV_B200_DEVICES")) {
            for (const char* p = s; *p; ) {
                char* end = nullptr;
                const long d = strtol(p, &end, 10);
                if (end == p) break;
                if (d >= 0 && d < n) v.push_back((int)d);
                p = *end ? end + 1 : end;
            }
        }
        return v;
    }();
    static std::atomic<unsigned> next_thread{0};
    if (!devices.empty()) HCK(cudaSetDevice(devices[next_thread.fetch_add(1) % devices.size()]));
    HCK(cudaGetDevice(&g_ctx.device));
    HCK(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
    return B200_OK;
}

// INTER_LANCZOS4: per-column / per-row coefficients exactly as resize.cpp:3899-3996 builds them (fp64 sin/cos of the host libm)
void build_lanczos4_tables(const ResizeArgs& pre, int dw, int dh, std::vector<float>& ltab)
{
    ltab.resize((size_t)(dw + dh) * 8);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * pre.scale_x - 0.5);
        fx -= (float)(int)std::floor(fx);
        lanczos4_coeffs_host(fx, &ltab[(size_t)dx * 8]);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * pre.scale_y - 0.5);
        fy -= (float)(int)std::floor(fy);
        lanczos4_coeffs_host(fy, &ltab[(size_t)(dw + dy) * 8]);
    }
}

// Device-resident INTER_LANCZOS4 calls need the same host-built tables on the device: kept per host thread for the last
// geometry (a video loop resizes every frame the same way), rebuilt and re-uploaded only when the geometry or device changes.
// The upload is ordered on the caller's stream; the table memory is this library's (one small cudaMalloc per geometry change).
struct LanczosCache {
    int device = -1, sw = 0, sh = 0, dw = 0, dh = 0; double sx = 0, sy = 0;
    DevBuf buf;
};
thread_local LanczosCache g_lcache;

int lanczos4_device_tables(const ResizeArgs& pre, int sw, int sh, int dw, int dh, cudaStream_t s, const float** tx, const float** ty)
{
    int dev = -1;
    CK(cudaGetDevice(&dev));
    LanczosCache& c = g_lcache;
    if (!(c.device == dev && c.sw == sw && c.sh == sh && c.dw == dw && c.dh == dh && c.sx == pre.scale_x && c.sy == pre.scale_y && c.buf.p)) {
        std::vector<float> ltab;
        build_lanczos4_tables(pre, dw, dh, ltab);
        if (c.device != dev) c.buf = DevBuf();                 // another device's allocation is not ours to reuse here
        CK(c.buf.reserve(ltab.size() * sizeof(float)));
        // pageable source: the runtime stages it before returning, so `ltab` may go out of scope
        CK(cudaMemcpyAsync(c.buf.p, ltab.data(), ltab.size() * sizeof(float), cudaMemcpyHostToDevice, s));
        c.device = dev; c.sw = sw; c.sh = sh; c.dw = dw; c.dh = dh; c.sx = pre.scale_x; c.sy = pre.scale_y;
    }
    *tx = (const float*)c.buf.p; *ty = *tx + (size_t)dw * 8;
    return B200_OK;
}

size_t pitch_for(size_t row_bytes) { return (row_bytes + 255) & ~(size_t)255; }

// ---- host frame streams: upload, kernels and download of consecutive chunks overlap on three streams ----
struct StreamCtx {
    static constexpr int NS = 3;                       // staging slots in flight
    int device = -1;
    cudaStream_t up = nullptr, run = nullptr, down = nullptr;
    cudaEvent_t up_done[NS] = {}, run_done[NS] = {}, down_done[NS] = {};
    DevBuf src[NS], dst[NS], mid, maps;
    void release() {
        if (!up) return;
        int cur = -1;
        if (cudaGetDevice(&cur) != cudaSuccess) return;
        if (cur != device && cudaSetDevice(device) != cudaSuccess) return;
        for (int i = 0; i < NS; i++) {
            if (src[i].p) cudaFree(src[i].p);
            if (dst[i].p) cudaFree(dst[i].p);
            src[i] = DevBuf(); dst[i] = DevBuf();
            cudaEventDestroy(up_done[i]); cudaEventDestroy(run_done[i]); cudaEventDestroy(down_done[i]);
        }
        if (mid.p) cudaFree(mid.p);
        if (maps.p) cudaFree(maps.p);
        mid = DevBuf(); maps = DevBuf();
        cudaStreamDestroy(up); cudaStreamDestroy(run); cudaStreamDestroy(down);
        up = run = down = nullptr;
        if (cur != device) cudaSetDevice(cur);
    }
    ~StreamCtx() { release(); }
};
thread_local StreamCtx g_sctx;

int stream_ctx_init()
{
    int cur = -1;
    CK(cudaGetDevice(&cur));
    if (g_sctx.up && cur == g_sctx.device) return B200_OK;
    g_sctx.release();
    g_sctx.device = cur;
    CK(cudaStreamCreateWithFlags(&g_sctx.up, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&g_sctx.run, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&g_sctx.down, cudaStreamNonBlocking));
    for (int i = 0; i < StreamCtx::NS; i++) {
        CK(cudaEventCreateWithFlags(&g_sctx.up_done[i], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&g_sctx.run_done[i], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&g_sctx.down_done[i], cudaEventDisableTiming));
    }
    return B200_OK;
}

// how a batch of host frames is laid out on the device: either the caller's own layout (one contiguous copy per
// chunk; see upload()) or rows padded to 256 bytes (one 2-D copy per frame)
struct FrameLayout { size_t step, frame_stride; bool block; };
FrameLayout frame_layout(size_t host_step, size_t host_frame_stride, size_t row_bytes, int rows, bool is_dst)
{
    const size_t tight = host_step * (size_t)rows;
    const bool rows_ok = is_dst ? (host_step == row_bytes && row_bytes % 16 == 0)
                                : (host_step % 16 == 0 && host_step >= row_bytes && host_step - row_bytes <= row_bytes / 4);
    const bool frames_ok = is_dst ? host_frame_stride == tight
                                  : (host_frame_stride % 16 == 0 && host_frame_stride >= tight && host_frame_stride - tight <= tight / 4);
    if (rows_ok && frames_ok) return {host_step, host_frame_stride, true};
    const size_t p = pitch_for(row_bytes);
    return {p, p * (size_t)rows, false};
}

long long min_pixels()
{
    long long v = g_min_pixels.load();
    if (v < 0) {
        const char* s = getenv("OPENCV_B200_MIN_PIXELS");
        v = s ? atoll(s) : (1 << 16);
        g_min_pixels.store(v);
    }
    return v;
}

// Which cells the per-image hooks take.  0 (default, env OPENCV_B200_POLICY unset or "profitable"): only cells in which the upload +
// kernel + download of one image beat the reference's CPU code on a many-core host (profiles/r2_perf_view.md: the reference's own perf
// grids on pageable memory, 16 host threads) -- every INTER_NEAREST cell, cv::resize with INTER_LINEAR / INTER_LINEAR_EXACT /
// the 2x2 INTER_AREA box and INTER_LINEAR remaps of more than 2^20 pixels (their maps cross PCIe with every call) cost the host
// threads less than the PCIe copies alone and are left to them.  1 (env "all", or b200_hal_set_policy(1)): every supported cell.
bool hal_policy_all()
{
    int v = g_policy.load();
    if (v < 0) {
        const char* s = getenv("OPENCV_B200_POLICY");
        v = (s && (s[0] == 'a' || s[0] == '1')) ? 1 : 0;
        g_policy.store(v);
    }
    return v != 0;
}

// Measured on B200 (PCIe Gen5): a 1080-row cudaMemcpy2DAsync with different host/device pitches runs at ~15 GB/s,
// one contiguous cudaMemcpyAsync at ~54 GB/s.  So when the caller's rows are (nearly) contiguous and 16-byte
// aligned the device copy keeps the caller's step and moves as ONE block; the bytes between rows of a ROI are
// only ever read (they belong to the parent image), never written back.
int upload(DevBuf& b, const void* host, size_t host_step, size_t row_bytes, int rows, size_t& pitch, cudaStream_t s)
{
    if ((host_step % 16) == 0 && host_step >= row_bytes && host_step - row_bytes <= row_bytes / 4) {
        pitch = host_step;
        const size_t bytes = host_step * (size_t)(rows - 1) + row_bytes;
        HCK(b.reserve(pitch * (size_t)rows));
        HCK(cudaMemcpyAsync(b.p, host, bytes, cudaMemcpyHostToDevice, s));
        return B200_OK;
    }
    pitch = pitch_for(row_bytes);
    HCK(b.reserve(pitch * (size_t)rows));
    HCK(cudaMemcpy2DAsync(b.p, pitch, host, host_step, row_bytes, rows, cudaMemcpyHostToDevice, s));
    return B200_OK;
}

// device pitch for the output and the copy back; contiguous destinations come back as one block
size_t out_pitch(size_t host_step, size_t row_bytes) { return (host_step == row_bytes && (row_bytes % 16) == 0) ? row_bytes : pitch_for(row_bytes); }
int download(void* host, size_t host_step, const void* dev, size_t pitch, size_t row_bytes, int rows, cudaStream_t s)
{
    if (pitch == host_step && host_step == row_bytes) CK(cudaMemcpyAsync(host, dev, row_bytes * (size_t)rows, cudaMemcpyDeviceToHost, s));
    else CK(cudaMemcpy2DAsync(host, host_step, dev, pitch, row_bytes, rows, cudaMemcpyDeviceToHost, s));
    return B200_OK;
}

bool hal_enabled()
{
    static int en = -1;
    if (en < 0) { const char* s = getenv("OPENCV_B200_HAL"); en = !(s && s[0] == '0'); }
    return en != 0;
}

// channels [c0, c0 + k) of an interleaved image <-> a packed k-channel image (elements of esz1 bytes): hal_warp runs images with more
// than 4 channels as groups of 4 -- remap's arithmetic does not depend on the channel count, and the border value of channel c is
// borderValue[c & 3] (imgwarp.cpp:278-279, 630-631), which a group starting at a multiple of 4 sees unchanged
template <typename U, bool GATHER>
__global__ void __launch_bounds__(256) channel_group_kernel(uint8_t* full, long long full_step, int cn, int c0, int k,
                                                            uint8_t* grp, long long grp_step, int w, int h)
{
    const int x = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (x >= w) return;
    U* f = reinterpret_cast<U*>(full + (long long)y * full_step) + (long long)x * cn + c0;
    U* g = reinterpret_cast<U*>(grp + (long long)y * grp_step) + (long long)x * k;
    for (int i = 0; i < k; i++) { if (GATHER) g[i] = f[i]; else f[i] = g[i]; }
}
template <bool GATHER>
int channel_group(int esz1, void* full, size_t full_step, int cn, int c0, int k, void* grp, size_t grp_step, int w, int h, cudaStream_t s)
{
    dim3 grid((w + 255) / 256, h);
#define B200_CG(U) channel_group_kernel<U, GATHER><<<grid, 256, 0, s>>>((uint8_t*)full, (long long)full_step, cn, c0, k, (uint8_t*)grp, (long long)grp_step, w, h)
    if (esz1 == 1) B200_CG(uint8_t); else if (esz1 == 2) B200_CG(uint16_t); else if (esz1 == 4) B200_CG(uint32_t); else B200_CG(unsigned long long);
#undef B200_CG
    return cudaGetLastError() == cudaSuccess ? B200_OK : B200_ERROR;
}

int hal_warp(WarpCall c)
{
    if (!hal_enabled()) return B200_NOT_IMPLEMENTED;
    if ((long long)c.dw * c.dh < min_pixels()) return B200_NOT_IMPLEMENTED;
    if (!hal_policy_all()) {
        int ip = c.interpolation & 7;                        // INTER_MAX: flag bits (WARP_RELATIVE_MAP ...) sit above
        if (ip == I_AREA) ip = I_LINEAR;                     // as prepare_warp
        if (ip == I_NEAREST) return B200_NOT_IMPLEMENTED;
        if (c.coord >= COORD_MAP32F && ip == I_LINEAR && (long long)c.dw * c.dh > (1LL << 20)) return B200_NOT_IMPLEMENTED;
    }
    const int depth0 = B200_DEPTH(c.src_type), cn0 = B200_CN(c.src_type);
    {   // validate before touching the GPU so NOT_IMPLEMENTED never costs a copy
        WarpArgs tmp; int d, n;
        WarpCall v = c;
        if (cn0 > 4) {                                  // more than 4 channels: groups of 4 (below); the rules are those of a 4-channel image
            // (INTER_CUBIC / INTER_LANCZOS4 with more than 4 channels: the reference asserts, imgwarp.cpp:1648-1652, :2374 -- left to it)
            const int ip = c.interpolation & 7;
            if (c.border == B_TRANSPARENT || cn0 > 512 || ip == I_CUBIC || ip == I_LANCZOS4) return B200_NOT_IMPLEMENTED;
            v.src_type = depth0 + (3 << 3);
        }
        int rc = prepare_warp(v, tmp, d, n, false);     // argument rules only: no CUDA call yet
        if (rc != B200_OK) return rc;
    }
    int rc = hal_ctx_init();
    if (rc != B200_OK) return rc;
    cudaStream_t s = g_ctx.stream;
    const int esz = elem_size1(B200_DEPTH(c.src_type)) * B200_CN(c.src_type);
    const void* h_src = c.src; void* h_dst = c.dst; const size_t h_dst_step = c.dst_step;
    size_t sp, dp, m1p = 0, m2p = 0;
    if ((rc = upload(g_ctx.src, h_src, c.src_step, (size_t)c.sw * esz, c.sh, sp, s)) != B200_OK) return rc;
    if (c.coord >= COORD_MAP32F) {
        size_t m1row = (size_t)c.dw * (c.coord == COORD_MAP32FC2 ? 8 : 4);
        if ((rc = upload(g_ctx.map1, c.map1, c.map1_step, m1row, c.dh, m1p, s)) != B200_OK) return rc;
        c.map1 = g_ctx.map1.p; c.map1_step = m1p;
        if (c.map2) {
            size_t m2row = (size_t)c.dw * (c.coord == COORD_MAP32F ? 4 : 2);
            if ((rc = upload(g_ctx.map2, c.map2, c.map2_step, m2row, c.dh, m2p, s)) != B200_OK) return rc;
            c.map2 = g_ctx.map2.p; c.map2_step = m2p;
        }
    }
    dp = out_pitch(h_dst_step, (size_t)c.dw * esz);
    HCK(g_ctx.dst.reserve(dp * (size_t)c.dh));
    if (c.border == B_TRANSPARENT) {
        // pixels whose source lies outside keep the caller's destination contents: bring those to the device first
        const size_t drow = (size_t)c.dw * esz;
        if (dp == h_dst_step && h_dst_step == drow) HCK(cudaMemcpyAsync(g_ctx.dst.p, h_dst, drow * (size_t)c.dh, cudaMemcpyHostToDevice, s));
        else HCK(cudaMemcpy2DAsync(g_ctx.dst.p, dp, h_dst, h_dst_step, drow, c.dh, cudaMemcpyHostToDevice, s));
    }
    c.src = g_ctx.src.p; c.src_step = sp; c.dst = g_ctx.dst.p; c.dst_step = dp;
    if (cn0 <= 4) rc = run_warp_dev(c, s);
    else {
        const int e1 = elem_size1(depth0);
        const size_t gsp = pitch_for((size_t)c.sw * 4 * e1), gdp = pitch_for((size_t)c.dw * 4 * e1);
        HCK(g_ctx.grp_src.reserve(gsp * (size_t)c.sh));
        HCK(g_ctx.grp_dst.reserve(gdp * (size_t)c.dh));
        for (int c0 = 0; c0 < cn0 && rc == B200_OK; c0 += 4) {
            const int k = cn0 - c0 < 4 ? cn0 - c0 : 4;
            WarpCall g = c;
            g.src_type = depth0 + ((k - 1) << 3);
            g.src = g_ctx.grp_src.p; g.src_step = gsp; g.dst = g_ctx.grp_dst.p; g.dst_step = gdp;
            rc = channel_group<true>(e1, g_ctx.src.p, sp, cn0, c0, k, g_ctx.grp_src.p, gsp, c.sw, c.sh, s);
            if (rc == B200_OK) rc = run_warp_dev(g, s);
            if (rc == B200_OK) rc = channel_group<false>(e1, g_ctx.dst.p, dp, cn0, c0, k, g_ctx.grp_dst.p, gdp, c.dw, c.dh, s);
        }
    }
    if (rc == B200_ERROR) return hal_decline("warp kernel launch", cudaGetLastError());   // nothing has reached the caller's dst yet
    if (rc != B200_OK) return rc;
    // uploads and kernel are checked BEFORE the copy back starts: an asynchronous failure up to here still declines
    HCK(cudaStreamSynchronize(s));
    // from here on the caller's dst is being written: a failure is a real error (CALL_HAL throws)
    if ((rc = download(h_dst, h_dst_step, g_ctx.dst.p, dp, (size_t)c.dw * esz, c.dh, s)) != B200_OK) return rc;
    CK(cudaStreamSynchronize(s));
    return B200_OK;
}

WarpCall make_warp(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step,
                   int dw, int dh, int coord, const double* M, const void* m1, size_t m1s, const void* m2, size_t m2s,
                   int interp, int border, const double* bv)
{
    WarpCall c;
    c.src_type = src_type; c.src = src; c.src_step = src_step; c.sw = sw; c.sh = sh;
    c.dst = dst; c.dst_step = dst_step; c.dw = dw; c.dh = dh; c.coord = coord; c.M = M;
    c.map1 = m1; c.map1_step = m1s; c.map2 = m2; c.map2_step = m2s;
    c.interpolation = interp; c.border = border; c.bv = bv;
    return c;
}

// ---------------------------------------------------------------- cv::convertMaps -------------
// imgwarp.cpp:1682-1986, one thread per map entry.  Fixed point: ix = cvRound(x*32) (x86 rule), xy = sat16(ix >> 5),
// fraction index = (iy & 31)*32 + (ix & 31); back: x = xy + (f & 31)*(1/32) in fp32.
enum MapConv { MC_F1F1_TO_S2 = 0, MC_F2_TO_S2 = 1, MC_S2_TO_F1F1 = 2, MC_S2_TO_F2 = 3, MC_F1F1_TO_F2 = 4, MC_F2_TO_F1F1 = 5,
               MC_COPY_S2 = 6, MC_COPY_F1F1 = 7, MC_COPY_F2 = 8 };
struct MapConvArgs {
    const uint8_t* s1; int s1_step; const uint8_t* s2; int s2_step;
    uint8_t* d1; int d1_step; uint8_t* d2; int d2_step;
    int w, h, mode, nn;
};
__global__ void convert_maps_kernel(const MapConvArgs a)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= a.w) return;
    const uint8_t* r1 = a.s1 + (long long)y * a.s1_step;
    const uint8_t* r2 = a.s2 ? a.s2 + (long long)y * a.s2_step : nullptr;
    uint8_t* o1 = a.d1 + (long long)y * a.d1_step;
    uint8_t* o2 = a.d2 ? a.d2 + (long long)y * a.d2_step : nullptr;
    switch (a.mode) {
    case MC_F1F1_TO_S2: case MC_F2_TO_S2: {
        float fx, fy;
        if (a.mode == MC_F1F1_TO_S2) { fx = reinterpret_cast<const float*>(r1)[x]; fy = reinterpret_cast<const float*>(r2)[x]; }
        else { const float2 v = reinterpret_cast<const float2*>(r1)[x]; fx = v.x; fy = v.y; }
        short2 xy; 
        if (a.nn) { xy.x = (short)sat_s16(rne_f(fx)); xy.y = (short)sat_s16(rne_f(fy)); }       // :1794-1798 saturate_cast<short>(float)
        else {                                                                                  // :1836-1842, :1889-1895
            const int ix = rne_f(fmul(fx, 32.f)), iy = rne_f(fmul(fy, 32.f));
            xy.x = (short)sat_s16(ix >> 5); xy.y = (short)sat_s16(iy >> 5);
            reinterpret_cast<unsigned short*>(o2)[x] = (unsigned short)((iy & 31) * 32 + (ix & 31));
        }
        reinterpret_cast<short2*>(o1)[x] = xy;
        break; }
    case MC_S2_TO_F1F1: case MC_S2_TO_F2: {                                                    // :1945-1950, :1974-1979
        const short2 xy = reinterpret_cast<const short2*>(r1)[x];
        const int f = r2 ? (reinterpret_cast<const unsigned short*>(r2)[x] & 1023) : 0;
        const float sc = 1.f / 32;
        const float fx = fadd((float)xy.x, fmul((float)(f & 31), sc)), fy = fadd((float)xy.y, fmul((float)(f >> 5), sc));
        if (a.mode == MC_S2_TO_F1F1) { reinterpret_cast<float*>(o1)[x] = fx; reinterpret_cast<float*>(o2)[x] = fy; }
        else reinterpret_cast<float2*>(o1)[x] = make_float2(fx, fy);
        break; }
    case MC_F1F1_TO_F2:                                                                        // merge, :1727-1732
        reinterpret_cast<float2*>(o1)[x] = make_float2(reinterpret_cast<const float*>(r1)[x], reinterpret_cast<const float*>(r2)[x]);
        break;
    case MC_F2_TO_F1F1: {                                                                      // split, :1734-1739
        const float2 v = reinterpret_cast<const float2*>(r1)[x];
        reinterpret_cast<float*>(o1)[x] = v.x;
        if (o2) reinterpret_cast<float*>(o2)[x] = v.y;                                          // nninterpolation: the y plane is dropped
        break; }
    case MC_COPY_S2:                                                                           // :1718-1725
        reinterpret_cast<short2*>(o1)[x] = reinterpret_cast<const short2*>(r1)[x];
        if (o2 && r2) reinterpret_cast<unsigned short*>(o2)[x] = reinterpret_cast<const unsigned short*>(r2)[x];
        break;
    case MC_COPY_F1F1:
        reinterpret_cast<float*>(o1)[x] = reinterpret_cast<const float*>(r1)[x];
        if (o2 && r2) reinterpret_cast<float*>(o2)[x] = reinterpret_cast<const float*>(r2)[x];
        break;
    default:
        reinterpret_cast<float2*>(o1)[x] = reinterpret_cast<const float2*>(r1)[x];
        break;
    }
}

// ---------------------------------------------------------------- synthetic frames ------------
__device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
// 4 bytes per thread: word j of row y of frame f = hash(seed, f, y, j)
__global__ void synth_frames_kernel(uint8_t* dst, long long frame_stride, long long step, int row_bytes, int height,
                                    int first_frame, unsigned seed)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;     // 32-bit word index in the row
    const int y = blockIdx.y;
    const int f = blockIdx.z;
    if (j * 4 >= row_bytes) return;
    uint32_t h = mix32(seed ^ mix32((uint32_t)(first_frame + f) * 0x9E3779B9U ^ mix32((uint32_t)y * 0x85EBCA6BU ^ (uint32_t)j)));
    uint8_t* p = dst + (long long)f * frame_stride + (long long)y * step + (long long)j * 4;
    if (j * 4 + 4 <= row_bytes && ((uintptr_t)p & 3) == 0) *reinterpret_cast<uint32_t*>(p) = h;
    else for (int k = 0; k < 4 && j * 4 + k < row_bytes; k++) p[k] = (uint8_t)(h >> (8 * k));
}

// 64-bit checksum of a pitched batch (order-independent sum of hashed (position,value) words)
__global__ void checksum_kernel(const uint8_t* p, long long frame_stride, long long step, int row_bytes, int height,
                                int first_frame, unsigned long long* out)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y, f = blockIdx.z;
    unsigned long long v = 0;
    if (j < row_bytes) {
        uint32_t b = p[(long long)f * frame_stride + (long long)y * step + j];
        uint32_t h = mix32((uint32_t)(first_frame + f) * 0x9E3779B9U ^ mix32((uint32_t)y * 0x85EBCA6BU ^ (uint32_t)j));
        v = (unsigned long long)mix32(h ^ b) * (b + 1);
    }
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(out, v);
}

} // namespace

// shared with the entry points that live in other files (map_gen.cu)
namespace b200 {
void count_launch() { g_launches++; }
int report_cuda_error(const char* what, cudaError_t e) { return fail(what, e); }
}

// =================================================================================== C ABI ====
// pyr_kernels.cu
namespace b200 {
int pyrdown_check(int type, int sw, int sh, int dw, int dh, int border);
int pyrup_check(int type, int sw, int sh, int dw, int dh, int border);
int launch_pyrup(int type, const void* src, size_t src_step, long long sfs, int sw, int sh,
                 void* dst, size_t dst_step, long long dfs, int dw, int dh, int n_frames, int border, cudaStream_t s);
int launch_pyrdown(int type, const void* src, size_t src_step, long long sfs, int sw, int sh,
                   void* dst, size_t dst_step, long long dfs, int dw, int dh, int n_frames, int border, cudaStream_t s);
}

namespace {
// Frame blocks whose device failed (lost device, out of memory, a bad ordinal) are run again, one after the other, on the first device
// whose own block went through: frames are independent, so the job still completes, later.  b200_last_error() then names the devices
// that were skipped.  With no surviving device the first failure is returned.
template <class RunBlock>
static int sched_redispatch(const int* devices, int n_devices, std::vector<int>& rcs, const std::vector<std::string>& errs, RunBlock run_block)
{
    int healthy = -1, first_bad = -1;
    for (int g = 0; g < n_devices; g++) {
        if (rcs[g] == B200_OK && healthy < 0) healthy = g;
        if (rcs[g] != B200_OK && first_bad < 0) first_bad = g;
    }
    if (first_bad < 0) return B200_OK;
    if (healthy < 0) { g_err = "device " + std::to_string(devices[first_bad]) + ": " + errs[first_bad]; return rcs[first_bad]; }
    std::string note;
    for (int g = 0; g < n_devices; g++) {
        if (rcs[g] == B200_OK) continue;
        cudaGetLastError();
        const int rc = run_block(g, devices[healthy]);
        if (rc != B200_OK) { g_err = "device " + std::to_string(devices[g]) + ": " + errs[g] + "; re-run on device " + std::to_string(devices[healthy]) + " failed too: " + g_err; return rc; }
        note += "device " + std::to_string(devices[g]) + " failed (" + errs[g] + "), its frames were re-run on device " + std::to_string(devices[healthy]) + "; ";
    }
    g_err = note;
    return B200_OK;
}

// A stream of frames in HOST memory through a device-side batch routine: upload, kernels and download of consecutive chunks overlap
// on three streams (three staging slots in flight).  `run(dsrc, src_layout, ddst, dst_layout, n, stream)` launches the work for one
// chunk; `mid_bytes_per_frame` asks for scratch (c.mid) sized for one chunk.
template <class Run>
static int stream_frames(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                         void* dst, size_t dfs, size_t dst_step, int dw, int dh,
                         int n_frames, int chunk_frames, size_t mid_bytes_per_frame, Run run)
{
    if (n_frames <= 0) return B200_OK;
    if (!src || !dst || sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0) return B200_NOT_IMPLEMENTED;
    const int depth = B200_DEPTH(src_type), cn = B200_CN(src_type);
    if (depth > D_32F || depth == 1 || depth == 4 || cn < 1 || cn > 4) return B200_NOT_IMPLEMENTED;
    const size_t esz = (size_t)(depth == D_8U ? 1 : depth == D_32F ? 4 : 2) * cn;
    const size_t srow = (size_t)sw * esz, drow = (size_t)dw * esz;
    if (src_step < srow || dst_step < drow || sfs < src_step * (size_t)(sh - 1) + srow || dfs < dst_step * (size_t)(dh - 1) + drow)
        return B200_NOT_IMPLEMENTED;
    int rc = stream_ctx_init();
    if (rc != B200_OK) return rc;
    StreamCtx& c = g_sctx;
    const FrameLayout ls = frame_layout(src_step, sfs, srow, sh, false), ld = frame_layout(dst_step, dfs, drow, dh, true);
    int chunk = chunk_frames;
    if (chunk <= 0) {                                   // ~24 MB of source per chunk, at least 3 chunks in flight
        chunk = (int)(((size_t)24 << 20) / (ls.frame_stride ? ls.frame_stride : 1));
        if (chunk > (n_frames + 2) / 3) chunk = (n_frames + 2) / 3;
    }
    if (chunk < 1) chunk = 1;
    if (chunk > n_frames) chunk = n_frames;
    if (chunk > 4096) chunk = 4096;
    if (mid_bytes_per_frame) CK(c.mid.reserve(mid_bytes_per_frame * (size_t)chunk));
    for (int i = 0; i < StreamCtx::NS; i++) {
        CK(c.src[i].reserve(ls.frame_stride * (size_t)chunk));
        CK(c.dst[i].reserve(ld.frame_stride * (size_t)chunk));
    }
    int k = 0;
    for (int f0 = 0; f0 < n_frames; f0 += chunk, k++) {
        const int nf = n_frames - f0 < chunk ? n_frames - f0 : chunk;
        const int slot = k % StreamCtx::NS;
        const char* hs = (const char*)src + (size_t)f0 * sfs;
        char* hd = (char*)dst + (size_t)f0 * dfs;
        // upload (waits until the kernels of the chunk that used this slot last have read it)
        if (k >= StreamCtx::NS) CK(cudaStreamWaitEvent(c.up, c.run_done[slot], 0));
        if (ls.block) CK(cudaMemcpyAsync(c.src[slot].p, hs, sfs * (size_t)(nf - 1) + src_step * (size_t)(sh - 1) + srow, cudaMemcpyHostToDevice, c.up));
        else for (int f = 0; f < nf; f++)
            CK(cudaMemcpy2DAsync((char*)c.src[slot].p + (size_t)f * ls.frame_stride, ls.step, hs + (size_t)f * sfs, src_step, srow, sh, cudaMemcpyHostToDevice, c.up));
        CK(cudaEventRecord(c.up_done[slot], c.up));
        // kernels (wait for the upload, and for the download that emptied this slot's output)
        CK(cudaStreamWaitEvent(c.run, c.up_done[slot], 0));
        if (k >= StreamCtx::NS) CK(cudaStreamWaitEvent(c.run, c.down_done[slot], 0));
        rc = run(c.src[slot].p, ls, c.dst[slot].p, ld, nf, c.run);
        if (rc != B200_OK) { cudaStreamSynchronize(c.up); cudaStreamSynchronize(c.run); cudaStreamSynchronize(c.down); return rc; }
        CK(cudaEventRecord(c.run_done[slot], c.run));
        // download
        CK(cudaStreamWaitEvent(c.down, c.run_done[slot], 0));
        if (ld.block) CK(cudaMemcpyAsync(hd, c.dst[slot].p, dfs * (size_t)nf, cudaMemcpyDeviceToHost, c.down));
        else for (int f = 0; f < nf; f++)
            CK(cudaMemcpy2DAsync(hd + (size_t)f * dfs, dst_step, (char*)c.dst[slot].p + (size_t)f * ld.frame_stride, ld.step, drow, dh, cudaMemcpyDeviceToHost, c.down));
        CK(cudaEventRecord(c.down_done[slot], c.down));
    }
    CK(cudaStreamSynchronize(c.down));
    CK(cudaStreamSynchronize(c.run));
    return B200_OK;
}

} // namespace

extern "C" {
#pragma GCC visibility push(default)

const char* b200_last_error(void) { return g_err.c_str(); }
const char* b200_version(void) { return "b200geom 0.1 (sm_100a)"; }
unsigned long long b200_launch_count(void) { return g_launches.load(); }
int b200_device_count(void) { int n = 0; return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0; }
void b200_profile_begin(void)
{
    for (auto& r : g_prof.recs) { g_prof.pool.push_back(r.e0); g_prof.pool.push_back(r.e1); }
    g_prof.recs.clear();
    g_prof.on = true;
}
int b200_profile_end(double* ms_by_kind, unsigned long long* launches_by_kind, int n_kinds)
{
    g_prof.on = false;
    for (int k = 0; k < n_kinds; k++) { if (ms_by_kind) ms_by_kind[k] = 0; if (launches_by_kind) launches_by_kind[k] = 0; }
    for (auto& r : g_prof.recs) {
        CK(cudaEventSynchronize(r.e1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, r.e0, r.e1));
        if (r.kind < n_kinds) { if (ms_by_kind) ms_by_kind[r.kind] += ms; if (launches_by_kind) launches_by_kind[r.kind]++; }
        g_prof.pool.push_back(r.e0); g_prof.pool.push_back(r.e1);
    }
    g_prof.recs.clear();
    return B200_OK;
}
int b200_set_force_generic(int on) { int old = g_force_generic.load(); g_force_generic.store(on ? 1 : 0); return old; }
long long b200_hal_set_min_pixels(long long v) { long long old = min_pixels(); g_min_pixels.store(v < 0 ? 0 : v); return old; }
int b200_hal_set_policy(int all) { const int old = hal_policy_all() ? 1 : 0; g_policy.store(all ? 1 : 0); return old; }

// ---- device-resident ----
int b200_warpAffine_dev(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step,
                        int dw, int dh, const double M[6], int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    return run_warp_dev(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_AFFINE, M, 0, 0, 0, 0, interp, border, bv), (cudaStream_t)stream);
}
int b200_warpPerspective_dev(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step,
                             int dw, int dh, const double M[9], int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    return run_warp_dev(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_PERSP, M, 0, 0, 0, 0, interp, border, bv), (cudaStream_t)stream);
}
int b200_remap32f_dev(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step, int dw, int dh,
                      const float* mapx, size_t mapx_step, const float* mapy, size_t mapy_step,
                      int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    if (!mapx || !mapy) return B200_NOT_IMPLEMENTED;
    return run_warp_dev(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_MAP32F, 0, mapx, mapx_step, mapy, mapy_step, interp, border, bv), (cudaStream_t)stream);
}
int b200_remap32fc2_dev(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step, int dw, int dh,
                        const float* map, size_t map_step, int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    if (!map) return B200_NOT_IMPLEMENTED;
    return run_warp_dev(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_MAP32FC2, 0, map, map_step, 0, 0, interp, border, bv), (cudaStream_t)stream);
}
int b200_remap16s_dev(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step, int dw, int dh,
                      const short* mapx, size_t mapx_step, const unsigned short* mapy, size_t mapy_step,
                      int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    if (!mapx) return B200_NOT_IMPLEMENTED;
    return run_warp_dev(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_MAP16S, 0, mapx, mapx_step, mapy, mapy_step, interp, border, bv), (cudaStream_t)stream);
}
int b200_convertMaps_dev(int map1_type, const void* map1, size_t map1_step, int map2_type, const void* map2, size_t map2_step,
                         int dstmap1_type, void* dstmap1, size_t dstmap1_step, void* dstmap2, size_t dstmap2_step,
                         int width, int height, int nninterpolation, void* stream)
{
    B200_TRACE();
    const int T_16SC2 = 3 + 8, T_32FC1 = 5, T_32FC2 = 5 + 8, T_16UC1 = 2, T_16SC1 = 3;
    if (width <= 0 || height <= 0 || !map1 || !dstmap1) return B200_NOT_IMPLEMENTED;
    const int nn = nninterpolation != 0;
    if (!map2) map2_type = -1;
    // the reference accepts the pair in either order (:1701-1705)
    if (map2_type == T_16SC2) { std::swap(map1, map2); std::swap(map1_step, map2_step); std::swap(map1_type, map2_type); }
    const bool ok_in = (map1_type == T_16SC2 && (nn || map2_type == T_16UC1 || map2_type == T_16SC1)) ||
                       (map1_type == T_32FC1 && map2_type == T_32FC1) || (map1_type == T_32FC2 && map2_type == -1);
    if (!ok_in) return B200_NOT_IMPLEMENTED;
    if (dstmap1_type <= 0) dstmap1_type = map1_type == T_16SC2 ? T_32FC2 : T_16SC2;                       // :1707-1708
    if (dstmap1_type != T_16SC2 && dstmap1_type != T_32FC1 && dstmap1_type != T_32FC2) return B200_NOT_IMPLEMENTED;
    const bool want2 = !nn && dstmap1_type != T_32FC2;                                                     // :1713-1719
    if (want2 && !dstmap2) return B200_NOT_IMPLEMENTED;
    if (!want2) dstmap2 = nullptr;
    if (map1_type == T_16SC2 && map2_type != T_16UC1 && map2_type != T_16SC1) map2 = nullptr;             // fraction plane absent
    MapConvArgs a;
    a.nn = nn;
    if (map1_type == dstmap1_type) a.mode = map1_type == T_16SC2 ? MC_COPY_S2 : map1_type == T_32FC1 ? MC_COPY_F1F1 : MC_COPY_F2;
    else if (map1_type == T_32FC1 && dstmap1_type == T_16SC2) a.mode = MC_F1F1_TO_S2;
    else if (map1_type == T_32FC2 && dstmap1_type == T_16SC2) a.mode = MC_F2_TO_S2;
    else if (map1_type == T_16SC2 && dstmap1_type == T_32FC2) a.mode = MC_S2_TO_F2;
    else if (map1_type == T_16SC2 && dstmap1_type == T_32FC1) { if (nn) return B200_NOT_IMPLEMENTED; a.mode = MC_S2_TO_F1F1; }   // no second plane to write
    else if (map1_type == T_32FC1 && dstmap1_type == T_32FC2) a.mode = MC_F1F1_TO_F2;
    else a.mode = MC_F2_TO_F1F1;
    if (a.mode == MC_COPY_S2 && map2_type != T_16UC1) map2 = nullptr;      // :1722-1723 copies the plane only when the types match
    // nninterpolate 16SC2 -> 32FC2 is convertTo (:1718-1723): the fraction plane is ignored
    if (nn && a.mode == MC_S2_TO_F2) map2 = nullptr;
    const int a1 = map1_type == T_32FC2 ? 8 : 4, ad = dstmap1_type == T_32FC2 ? 8 : 4;
    if (!aligned_to(map1, map1_step, a1) || !aligned_to(dstmap1, dstmap1_step, ad)) return B200_NOT_IMPLEMENTED;
    if (map2 && !aligned_to(map2, map2_step, map1_type == T_32FC1 ? 4 : 2)) return B200_NOT_IMPLEMENTED;
    if (dstmap2 && !aligned_to(dstmap2, dstmap2_step, dstmap1_type == T_32FC1 ? 4 : 2)) return B200_NOT_IMPLEMENTED;
    if (map1_step > 0x7fffffffu || map2_step > 0x7fffffffu || dstmap1_step > 0x7fffffffu || dstmap2_step > 0x7fffffffu || height > 65535) return B200_NOT_IMPLEMENTED;
    a.s1 = (const uint8_t*)map1; a.s1_step = (int)map1_step; a.s2 = (const uint8_t*)map2; a.s2_step = (int)map2_step;
    a.d1 = (uint8_t*)dstmap1; a.d1_step = (int)dstmap1_step; a.d2 = (uint8_t*)dstmap2; a.d2_step = (int)dstmap2_step;
    a.w = width; a.h = height;
    ProfScope prof(PK_OTHER, (cudaStream_t)stream);
    convert_maps_kernel<<<dim3((width + 255) / 256, height), 256, 0, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("convertMaps kernel launch", e);
    g_launches++;
    return B200_OK;
}

int b200_blobFromImagesWithParams_dev(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                                      float* blob, int bw, int bh, int n_frames,
                                      const double scalefactor[4], const double mean[4], int swapRB,
                                      int paddingmode, const double borderValue[4], void* stream)
{
    B200_TRACE();
    if (n_frames <= 0) return B200_OK;
    const int depth = B200_DEPTH(src_type), cn = B200_CN(src_type);
    if (depth != D_8U || !(cn == 1 || cn == 3 || cn == 4) || !src || !blob || n_frames > 65535) return B200_NOT_IMPLEMENTED;
    if (sw <= 0 || sh <= 0 || bw <= 0 || bh <= 0 || ((uintptr_t)blob & 3)) return B200_NOT_IMPLEMENTED;
    if (paddingmode < 0 || paddingmode > 2) return B200_NOT_IMPLEMENTED;
    // size of the image cv::resize produces and where the blob window sits in it (dnn_utils.cpp:303-337)
    int rw = bw, rh = bh, ox = 0, oy = 0;
    double inv = 0.0;
    if (bw == sw && bh == sh) paddingmode = 0;              // :307 `size != imgSize` guards all three branches
    if (paddingmode == 1) {                                 // DNN_PMODE_CROP_CENTER: resize(img, Size(), f, f), f = max ratio in float
        const float fa = bw / (float)sw, fb = bh / (float)sh, rf = fa > fb ? fa : fb;
        inv = rf;
        rw = host_rne(sw * inv); rh = host_rne(sh * inv);   // resize.cpp:4015-4016
        ox = (int)(0.5 * (rw - bw)); oy = (int)(0.5 * (rh - bh));
        if (rw <= 0 || rh <= 0 || ox < 0 || oy < 0 || ox + bw > rw || oy + bh > rh) return B200_NOT_IMPLEMENTED;
    } else if (paddingmode == 2) {                          // DNN_PMODE_LETTERBOX: resize(img, Size(int(w*f), int(h*f))), f = min ratio
        const float fa = bw / (float)sw, fb = bh / (float)sh, rf = fa < fb ? fa : fb;
        rh = (int)(sh * rf); rw = (int)(sw * rf);
        if (rw <= 0 || rh <= 0) return B200_NOT_IMPLEMENTED;
        oy = -((bh - rh) / 2); ox = -((bw - rw) / 2);
    }
    ResizeArgs a; int mode = -2;
    if (rw != sw || rh != sh) {
        // the scratch pointer only satisfies prepare_resize's checks; the kernel writes the blob
        ResizeCall rc{src_type, src, src_step, sw, sh, blob, (size_t)rw * cn, rw, rh, inv, inv, I_LINEAR};
        rc.src_frame_stride = sfs; rc.n_frames = n_frames;
        int d2, c2;
        int r = prepare_resize(rc, a, mode, d2, c2);
        if (r != B200_OK) return r;
        if (mode == RS_AREA_FAST && !(a.iscale_x == 2 && a.iscale_y == 2 && sw == 2 * rw && sh == 2 * rh)) return B200_NOT_IMPLEMENTED;
        if (mode != RS_LINEAR && mode != RS_AREA_FAST) return B200_NOT_IMPLEMENTED;
    } else {
        if (src_step > 0x7fffffffu) return B200_NOT_IMPLEMENTED;
        a.src = Image{(const uint8_t*)src, (int)src_step, sw, sh};
        a.dst = ImageOut{(uint8_t*)blob, rw * cn, rw, rh};
        a.src_frame_stride = (long long)sfs; a.dst_frame_stride = 0;
    }
    BlobArgs b;
    b.blob = blob; b.swap_rb = swapRB != 0 && cn >= 3;
    b.bw = bw; b.bh = bh; b.ox = ox; b.oy = oy;
    for (int k = 0; k < 4; k++) {
        b.mean[k] = mean ? (float)mean[k] : 0.f; b.scale[k] = scalefactor ? (float)scalefactor[k] : 1.f;
        const int pv = host_rne(borderValue ? borderValue[k] : 0.0);        // saturate_cast<uchar>(double)
        b.pad[k] = (uint8_t)(pv < 0 ? 0 : pv > 255 ? 255 : pv);
    }
    ProfScope prof(PK_OTHER, (cudaStream_t)stream);
    cudaError_t e = (cudaError_t)launch_blob_u8(a, b, mode, cn, n_frames, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail("blobFromImages kernel launch", e);
    g_launches++;
    return B200_OK;
}

int b200_blobFromImages_dev(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                            float* blob, int bw, int bh, int n_frames,
                            const double scalefactor[4], const double mean[4], int swapRB, void* stream)
{
    return b200_blobFromImagesWithParams_dev(src_type, src, sfs, src_step, sw, sh, blob, bw, bh, n_frames, scalefactor, mean, swapRB, 0, nullptr, stream);
}

int b200_resize_dev(int src_type, const void* src, size_t src_step, int sw, int sh, void* dst, size_t dst_step, int dw, int dh,
                    double inv_scale_x, double inv_scale_y, int interp, void* stream)
{
    B200_TRACE();
    ResizeCall c{src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y, interp};
    return run_resize_dev(c, (cudaStream_t)stream);
}

// ---- batched ----
int b200_resize_batch_dev(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                          void* dst, size_t dfs, size_t dst_step, int dw, int dh,
                          int n_frames, double inv_scale_x, double inv_scale_y, int interp, void* stream)
{
    B200_TRACE();
    if (n_frames <= 0) return B200_OK;
    if (n_frames > 65535) return B200_NOT_IMPLEMENTED;
    ResizeCall c{src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y, interp};
    c.src_frame_stride = sfs; c.dst_frame_stride = dfs; c.n_frames = n_frames;
    return run_resize_dev(c, (cudaStream_t)stream);
}
int b200_warpAffine_batch_dev(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                              void* dst, size_t dfs, size_t dst_step, int dw, int dh,
                              int n_frames, const double M[6], int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    if (n_frames <= 0) return B200_OK;
    if (n_frames > 65535) return B200_NOT_IMPLEMENTED;
    WarpCall c = make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_AFFINE, M, 0, 0, 0, 0, interp, border, bv);
    c.src_frame_stride = sfs; c.dst_frame_stride = dfs; c.n_frames = n_frames;
    return run_warp_dev(c, (cudaStream_t)stream);
}
int b200_warpPerspective_batch_dev(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                                   void* dst, size_t dfs, size_t dst_step, int dw, int dh,
                                   int n_frames, const double M[9], int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    if (n_frames <= 0) return B200_OK;
    if (n_frames > 65535) return B200_NOT_IMPLEMENTED;
    WarpCall c = make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_PERSP, M, 0, 0, 0, 0, interp, border, bv);
    c.src_frame_stride = sfs; c.dst_frame_stride = dfs; c.n_frames = n_frames;
    return run_warp_dev(c, (cudaStream_t)stream);
}
int b200_remap_batch_dev(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                         void* dst, size_t dfs, size_t dst_step, int dw, int dh, int n_frames,
                         int map1_type, const void* map1, size_t map1_step, const void* map2, size_t map2_step,
                         int interp, int border, const double bv[4], void* stream)
{
    B200_TRACE();
    if (n_frames <= 0) return B200_OK;
    if (n_frames > 65535 || !map1) return B200_NOT_IMPLEMENTED;
    int coord;
    if (map1_type == 5 && map2) coord = COORD_MAP32F;                  // CV_32FC1 pair    (cv_hal_remap32f)
    else if (map1_type == 13) { coord = COORD_MAP32FC2; map2 = nullptr; map2_step = 0; }   // CV_32FC2 (cv_hal_remap32fc2)
    else if (map1_type == 11) coord = COORD_MAP16S;                    // CV_16SC2 [+ CV_16UC1] (cv_hal_remap16s)
    else return B200_NOT_IMPLEMENTED;
    WarpCall c = make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, coord, 0, map1, map1_step, map2, map2_step, interp, border, bv);
    c.src_frame_stride = sfs; c.dst_frame_stride = dfs; c.n_frames = n_frames;
    return run_warp_dev(c, (cudaStream_t)stream);
}
int b200_resize_warpAffine_batch_dev(int src_type,
                                     const void* src, size_t sfs, size_t src_step, int sw, int sh,
                                     void* mid, size_t mfs, size_t mid_step, int mw, int mh,
                                     void* dst, size_t dfs, size_t dst_step, int dw, int dh,
                                     int n_frames, int resize_interp, const double M[6], int warp_interp, int border,
                                     const double bv[4], void* stream)
{
    B200_TRACE();
    if (!mid) return B200_NOT_IMPLEMENTED;
    if (n_frames <= 0) return B200_OK;
    if (n_frames > 65535) return B200_NOT_IMPLEMENTED;
    // Two batched launches.  Both write with evict-first stores: measured on B200, leaving the intermediate
    // frames in L2 with plain stores (even in L2-sized sub-batches) costs the box kernel 40 % of its DRAM
    // throughput, more than the saved re-read of `mid` is worth (profiles/r1_notes.md).
    ResizeCall rc_{src_type, src, src_step, sw, sh, mid, mid_step, mw, mh, 0.0, 0.0, resize_interp};
    rc_.src_frame_stride = sfs; rc_.dst_frame_stride = mfs; rc_.n_frames = n_frames;
    WarpCall wc = make_warp(src_type, mid, mid_step, mw, mh, dst, dst_step, dw, dh, COORD_AFFINE, M, 0, 0, 0, 0, warp_interp, border, bv);
    wc.src_frame_stride = mfs; wc.dst_frame_stride = dfs; wc.n_frames = n_frames;
    {
        // 8U, exact 2x2 box, bilinear warp: ONE persistent launch whose half-size frames stay in L2 (fast_kernels.cu)
        ResizeArgs ra; WarpArgs wa; int mode, depth, cn, wdepth, wcn;
        if (prepare_resize(rc_, ra, mode, depth, cn) == B200_OK && prepare_warp(wc, wa, wdepth, wcn) == B200_OK &&
            depth == D_8U && mode == RS_AREA_FAST && ra.iscale_x == 2 && ra.iscale_y == 2) {
            ProfScope prof(PK_WARP, (cudaStream_t)stream);
            const int r = try_launch_pipe_box_warp_u8(ra, wa, cn, n_frames, (size_t)n_frames * mfs, (cudaStream_t)stream);
            if (r == 0) { g_launches++; return B200_OK; }
            prof.cancel();
            if (r > 0) return fail("pipeline kernel launch", (cudaError_t)r);
        }
    }
    int rc = run_resize_dev(rc_, (cudaStream_t)stream);
    if (rc != B200_OK) return rc;
    rc = run_warp_dev(wc, (cudaStream_t)stream);
    if (rc != B200_OK) return rc;
    return B200_OK;
}

int b200_hal_resize_warpAffine_batch(int src_type,
                                     const void* src, size_t sfs, size_t src_step, int sw, int sh,
                                     int mw, int mh,
                                     void* dst, size_t dfs, size_t dst_step, int dw, int dh,
                                     int n_frames, int resize_interp, const double M[6], int warp_interp, int border,
                                     const double bv[4], int chunk_frames)
{
    B200_TRACE();
    if (mw <= 0 || mh <= 0) return B200_NOT_IMPLEMENTED;
    // BORDER_TRANSPARENT keeps the caller's dst where the source lies outside; the staging slots of these entries never see
    // the caller's dst, so it is declined here (the per-image hook b200_hal_warpAffine uploads dst and handles it)
    if (border == B_TRANSPARENT) return B200_NOT_IMPLEMENTED;
    const int depth = B200_DEPTH(src_type), cn = B200_CN(src_type);
    const size_t esz = (size_t)(depth == D_8U ? 1 : depth == D_32F ? 4 : 2) * (cn > 0 ? cn : 1);
    const size_t mstep = pitch_for((size_t)mw * esz), mfs = mstep * (size_t)mh;
    return stream_frames(src_type, src, sfs, src_step, sw, sh, dst, dfs, dst_step, dw, dh, n_frames, chunk_frames, mfs,
        [&](void* ds, const FrameLayout& ls, void* dd, const FrameLayout& ld, int nf, cudaStream_t st) {
            return b200_resize_warpAffine_batch_dev(src_type, ds, ls.frame_stride, ls.step, sw, sh, g_sctx.mid.p, mfs, mstep, mw, mh,
                                                    dd, ld.frame_stride, ld.step, dw, dh, nf, resize_interp, M, warp_interp, border, bv, st);
        });
}

// the same chunked host pipeline around the single batched operations (BASELINE configs 2-4 as frame streams)
int b200_hal_resize_batch(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                          void* dst, size_t dfs, size_t dst_step, int dw, int dh, int n_frames,
                          double inv_scale_x, double inv_scale_y, int interpolation, int chunk_frames)
{
    B200_TRACE();
    return stream_frames(src_type, src, sfs, src_step, sw, sh, dst, dfs, dst_step, dw, dh, n_frames, chunk_frames, 0,
        [&](void* ds, const FrameLayout& ls, void* dd, const FrameLayout& ld, int nf, cudaStream_t st) {
            return b200_resize_batch_dev(src_type, ds, ls.frame_stride, ls.step, sw, sh, dd, ld.frame_stride, ld.step, dw, dh, nf,
                                         inv_scale_x, inv_scale_y, interpolation, st);
        });
}

int b200_hal_warpAffine_batch(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                              void* dst, size_t dfs, size_t dst_step, int dw, int dh, int n_frames,
                              const double M[6], int interpolation, int border, const double bv[4], int chunk_frames)
{
    B200_TRACE();
    if (border == B_TRANSPARENT) return B200_NOT_IMPLEMENTED;
    return stream_frames(src_type, src, sfs, src_step, sw, sh, dst, dfs, dst_step, dw, dh, n_frames, chunk_frames, 0,
        [&](void* ds, const FrameLayout& ls, void* dd, const FrameLayout& ld, int nf, cudaStream_t st) {
            return b200_warpAffine_batch_dev(src_type, ds, ls.frame_stride, ls.step, sw, sh, dd, ld.frame_stride, ld.step, dw, dh, nf, M, interpolation, border, bv, st);
        });
}

int b200_hal_warpPerspective_batch(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                                   void* dst, size_t dfs, size_t dst_step, int dw, int dh, int n_frames,
                                   const double M[9], int interpolation, int border, const double bv[4], int chunk_frames)
{
    B200_TRACE();
    if (border == B_TRANSPARENT) return B200_NOT_IMPLEMENTED;
    return stream_frames(src_type, src, sfs, src_step, sw, sh, dst, dfs, dst_step, dw, dh, n_frames, chunk_frames, 0,
        [&](void* ds, const FrameLayout& ls, void* dd, const FrameLayout& ld, int nf, cudaStream_t st) {
            return b200_warpPerspective_batch_dev(src_type, ds, ls.frame_stride, ls.step, sw, sh, dd, ld.frame_stride, ld.step, dw, dh, nf, M, interpolation, border, bv, st);
        });
}

// frames through ONE map held in host memory: the map is uploaded once (c.mid), the frames stream
int b200_hal_remap_batch(int src_type, const void* src, size_t sfs, size_t src_step, int sw, int sh,
                         void* dst, size_t dfs, size_t dst_step, int dw, int dh, int n_frames,
                         int map1_type, const void* map1, size_t map1_step, const void* map2, size_t map2_step,
                         int interpolation, int border, const double bv[4], int chunk_frames)
{
    B200_TRACE();
    if (border == B_TRANSPARENT || !map1 || dw <= 0 || dh <= 0) return B200_NOT_IMPLEMENTED;
    if (!(map1_type == 5 || map1_type == 13 || map1_type == 11) || (map1_type == 5 && !map2)) return B200_NOT_IMPLEMENTED;
    const size_t r1 = (size_t)dw * (map1_type == 13 ? 8 : 4), r2 = map2 ? (size_t)dw * (map1_type == 5 ? 4 : 2) : 0;
    if (map1_step < r1 || (map2 && map2_step < r2)) return B200_NOT_IMPLEMENTED;
    int rc = stream_ctx_init();
    if (rc != B200_OK) return rc;
    StreamCtx& c = g_sctx;
    const size_t p1 = pitch_for(r1), p2 = map2 ? pitch_for(r2) : 0, off2 = (p1 * (size_t)dh + 255) & ~(size_t)255;
    CK(c.maps.reserve(off2 + p2 * (size_t)dh));
    CK(cudaMemcpy2DAsync(c.maps.p, p1, map1, map1_step, r1, dh, cudaMemcpyHostToDevice, c.run));
    if (map2) CK(cudaMemcpy2DAsync((char*)c.maps.p + off2, p2, map2, map2_step, r2, dh, cudaMemcpyHostToDevice, c.run));
    return stream_frames(src_type, src, sfs, src_step, sw, sh, dst, dfs, dst_step, dw, dh, n_frames, chunk_frames, 0,
        [&](void* ds, const FrameLayout& ls, void* dd, const FrameLayout& ld, int nf, cudaStream_t st) {
            return b200_remap_batch_dev(src_type, ds, ls.frame_stride, ls.step, sw, sh, dd, ld.frame_stride, ld.step, dw, dh, nf,
                                        map1_type, c.maps.p, p1, map2 ? (char*)c.maps.p + off2 : nullptr, p2, interpolation, border, bv, st);
        });
}

int b200_synth_frames_dev(void* dst, size_t frame_stride, size_t step, int width, int height, int cn,
                          int first_frame, int n_frames, unsigned int seed, void* stream)
{
    if (n_frames <= 0) return B200_OK;
    const int row_bytes = width * cn;
    dim3 block(256), grid(((row_bytes + 3) / 4 + 255) / 256, height, n_frames);
    synth_frames_kernel<<<grid, block, 0, (cudaStream_t)stream>>>((uint8_t*)dst, (long long)frame_stride, (long long)step, row_bytes, height, first_frame, seed);
    CK(cudaGetLastError());
    g_launches++;
    return B200_OK;
}

// ---- HAL hooks (host pointers) ----
int b200_hal_warpAffine(int src_type, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step,
                        int dw, int dh, const double M[6], int interp, int border, const double bv[4])
{
    B200_TRACE();
    return hal_warp(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_AFFINE, M, 0, 0, 0, 0, interp, border, bv));
}
int b200_hal_warpPerspective(int src_type, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step,
                             int dw, int dh, const double M[9], int interp, int border, const double bv[4])
{
    B200_TRACE();
    return hal_warp(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_PERSP, M, 0, 0, 0, 0, interp, border, bv));
}
int b200_hal_remap32f(int src_type, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step, int dw, int dh,
                      float* mapx, size_t mapx_step, float* mapy, size_t mapy_step, int interp, int border, const double bv[4])
{
    B200_TRACE();
    if (!mapx || !mapy) return B200_NOT_IMPLEMENTED;
    return hal_warp(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_MAP32F, 0, mapx, mapx_step, mapy, mapy_step, interp, border, bv));
}
int b200_hal_remap32fc2(int src_type, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step, int dw, int dh,
                        float* map, size_t map_step, int interp, int border, const double bv[4])
{
    B200_TRACE();
    if (!map) return B200_NOT_IMPLEMENTED;
    return hal_warp(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_MAP32FC2, 0, map, map_step, 0, 0, interp, border, bv));
}
int b200_hal_remap16s(int src_type, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step, int dw, int dh,
                      short* mapx, size_t mapx_step, unsigned short* mapy, size_t mapy_step, int interp, int border, const double bv[4])
{
    B200_TRACE();
    if (!mapx) return B200_NOT_IMPLEMENTED;
    return hal_warp(make_warp(src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, COORD_MAP16S, 0, mapx, mapx_step, mapy, mapy_step, interp, border, bv));
}
int b200_hal_resize(int src_type, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step,
                    int dw, int dh, double inv_scale_x, double inv_scale_y, int interp)
{
    B200_TRACE();
    if (!hal_enabled()) return B200_NOT_IMPLEMENTED;
    if ((long long)dw * dh < min_pixels()) return B200_NOT_IMPLEMENTED;
    ResizeCall c{src_type, src, src_step, sw, sh, dst, dst_step, dw, dh, inv_scale_x, inv_scale_y, interp};
    const int depth0 = B200_DEPTH(src_type), cn0 = B200_CN(src_type);
    ResizeArgs pre; int pre_mode, pre_depth, pre_cn;
    {
        // more than 4 channels: groups of 4 (below), under the rules of a 4-channel image
        if (cn0 > 4) c.src_type = depth0 + (3 << 3);
        int rc = prepare_resize(c, pre, pre_mode, pre_depth, pre_cn);
        if (rc != B200_OK) return rc;
        // ... where the reference treats channels alike whatever their number (tests/test_oracle_pinned.py pins which modes do).  Left to
        // the CPU: the 2x2 decimation (the vector path of 1 / 3 / 4 channels rounds half up, the generic one the reference takes for
        // more channels rounds half to even), fractional INTER_AREA (CV_Assert(cn <= 4), resize.cpp:3851), and floating-point CUBIC /
        // LANCZOS4 (the split between vector body and scalar tail, which evaluate in different orders, moves with width * cn)
        if (cn0 > 4 && (cn0 > 512 || pre_mode == RS_AREA || (pre_mode == RS_AREA_FAST && pre.iscale_x == 2 && pre.iscale_y == 2) ||
                        ((pre_mode == RS_CUBIC || pre_mode == RS_LANCZOS4) && depth0 >= D_32F))) return B200_NOT_IMPLEMENTED;
    }
    // (integer-scale INTER_AREA: only the 2x2 case has the reference's vector fast path; 4x4 at 1080p costs the host 1.3 ms, the hook 0.4)
    if (!hal_policy_all() && pre_mode != RS_AREA && pre_mode != RS_CUBIC && pre_mode != RS_LANCZOS4 &&
        !(pre_mode == RS_AREA_FAST && !(pre.iscale_x == 2 && pre.iscale_y == 2))) return B200_NOT_IMPLEMENTED;
    int rc = hal_ctx_init();
    if (rc != B200_OK) return rc;
    cudaStream_t s = g_ctx.stream;
    std::vector<float> ltab;
    if (pre_mode == RS_LANCZOS4) {
        build_lanczos4_tables(pre, dw, dh, ltab);
        HCK(g_ctx.map1.reserve(ltab.size() * sizeof(float)));
        HCK(cudaMemcpyAsync(g_ctx.map1.p, ltab.data(), ltab.size() * sizeof(float), cudaMemcpyHostToDevice, s));
        c.ltab_x = (const float*)g_ctx.map1.p; c.ltab_y = c.ltab_x + (size_t)dw * 8;
    }
    const int e1 = elem_size1(depth0), esz = e1 * cn0;
    size_t sp, dp;
    if ((rc = upload(g_ctx.src, src, src_step, (size_t)sw * esz, sh, sp, s)) != B200_OK) return rc;
    dp = out_pitch(dst_step, (size_t)dw * esz);
    HCK(g_ctx.dst.reserve(dp * (size_t)dh));
    c.src = g_ctx.src.p; c.src_step = sp; c.dst = g_ctx.dst.p; c.dst_step = dp;
    if (cn0 <= 4) rc = run_resize_dev(c, s);
    else {
        const size_t gsp = pitch_for((size_t)sw * 4 * e1), gdp = pitch_for((size_t)dw * 4 * e1);
        HCK(g_ctx.grp_src.reserve(gsp * (size_t)sh));
        HCK(g_ctx.grp_dst.reserve(gdp * (size_t)dh));
        for (int c0 = 0; c0 < cn0 && rc == B200_OK; c0 += 4) {
            const int k = cn0 - c0 < 4 ? cn0 - c0 : 4;
            ResizeCall g = c;
            g.src_type = depth0 + ((k - 1) << 3);
            g.src = g_ctx.grp_src.p; g.src_step = gsp; g.dst = g_ctx.grp_dst.p; g.dst_step = gdp;
            rc = channel_group<true>(e1, g_ctx.src.p, sp, cn0, c0, k, g_ctx.grp_src.p, gsp, sw, sh, s);
            if (rc == B200_OK) rc = run_resize_dev(g, s);
            if (rc == B200_OK) rc = channel_group<false>(e1, g_ctx.dst.p, dp, cn0, c0, k, g_ctx.grp_dst.p, gdp, dw, dh, s);
        }
    }
    if (rc == B200_ERROR) return hal_decline("resize kernel launch", cudaGetLastError());
    if (rc != B200_OK) return rc;
    HCK(cudaStreamSynchronize(s));                      // see hal_warp: everything before the copy back may still decline
    if ((rc = download(dst, dst_step, g_ctx.dst.p, dp, (size_t)dw * esz, dh, s)) != B200_OK) return rc;
    CK(cudaStreamSynchronize(s));
    return B200_OK;
}

// cv_hal_pyrdown / cv_hal_pyrup: upload, one kernel, download; every decline happens before dst is written
static int hal_pyr(bool up, const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step, int dw, int dh,
                   int depth, int cn, int border_type)
{
    if (!hal_enabled()) return B200_NOT_IMPLEMENTED;
    if ((long long)dw * dh < min_pixels()) return B200_NOT_IMPLEMENTED;
    if (!src || !dst || cn < 1 || cn > 4 || depth < 0 || depth > 7) return B200_NOT_IMPLEMENTED;
    // default policy (measured on the reference's perf_pyramids grid, profiles/r2_perf_view.md): the reference's pyrDown has vector code
    // for 1, 2 and 4 channels that 16 host threads run faster than a pageable round trip over PCIe; 8UC3 is the case the GPU wins
    // (1.3-2.1x).  pyrUp wins in every cell (1.2-6.4x).
    if (!up && !hal_policy_all() && !(cn == 3 && depth == D_8U)) return B200_NOT_IMPLEMENTED;
    const int type = depth + ((cn - 1) << 3);
    int rc = up ? b200::pyrup_check(type, sw, sh, dw, dh, border_type) : b200::pyrdown_check(type, sw, sh, dw, dh, border_type);
    if (rc != B200_OK) return rc;
    if ((rc = hal_ctx_init()) != B200_OK) return rc;
    cudaStream_t s = g_ctx.stream;
    const int esz = elem_size1(depth) * cn;
    size_t sp, dp;
    if ((rc = upload(g_ctx.src, src, src_step, (size_t)sw * esz, sh, sp, s)) != B200_OK) return rc;
    dp = out_pitch(dst_step, (size_t)dw * esz);
    HCK(g_ctx.dst.reserve(dp * (size_t)dh));
    rc = up ? b200::launch_pyrup(type, g_ctx.src.p, sp, 0, sw, sh, g_ctx.dst.p, dp, 0, dw, dh, 1, border_type, s)
            : b200::launch_pyrdown(type, g_ctx.src.p, sp, 0, sw, sh, g_ctx.dst.p, dp, 0, dw, dh, 1, border_type, s);
    if (rc == B200_ERROR) return hal_decline(up ? "pyrUp kernel launch" : "pyrDown kernel launch", cudaGetLastError());
    if (rc != B200_OK) return rc;
    HCK(cudaStreamSynchronize(s));
    if ((rc = download(dst, dst_step, g_ctx.dst.p, dp, (size_t)dw * esz, dh, s)) != B200_OK) return rc;
    CK(cudaStreamSynchronize(s));
    return B200_OK;
}

// replaces cv_hal_pyrdown (hal_replacement.hpp:1461-1464; called from pyramids.cpp:1292)
int b200_hal_pyrdown(const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step, int dw, int dh,
                     int depth, int cn, int border_type)
{
    B200_TRACE();
    return hal_pyr(false, src, src_step, sw, sh, dst, dst_step, dw, dh, depth, cn, border_type);
}

// replaces cv_hal_pyrup (hal_replacement.hpp:1508-1511; called from pyramids.cpp:1391)
int b200_hal_pyrup(const unsigned char* src, size_t src_step, int sw, int sh, unsigned char* dst, size_t dst_step, int dw, int dh,
                   int depth, int cn, int border_type)
{
    B200_TRACE();
    return hal_pyr(true, src, src_step, sw, sh, dst, dst_step, dw, dh, depth, cn, border_type);
}

// ---- multi-GPU frame scheduler ----
void b200_sched_partition(int n_frames, int world, int rank, int* first, int* count)
{
    long long a = (long long)n_frames * rank / world, b = (long long)n_frames * (rank + 1) / world;
    if (first) *first = (int)a;
    if (count) *count = (int)(b - a);
}

int b200_sched_run_pipeline(const int* devices, int n_devices, int n_frames, int chunk,
                            int sw, int sh, int cn, int mw, int mh, int dw, int dh, const double M[6], unsigned int seed,
                            float* elapsed_ms, unsigned long long* checksums)
{
    B200_TRACE();
    if (n_devices <= 0 || n_frames <= 0 || chunk <= 0 || cn < 1 || cn > 4) return B200_NOT_IMPLEMENTED;
    std::vector<int> rcs(n_devices, B200_OK);
    std::vector<std::string> errs(n_devices);
    std::vector<std::thread> th;
    const int type = D_8U + ((cn - 1) << 3);
    // frame block g of the partition, run on device `dev` (its own, or a surviving one when its own failed)
    auto run_block = [&](int g, int dev) -> int {
        {
            {
                CK(cudaSetDevice(dev));
                int first, count;
                b200_sched_partition(n_frames, n_devices, g, &first, &count);
                cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
                const size_t sstep = pitch_for((size_t)sw * cn), mstep = pitch_for((size_t)mw * cn), dstep = pitch_for((size_t)dw * cn);
                const size_t sfs = sstep * sh, mfs = mstep * mh, dfs = dstep * dh;
                const int nb = std::min(chunk, std::max(count, 1));
                void *d_src = 0, *d_mid = 0, *d_dst = 0; unsigned long long* d_sum = 0;
                cudaEvent_t e0 = nullptr, e1 = nullptr;
                struct Cleanup {                       // every early return below releases what was created so far
                    void*& a; void*& b; void*& c; unsigned long long*& d; cudaEvent_t& e0; cudaEvent_t& e1; cudaStream_t& s;
                    ~Cleanup() {
                        cudaFree(a); cudaFree(b); cudaFree(c); cudaFree(d);
                        if (e0) cudaEventDestroy(e0);
                        if (e1) cudaEventDestroy(e1);
                        cudaStreamDestroy(s);
                    }
                } cleanup{d_src, d_mid, d_dst, d_sum, e0, e1, s};
                CK(cudaMalloc(&d_src, sfs * nb)); CK(cudaMalloc(&d_mid, mfs * nb)); CK(cudaMalloc(&d_dst, dfs * nb));
                CK(cudaMalloc(&d_sum, 8)); CK(cudaMemsetAsync(d_sum, 0, 8, s));
                CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
                float total = 0.f;
                const double bv[4] = {0, 0, 0, 0};
                for (int f = 0; f < count; f += nb) {
                    const int n = std::min(nb, count - f);
                    int rc = b200_synth_frames_dev(d_src, sfs, sstep, sw, sh, cn, first + f, n, seed, s);
                    if (rc != B200_OK) return rc;
                    CK(cudaEventRecord(e0, s));
                    rc = b200_resize_warpAffine_batch_dev(type, d_src, sfs, sstep, sw, sh, d_mid, mfs, mstep, mw, mh,
                                                          d_dst, dfs, dstep, dw, dh, n, I_LINEAR, M, I_LINEAR, B_CONSTANT, bv, s);
                    if (rc != B200_OK) return rc;
                    CK(cudaEventRecord(e1, s));
                    dim3 block(256), grid((dw * cn + 255) / 256, dh, n);
                    checksum_kernel<<<grid, block, 0, s>>>((const uint8_t*)d_dst, (long long)dfs, (long long)dstep, dw * cn, dh, first + f, d_sum);
                    CK(cudaGetLastError());
                    g_launches++;
                    CK(cudaStreamSynchronize(s));
                    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
                    total += ms;
                }
                if (checksums) CK(cudaMemcpy(&checksums[g], d_sum, 8, cudaMemcpyDeviceToHost));
                if (elapsed_ms) elapsed_ms[g] = total;
                return B200_OK;
            }
        }
    };
    for (int g = 0; g < n_devices; g++)
        th.emplace_back([&, g]() {
            rcs[g] = run_block(g, devices[g]);
            if (rcs[g] != B200_OK) errs[g] = g_err;
        });
    for (auto& t : th) t.join();
    return sched_redispatch(devices, n_devices, rcs, errs, run_block);
}

// Host frames over several GPUs: device g gets the contiguous block [g*n/G, (g+1)*n/G) (b200_sched_partition) and streams it
// through b200_hal_resize_warpAffine_batch on its own host thread (its own PCIe link, three streams, no collective).
int b200_sched_run_pipeline_host(const int* devices, int n_devices, int src_type,
                                 const void* src, size_t sfs, size_t src_step, int sw, int sh, int mw, int mh,
                                 void* dst, size_t dfs, size_t dst_step, int dw, int dh, int n_frames,
                                 int resize_interp, const double M[6], int warp_interp, int border, const double bv[4],
                                 float* elapsed_ms)
{
    B200_TRACE();
    if (n_devices <= 0 || !devices) return B200_NOT_IMPLEMENTED;
    if (n_frames <= 0) return B200_OK;
    std::vector<int> rcs(n_devices, B200_OK);
    std::vector<std::string> errs(n_devices);
    std::vector<std::thread> th;
    auto run_block = [&](int g, int dev) -> int {
        {
            {
                CK(cudaSetDevice(dev));
                int first, count;
                b200_sched_partition(n_frames, n_devices, g, &first, &count);
                const auto t0 = std::chrono::steady_clock::now();
                const int rc = b200_hal_resize_warpAffine_batch(src_type, (const char*)src + (size_t)first * sfs, sfs, src_step, sw, sh, mw, mh,
                                                                (char*)dst + (size_t)first * dfs, dfs, dst_step, dw, dh, count,
                                                                resize_interp, M, warp_interp, border, bv, 0);
                if (elapsed_ms) elapsed_ms[g] = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
                return rc;
            }
        }
    };
    for (int g = 0; g < n_devices; g++)
        th.emplace_back([&, g]() {
            rcs[g] = run_block(g, devices[g]);
            if (rcs[g] != B200_OK) errs[g] = g_err;
        });
    for (auto& t : th) t.join();
    return sched_redispatch(devices, n_devices, rcs, errs, run_block);
}

#pragma GCC visibility pop
} // extern "C"
