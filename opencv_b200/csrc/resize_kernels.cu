// cv::resize: INTER_NEAREST, INTER_LINEAR (generic 2-tap, plus INTER_AREA's "area mode" when a scale
// is < 1), integer-scale area-fast box (incl. INTER_LINEAR at exact 2x) and fractional INTER_AREA.
// Generic kernels, one thread per destination pixel; every per-column / per-row coefficient the
// reference precomputes on the host (xofs/alpha/yofs/beta, DecimateAlpha tables) is evaluated inside
// the kernel with the same fp64/fp32 operation sequence, so no host tables and no scratch memory.
// Reference: modules/imgproc/src/resize.cpp:1026-1172 (NN), :1877-1989,:2166-2266,:3876-3996 (linear),
// :2919-3069,:3824-3848 (area fast), :3183-3373,:3850-3872 (area).
#include "resize_kernels.cuh"
#include "fast_kernels.cuh"
#include <float.h>
#include <math.h>
#include <type_traits>
#include <algorithm>

namespace b200 {

constexpr int RT_W = 32, RT_H = 8;

// ------------------------------------------------------------------------------- NEAREST --------
// sx = min(cvFloor(x*ifx), W-1), sy = min(cvFloor(y*ify), H-1)  (resize.cpp:1052,1132-1133); element copy
// INTER_NEAREST_EXACT (EXACT = true): sx = min((ifx*x + ifx0) >> 16, W-1) in 16.16 integer arithmetic, pixel centres
// (resizeNN_bitexact, resize.cpp:1267-1288)
template <typename U, int N, bool EXACT>
__global__ void __launch_bounds__(RT_W * RT_H) resize_nn_kernel(const ResizeArgs a)
{
    const int x = blockIdx.x * RT_W + threadIdx.x, y = blockIdx.y * RT_H + threadIdx.y;
    if (x >= a.dst.w || y >= a.dst.h) return;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride;
    int sx, sy;
    if (EXACT) {
        sx = min((int)((unsigned)a.ifx * (unsigned)x + (unsigned)a.ifx0) >> 16, a.src.w - 1);
        sy = min((int)((unsigned)a.ify * (unsigned)y + (unsigned)a.ify0) >> 16, a.src.h - 1);
    } else {
        sx = min(floor_d(__dmul_rn((double)x, a.ifxd)), a.src.w - 1);
        sy = min(floor_d(__dmul_rn((double)y, a.ifyd)), a.src.h - 1);
    }
    const U* S = reinterpret_cast<const U*>(src + (long long)sy * a.src.step) + (long long)sx * N;
    U* D = reinterpret_cast<U*>(dst + (long long)y * a.dst.step) + (long long)x * N;
#pragma unroll
    for (int k = 0; k < N; k++) store_out(&D[k], S[k], a.keep_in_l2);
}

// ------------------------------------------------------------------------------- LINEAR ---------
// per-axis source index and fraction (resize.cpp:3899-3916 for x, :3955-3967 for y)
__device__ __forceinline__ void lin_coord(int d, double scale, double inv_scale, int area_mode, int& s, float& f)
{
    if (!area_mode) {
        f = __double2float_rn(__dsub_rn(__dmul_rn(__dadd_rn((double)d, 0.5), scale), 0.5));
        s = floor_f(f);
        f = fsub(f, (float)s);
    } else {
        s = floor_d(__dmul_rn((double)d, scale));
        f = __double2float_rn(__dsub_rn((double)(d + 1), __dmul_rn((double)(s + 1), inv_scale)));
        f = f <= 0 ? 0.f : fsub(f, (float)floor_f(f));
    }
}

template <typename T> struct LinOps;
// 8U: int32 rows scaled by 2^11, V pass (((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2  (resize.cpp:1961-1989)
template <> struct LinOps<uint8_t> {
    typedef uint16_t HT;                               // what the separable path keeps of a horizontal result: h >> 4 (<= 32640)
    int a0, a1, b0, b1;
    __device__ __forceinline__ LinOps(float fx) {
        a0 = sat_s16(rne_f(fmul(fsub(1.f, fx), 2048.f))); a1 = sat_s16(rne_f(fmul(fx, 2048.f)));   // :3947
    }
    // vertical coefficients arrive pre-converted from the tile's row table (ib = b0 | b1 << 16, :3989)
    static __device__ __forceinline__ int pack_row(float fy) {
        return sat_s16(rne_f(fmul(fsub(1.f, fy), 2048.f))) | (sat_s16(rne_f(fmul(fy, 2048.f))) << 16);
    }
    __device__ __forceinline__ void set_row(float, int ib) { b0 = (short)(ib & 0xFFFF); b1 = ib >> 16; }
    __device__ __forceinline__ HT hpass(uint8_t s0, uint8_t s1, bool edge) const {                  // :1898-1925
        return (HT)((edge ? s0 * 2048 : s0 * a0 + s1 * a1) >> 4);
    }
    __device__ __forceinline__ uint8_t vpass(int h0, int h1) const { return (uint8_t)((((b0 * h0) >> 16) + ((b1 * h1) >> 16) + 2) >> 2); }
    __device__ __forceinline__ uint8_t operator()(uint8_t s00, uint8_t s01, uint8_t s10, uint8_t s11, bool edge) const {
        return vpass(hpass(s00, s01, edge), hpass(s10, s11, edge));
    }
};
template <typename T> struct LinOpsF {
    typedef float HT;
    float a0, a1, b0, b1;
    __device__ __forceinline__ LinOpsF(float fx) : a0(fsub(1.f, fx)), a1(fx), b0(0.f), b1(0.f) {}   // a0 hoisted out of the row loop
    static __device__ __forceinline__ int pack_row(float) { return 0; }
    __device__ __forceinline__ void set_row(float fy, int) { b0 = fsub(1.f, fy); b1 = fy; }
    __device__ __forceinline__ float hpass(T s0, T s1, bool edge) const {
        return edge ? Elem<T>::to_f(s0) : fadd(fmul(Elem<T>::to_f(s0), a0), fmul(Elem<T>::to_f(s1), a1));
    }
    __device__ __forceinline__ T vpass(float h0, float h1) const { return Elem<T>::from_f(fadd(fmul(h0, b0), fmul(h1, b1))); }   // :1950-1959
    __device__ __forceinline__ T operator()(T s00, T s01, T s10, T s11, bool edge) const {
        return vpass(hpass(s00, s01, edge), hpass(s10, s11, edge));
    }
};
template <> struct LinOps<uint16_t> : LinOpsF<uint16_t> { __device__ __forceinline__ LinOps(float fx) : LinOpsF<uint16_t>(fx) {} };
template <> struct LinOps<int16_t>  : LinOpsF<int16_t>  { __device__ __forceinline__ LinOps(float fx) : LinOpsF<int16_t>(fx) {} };
template <> struct LinOps<float>    : LinOpsF<float>    { __device__ __forceinline__ LinOps(float fx) : LinOpsF<float>(fx) {} };
// 64F: HResizeLinear<double,double,float> / VResizeLinear<double,double,float,Cast<double,double>>: float coefficients, double arithmetic
template <> struct LinOps<double> {
    typedef double HT;
    double a0, a1, b0, b1;
    __device__ __forceinline__ LinOps(float fx) : a0((double)fsub(1.f, fx)), a1((double)fx), b0(0.0), b1(0.0) {}
    static __device__ __forceinline__ int pack_row(float) { return 0; }
    __device__ __forceinline__ void set_row(float fy, int) { b0 = (double)fsub(1.f, fy); b1 = (double)fy; }
    __device__ __forceinline__ double hpass(double s0, double s1, bool edge) const {
        return edge ? s0 : __dadd_rn(__dmul_rn(s0, a0), __dmul_rn(s1, a1));
    }
    __device__ __forceinline__ double vpass(double h0, double h1) const { return __dadd_rn(__dmul_rn(h0, b0), __dmul_rn(h1, b1)); }
    __device__ __forceinline__ double operator()(double s00, double s01, double s10, double s11, bool edge) const {
        return vpass(hpass(s00, s01, edge), hpass(s10, s11, edge));
    }
};

// Tile = LT_W columns x LT_H rows, one thread per column.  The column terms (source index, both horizontal
// coefficients, the x-border flag) are evaluated ONCE per thread and stay in registers; the row terms of the tile
// (two clipped source rows + vertical coefficients, :2211,:3955-3996) are evaluated once per tile into shared memory.
// The per-pixel work is then only the taps, the blend and the store -- the first version re-derived both axes per
// pixel in fp64 and was instruction-issue bound (profiles/r1_step2_ops_ncu_full.csv).
constexpr int LT_W = 128;     // LT_H (rows per thread) is a template parameter: fewer rows for small images, to keep the GPU full

struct LinRow { int r0, r1; float fy; int ib; };

struct LinCol { int sx; float fx; };

// CTA = LT_W columns x RG row groups; a thread owns one column and LT_H consecutive rows.  The column terms are evaluated by
// the first row group only and shared through shared memory: with one row group per CTA the per-thread column set-up (fp64,
// ~70 instructions) was a third of all instructions of the 32FC1 upscale (ncu: 55 instructions per pixel at 8 rows per
// thread), and making the threads longer instead (16 rows) ran slower -- these kernels live on many short threads.
template <typename T, int CN, int LT_H, int RG>
__global__ void __launch_bounds__(LT_W * RG) resize_linear_kernel(const ResizeArgs a)
{
    __shared__ LinRow rows_all[LT_H * RG];
    __shared__ LinCol cols[LT_W];
    const int tid = threadIdx.y * LT_W + threadIdx.x;
    const int x = blockIdx.x * LT_W + threadIdx.x, y0 = (blockIdx.y * RG + threadIdx.y) * LT_H;
    const int W = a.src.w, H = a.src.h;
    if (tid < LT_H * RG) {
        int sy; float fy;
        lin_coord(blockIdx.y * RG * LT_H + tid, a.scale_y, a.inv_scale_y, a.area_mode, sy, fy);
        rows_all[tid] = LinRow{clipi(sy, 0, H), clipi(sy + 1, 0, H), fy, LinOps<T>::pack_row(fy)};   // unmodified fy: y gets no zeroing
    }
    if (threadIdx.y == RG - 1) {                                      // (the last row group: its warps do not also build the row table)
        int sx; float fx;
        lin_coord(x, a.scale_x, a.inv_scale_x, a.area_mode, sx, fx);
        if (sx < 0) { fx = 0; sx = 0; }                               // :3918-3923
        if (sx >= W - 1) { fx = 0; sx = W - 1; }                      // :3928-3929 (sx + 1 >= W marks the x border below)
        cols[threadIdx.x] = LinCol{sx, fx};
    }
    __syncthreads();
    if (x >= a.dst.w || y0 >= a.dst.h) return;
    const LinRow* rows = rows_all + threadIdx.y * LT_H;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride;
    const LinCol col = cols[threadIdx.x];
    const int sx = col.sx; const float fx = col.fx;
    const bool edge = sx + 1 >= W;                                    // dx >= xmax (:3925-3927; sx is monotone in dx)
    const int o1 = edge ? 0 : CN;
    const uint8_t* scol = src + (long long)sx * CN * sizeof(T);
    uint8_t* drow = dst + (long long)x * CN * sizeof(T) + (long long)y0 * a.dst.step;
    const int ny = min(LT_H, a.dst.h - y0);
    const int sstep = a.src.step, dstep = a.dst.step;
    LinOps<T> ops(fx);
    // Separable path (the reference's own order: HResize per source row, then VResize, resize.cpp:1877-1989, :2166-2266).
    // When consecutive destination rows share source rows (vertical scale < 2: every upscale), a thread first evaluates the
    // horizontal result of EVERY source row the tile touches -- all their tap loads in flight together -- into its own slots
    // of shared memory (no barrier: nobody else reads them), then runs the vertical pass from there.  A 2x upscale then
    // loads one tap pair per two destination pixels instead of four taps per pixel.
    typedef typename LinOps<T>::HT HT;
    constexpr int SRC_CAP = LT_H + 2;
    extern __shared__ __align__(16) uint8_t lin_smem[];
    const int rmin = rows[0].r0, nsrc = rows[ny - 1].r1 - rmin + 1;
    if (a.lin_separable && nsrc <= SRC_CAP && nsrc < 2 * ny) {
        constexpr int NTH = LT_W * RG;
        HT* hb = reinterpret_cast<HT*>(lin_smem) + tid;                           // [SRC_CAP][CN][threads], this thread's slots
        const uint8_t* sp = scol + (long long)rmin * sstep;
#pragma unroll 4
        for (int r = 0; r < nsrc; r++) {
            const T* t0 = reinterpret_cast<const T*>(sp + (long long)r * sstep);
#pragma unroll
            for (int k = 0; k < CN; k++) hb[(r * CN + k) * NTH] = ops.hpass(__ldg(t0 + k), __ldg(t0 + k + o1), edge);
        }
        for (int j = 0; j < ny; j++) {
            const LinRow rw = rows[j];
            ops.set_row(rw.fy, rw.ib);
            T* D = reinterpret_cast<T*>(drow);
#pragma unroll
            for (int k = 0; k < CN; k++)
                store_out(&D[k], ops.vpass(hb[((rw.r0 - rmin) * CN + k) * NTH], hb[((rw.r1 - rmin) * CN + k) * NTH]), 0);
            drow += dstep;
        }
        return;
    }
    // rows in batches of RB (1-2 channels): all taps of a batch are loaded (read-only path) before the first store, so
    // RB x 4 x CN loads are in flight per thread.  The first version loaded, blended and stored one row at a time; the stores may alias
    // the source as far as the compiler knows, so nothing was hoisted and ncu showed 8.7 warps per issue slot waiting
    // on the long scoreboard at 21 % DRAM throughput (profiles/r1_notes.md).
    constexpr int RB = CN <= 2 ? (RG > 1 ? 2 : (4 < LT_H ? 4 : LT_H)) : 1;          // measured: batching helps 1-2 channels (32FC1 18.4 -> 16.5 us), costs 8UC3 / 8UC4 10-15 %
    static_assert(LT_H % RB == 0, "row batches");
#pragma unroll
    for (int rb = 0; rb < LT_H; rb += RB) {
        if (rb >= ny) break;
        T v[RB][4][CN];
#pragma unroll
        for (int j = 0; j < RB; j++) {
            if (rb + j < ny) {
                const LinRow rw = rows[rb + j];
                const T* r0 = reinterpret_cast<const T*>(scol + (long long)rw.r0 * sstep);
                const T* r1 = reinterpret_cast<const T*>(scol + (long long)rw.r1 * sstep);
#pragma unroll
                for (int k = 0; k < CN; k++) {
                    v[j][0][k] = __ldg(r0 + k); v[j][1][k] = __ldg(r0 + k + o1);
                    v[j][2][k] = __ldg(r1 + k); v[j][3][k] = __ldg(r1 + k + o1);
                }
            }
        }
#pragma unroll
        for (int j = 0; j < RB; j++) {
            if (rb + j < ny) {
                const LinRow rw = rows[rb + j];
                ops.set_row(rw.fy, rw.ib);
                T* D = reinterpret_cast<T*>(drow);
#pragma unroll
                for (int k = 0; k < CN; k++) store_out(&D[k], ops(v[j][0][k], v[j][1][k], v[j][2][k], v[j][3][k], edge), 0);
                drow += dstep;
            }
        }
    }
}

// 32FC1 (the 2x upscale of BASELINE config 2 and every other single-channel float resize): one thread owns FOUR adjacent
// destination columns and LT_H rows and writes 16-byte vectors.  The generic column kernel above spends 53 instructions per
// pixel on this case (ncu r2_resize: 24 in the vertical loop, 14 per source row in the horizontal one, all scalar); here the
// horizontal results of the two current source rows stay in registers and are re-used while consecutive destination rows
// keep them (every branch depends on the row table only, so a warp never diverges): ~9 instructions per pixel.
// Arithmetic = HResizeLinear / VResizeLinear<float> (resize.cpp:1877-1959): h = S[sx]*(1-fx) + S[sx+1]*fx, or S[sx] itself
// at the right border; D = h0*(1-fy) + h1*fy, every product and sum rounded separately.
constexpr int LV_W = 128;                         // threads per row group: 512 destination columns per CTA

template <int LT_H, int RG>
__global__ void __launch_bounds__(LV_W * RG) resize_linear_f32c1_x4_kernel(const ResizeArgs a)
{
    __shared__ LinRow rows_all[LT_H * RG];
    __shared__ LinCol cols[LV_W * 4];
    const int tid = threadIdx.y * LV_W + threadIdx.x;
    const int xb = blockIdx.x * (LV_W * 4), y0 = (blockIdx.y * RG + threadIdx.y) * LT_H;
    const int W = a.src.w, H = a.src.h;
    if (tid < LT_H * RG) {
        int sy; float fy;
        lin_coord(blockIdx.y * RG * LT_H + tid, a.scale_y, a.inv_scale_y, a.area_mode, sy, fy);
        rows_all[tid] = LinRow{clipi(sy, 0, H), clipi(sy + 1, 0, H), fy, 0};
    }
    for (int c = tid; c < LV_W * 4; c += LV_W * RG) {
        int sx; float fx;
        lin_coord(xb + c, a.scale_x, a.inv_scale_x, a.area_mode, sx, fx);
        if (sx < 0) { fx = 0; sx = 0; }                               // resize.cpp:3918-3923
        if (sx >= W - 1) { fx = 0; sx = W - 1; }                      // :3928-3929
        cols[c] = LinCol{sx, fx};
    }
    __syncthreads();
    const int x = xb + threadIdx.x * 4;
    if (x >= a.dst.w || y0 >= a.dst.h) return;
    const LinRow* rows = rows_all + threadIdx.y * LT_H;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    int o0[4], o1[4]; float a0[4], a1[4]; bool edge[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const LinCol c = cols[threadIdx.x * 4 + k];
        edge[k] = c.sx + 1 >= W;                                      // dx >= xmax (:3925-3927)
        o0[k] = c.sx; o1[k] = c.sx + (edge[k] ? 0 : 1);
        a1[k] = c.fx; a0[k] = fsub(1.f, c.fx);
    }
    float4* D = reinterpret_cast<float4*>(a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y0 * a.dst.step) + (x >> 2);
    const int ny = min(LT_H, a.dst.h - y0);
    float h0[4], h1[4];
    int i0 = -1, i1 = -1;
    auto hrow = [&](float* h, int r) {
        const float* S = reinterpret_cast<const float*>(src + (long long)r * a.src.step);
        float s0[4], s1[4];
#pragma unroll
        for (int k = 0; k < 4; k++) { s0[k] = __ldg(S + o0[k]); s1[k] = __ldg(S + o1[k]); }
#pragma unroll
        for (int k = 0; k < 4; k++) h[k] = edge[k] ? s0[k] : fadd(fmul(s0[k], a0[k]), fmul(s1[k], a1[k]));
    };
    for (int j = 0; j < ny; j++) {
        const LinRow rw = rows[j];
        if (rw.r0 == i1) {
#pragma unroll
            for (int k = 0; k < 4; k++) h0[k] = h1[k];
            i0 = i1;
        } else if (rw.r0 != i0) { hrow(h0, rw.r0); i0 = rw.r0; }
        if (rw.r1 == i0) {
#pragma unroll
            for (int k = 0; k < 4; k++) h1[k] = h0[k];
            i1 = i0;
        } else if (rw.r1 != i1) { hrow(h1, rw.r1); i1 = rw.r1; }
        const float b1 = rw.fy, b0 = fsub(1.f, rw.fy);
        float4 v;
        v.x = fadd(fmul(h0[0], b0), fmul(h1[0], b1)); v.y = fadd(fmul(h0[1], b0), fmul(h1[1], b1));
        v.z = fadd(fmul(h0[2], b0), fmul(h1[2], b1)); v.w = fadd(fmul(h0[3], b0), fmul(h1[3], b1));
        __stcs(D, v);
        D = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(D) + a.dst.step);
    }
}

static bool linear_f32c1_x4_applicable(const ResizeArgs& a)
{
    return !a.force_generic && (a.dst.w & 3) == 0 && ((uintptr_t)a.dst.p & 15) == 0 && (a.dst.step & 15) == 0 && (a.dst_frame_stride & 15) == 0 &&
           ((uintptr_t)a.src.p & 3) == 0 && (a.src.step & 3) == 0 && (a.src_frame_stride & 3) == 0;
}

static cudaError_t launch_linear_f32c1_x4(const ResizeArgs& a, int n_frames, cudaStream_t s)
{
    const long long px = (long long)a.dst.w * a.dst.h * n_frames;
    if (px >= 4000000) {
        dim3 g((a.dst.w + LV_W * 4 - 1) / (LV_W * 4), (a.dst.h + 31) / 32, n_frames);
        resize_linear_f32c1_x4_kernel<8, 4><<<g, dim3(LV_W, 4), 0, s>>>(a);
    } else {
        dim3 g((a.dst.w + LV_W * 4 - 1) / (LV_W * 4), (a.dst.h + 7) / 8, n_frames);
        resize_linear_f32c1_x4_kernel<2, 4><<<g, dim3(LV_W, 4), 0, s>>>(a);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------- AREA FAST ------
// integer scales (resize.cpp:2919-3059).  dwidth1 = (W/scale_x)*cn full-window elements per row;
// 2x2 with cn in {1,3,4}: 8U/16U/16S (a+b+c+d+2)>>2 (:2950-2980); 32F cn in {1,4}: vector body
// ((a+b)+(c+d))*0.25f for the first (dwidth1 & ~3) elements (:2857-2908, 4 lanes at the SSE3 baseline);
// otherwise sum in window raster order, unrolled by 4 (:3015-3025), times 1.f/area; ragged right/bottom
// edge: sum/count over the in-bounds part (:3028-3050).
template <typename T> struct AreaSum;       // WT accumulation of the generic body (areafast_tab :3745-3755)
template <> struct AreaSum<uint8_t> {       // WT = int
    int s = 0;
    __device__ __forceinline__ void add4(int a, int b, int c, int d) { s += a + b + c + d; }
    __device__ __forceinline__ void add1(int a) { s += a; }
    __device__ __forceinline__ float value() const { return (float)s; }
};
template <typename T> struct AreaSumIntGroups {   // WT = float, 4-tuples added in int (integer promotion)
    float s = 0.f;
    __device__ __forceinline__ void add4(int a, int b, int c, int d) { s = fadd(s, (float)(a + b + c + d)); }
    __device__ __forceinline__ void add1(int a) { s = fadd(s, (float)a); }
    __device__ __forceinline__ float value() const { return s; }
};
template <> struct AreaSum<uint16_t> : AreaSumIntGroups<uint16_t> {};
template <> struct AreaSum<int16_t>  : AreaSumIntGroups<int16_t> {};
template <> struct AreaSum<float> {
    float s = 0.f;
    __device__ __forceinline__ void add4(float a, float b, float c, float d) { s = fadd(s, fadd(fadd(fadd(a, b), c), d)); }
    __device__ __forceinline__ void add1(float a) { s = fadd(s, a); }
    __device__ __forceinline__ float value() const { return s; }
};

template <typename T, int CN>
__global__ void __launch_bounds__(RT_W * RT_H) resize_area_fast_kernel(const ResizeArgs a)
{
    const int x = blockIdx.x * RT_W + threadIdx.x, y = blockIdx.y * RT_H + threadIdx.y;
    if (x >= a.dst.w || y >= a.dst.h) return;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride;
    const int W = a.src.w, H = a.src.h, kx = a.iscale_x, ky = a.iscale_y;
    T* D = reinterpret_cast<T*>(dst + (long long)y * a.dst.step) + (long long)x * CN;
    const int sy0 = y * ky;
    if (sy0 >= H) {                                                    // :2998-3003
#pragma unroll
        for (int k = 0; k < CN; k++) D[k] = (T)0;
        return;
    }
    const int full_px = (sy0 + ky <= H) ? W / kx : 0;                  // w = dwidth1 or 0 (:2995)
    const long long sstep = a.src.step;
    const T* S = reinterpret_cast<const T*>(src + (long long)sy0 * sstep) + (long long)x * kx * CN;
    if (x < full_px) {
        const bool is2x2 = kx == 2 && ky == 2 && (CN == 1 || CN == 3 || CN == 4);
        const T* S1 = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(S) + sstep);
        if constexpr (!std::is_same<T, float>::value) {
            if (is2x2) {
#pragma unroll
                for (int k = 0; k < CN; k++) store_out(&D[k], (T)(((int)S[k] + (int)S[k + CN] + (int)S1[k] + (int)S1[k + CN] + 2) >> 2), a.keep_in_l2);
                return;
            }
        } else {
            if (kx == 2 && ky == 2 && (CN == 1 || CN == 4) && (x * CN < ((full_px * CN) & ~3))) {
#pragma unroll
                for (int k = 0; k < CN; k++) store_out(&D[k], fmul(fadd(fadd(S[k], S[k + CN]), fadd(S1[k], S1[k + CN])), 0.25f), a.keep_in_l2);
                return;
            }
        }
        const int area = kx * ky;
        const float scale = 1.f / (float)area;
#pragma unroll
        for (int c = 0; c < CN; c++) {
            AreaSum<T> acc;
            int k = 0;
            // ofs[k] = (k / kx) rows down, (k % kx) pixels right (:3832-3834)
            auto at = [&](int kk) { return reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(S) + (long long)(kk / kx) * sstep)[(kk % kx) * CN + c]; };
            for (; k <= area - 4; k += 4) acc.add4(at(k), at(k + 1), at(k + 2), at(k + 3));
            for (; k < area; k++) acc.add1(at(k));
            store_out(&D[c], Elem<T>::from_f(fmul(acc.value(), scale)), a.keep_in_l2);
        }
        return;
    }
    // ragged edge (:3028-3050): WT sum over the in-bounds part, divided by the tap count
#pragma unroll
    for (int c = 0; c < CN; c++) {
        const int sx0 = x * kx * CN + c;
        int count = 0;
        float fsum = 0.f; int isum = 0;
        for (int sy = 0; sy < ky; sy++) {
            if (sy0 + sy >= H) break;
            const T* R = reinterpret_cast<const T*>(src + (long long)(sy0 + sy) * sstep);
            for (int sxo = 0; sxo < kx * CN; sxo += CN) {
                if (sx0 + sxo >= W * CN) break;
                if constexpr (std::is_same<T, uint8_t>::value) isum += R[sx0 + sxo]; else fsum = fadd(fsum, (float)R[sx0 + sxo]);
                count++;
            }
        }
        float num = std::is_same<T, uint8_t>::value ? (float)isum : fsum;
        store_out(&D[c], Elem<T>::from_f(__fdiv_rn(num, (float)count)), a.keep_in_l2);
    }
}

// ------------------------------------------------------------------------------- AREA -----------
// fractional downscale (resize.cpp:3183-3304, tables :3334-3373).  Each tile evaluates the DecimateAlpha
// entries of its 32 columns and 8 rows in fp64 into shared memory (first source index, count, weights:
// the entries of one dst index are consecutive source indices), then every thread accumulates in the
// reference's order: per source row buf = sum_x S*alpha (in table order), sum = beta*buf for the first
// source row of a dst row, sum += beta*buf afterwards.  WT = float for every depth (area_tab :3757-3762).
struct AreaAxis { int first, n; };

__device__ __forceinline__ void area_axis_setup(int d, int ssize, double scale, AreaAxis& ax, float* w)
{
    double fsx1 = __dmul_rn((double)d, scale);
    double fsx2 = __dadd_rn(fsx1, scale);
    double cell = fmin(scale, __dsub_rn((double)ssize, fsx1));
    int sx1 = ceil_d(fsx1), sx2 = floor_d(fsx2);
    sx2 = min(sx2, ssize - 1);
    sx1 = min(sx1, sx2);
    int n = 0;
    ax.first = sx1;
    if (__dsub_rn((double)sx1, fsx1) > 1e-3) {
        ax.first = sx1 - 1;
        w[n++] = __double2float_rn(__ddiv_rn(__dsub_rn((double)sx1, fsx1), cell));
    }
    const float wfull = __double2float_rn(__ddiv_rn(1.0, cell));
    for (int sx = sx1; sx < sx2; sx++) w[n++] = wfull;
    if (__dsub_rn(fsx2, (double)sx2) > 1e-3)
        w[n++] = __double2float_rn(__ddiv_rn(fmin(fmin(__dsub_rn(fsx2, (double)sx2), 1.0), cell), cell));
    ax.n = n;
}

// One thread per destination COLUMN, AT_H destination rows per tile.  The source rows a tile touches are consecutive (the
// windows of consecutive destination rows abut or share one row), so the thread walks them once: per source row ONE horizontal
// sum buf = sum_x S*alpha in table order (hresize, resize.cpp:3216-3262), applied to the one or two destination rows that
// contain the row (vresize, :3264-3295: sum = beta*buf for the first row of a destination row, sum += beta*buf after).  The
// first version ran one thread per destination PIXEL and redid the horizontal sum of a shared row for both of its users
// (35.8 us at 4K -> 1600x900 8UC3, ncu: issue bound).
// KX > 0: at most KX taps per row (compile-time bound, alpha in registers); KX = 0 keeps run-time loops (any scale).
constexpr int AT_W = 128;

template <typename T, int CN, int KX, int AT_H, int RG>
__global__ void __launch_bounds__(AT_W * RG) resize_area_kernel(const ResizeArgs a)
{
    extern __shared__ float smem_w[];                  // [AT_W][kmax_x] then [AT_H * RG][kmax_y]
    __shared__ AreaAxis axy_all[AT_H * RG], axx[AT_W];
    const int tid = threadIdx.y * AT_W + threadIdx.x;
    const int x = blockIdx.x * AT_W + threadIdx.x, y0 = (blockIdx.y * RG + threadIdx.y) * AT_H;
    float* wx = smem_w + threadIdx.x * a.kmax_x;
    float* wy_all = smem_w + AT_W * a.kmax_x;
    // column tables by the last row group only (three fp64 divisions each), row tables by the first threads; everybody reads them
    if (threadIdx.y == RG - 1) {
        AreaAxis c{0, 0};
        if (x < a.dst.w) area_axis_setup(x, a.src.w, a.scale_x, c, wx);
        axx[threadIdx.x] = c;
    }
    if (tid < AT_H * RG && blockIdx.y * RG * AT_H + tid < a.dst.h)
        area_axis_setup(blockIdx.y * RG * AT_H + tid, a.src.h, a.scale_y, axy_all[tid], wy_all + tid * a.kmax_y);
    __syncthreads();
    if (x >= a.dst.w || y0 >= a.dst.h) return;
    const AreaAxis cx = axx[threadIdx.x];
    const AreaAxis* axy = axy_all + threadIdx.y * AT_H;
    const float* wy = wy_all + threadIdx.y * AT_H * a.kmax_y;
    const int nrows = min(AT_H, a.dst.h - y0);
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* drow = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y0 * a.dst.step + (long long)x * CN * sizeof(T);
    float al[KX > 0 ? KX : 1];
    if constexpr (KX > 0) {
#pragma unroll
        for (int i = 0; i < KX; i++) al[i] = i < cx.n ? wx[i] : 0.f;
    }
    const int s_first = axy[0].first, s_last = axy[nrows - 1].first + axy[nrows - 1].n - 1;
    const uint8_t* scol = src + (long long)cx.first * CN * sizeof(T);
    // 8U: the KX*CN tap bytes of a row are contiguous: fetch them as aligned words (a third of the load instructions and
    // L1 wavefronts of byte loads), shift the window into place with funnel shifts and turn byte k into a float with one
    // PRMT (into the mantissa of 2^23) and one FADD.  Only while the window stays inside the row pitch, so that it can never
    // leave the allocation.
    constexpr bool WORDS = KX > 0 && std::is_same<T, uint8_t>::value;
    constexpr int NB = (KX > 0 ? KX : 1) * CN, NW = (NB + 3) / 4 + 1;
    const bool words_ok = WORDS && (long long)cx.first * CN + NW * 4 <= (long long)a.src.step && (a.src.step & 3) == 0 &&
                          (((uintptr_t)src & 3) == 0);
    const int off = (cx.first * CN) & 3;                // rows are word aligned on this path
    int j = 0;                                         // destination row being accumulated
    AreaAxis cy = axy[0];
    const float* ay = wy;
    float sum[CN];
#pragma unroll
    for (int c = 0; c < CN; c++) sum[c] = 0.f;
    // (a queue of 2-4 source rows of tap words in flight, 16 rows per thread, and 4 row groups per CTA sharing the column
    // tables were all measured slower than this plain form: 36-37 us vs 28.5 us at 4K -> 1600x900 8UC3)
    for (int sr = s_first; sr <= s_last; sr++) {
        const T* S = reinterpret_cast<const T*>(scol + (long long)sr * a.src.step);
        float buf[CN];
#pragma unroll
        for (int c = 0; c < CN; c++) buf[c] = 0.f;
        if (WORDS && words_ok) {
            const uint32_t* wp = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(S) - off);
            uint32_t w[NW];
#pragma unroll
            for (int k = 0; k < NW; k++) w[k] = __ldg(wp + k);
#pragma unroll
            for (int k = 0; k < NW - 1; k++) w[k] = __funnelshift_r(w[k], w[k + 1], off * 8);
#pragma unroll
            for (int i = 0; i < (KX > 0 ? KX : 1); i++) {
                if (i < cx.n) {
#pragma unroll
                    for (int c = 0; c < CN; c++) {
                        const int k = i * CN + c;
                        const float v = __fsub_rn(__uint_as_float(__byte_perm(w[k >> 2], 0x4B000000u, 0x7650 + (k & 3))), 8388608.0f);
                        buf[c] = fadd(buf[c], fmul(v, al[i]));
                    }
                }
            }
        } else if constexpr (KX > 0) {
            T v[KX][CN];
#pragma unroll
            for (int i = 0; i < KX; i++)
#pragma unroll
                for (int c = 0; c < CN; c++) v[i][c] = i < cx.n ? __ldg(S + i * CN + c) : (T)0;
#pragma unroll
            for (int i = 0; i < KX; i++) {
                if (i < cx.n) {
#pragma unroll
                    for (int c = 0; c < CN; c++) buf[c] = fadd(buf[c], fmul(Elem<T>::to_f(v[i][c]), al[i]));
                }
            }
        } else {
            for (int i = 0; i < cx.n; i++) {
                const float alpha = wx[i];
#pragma unroll
                for (int c = 0; c < CN; c++) buf[c] = fadd(buf[c], fmul(Elem<T>::to_f(__ldg(S + i * CN + c)), alpha));
            }
        }
        // every destination row whose window contains source row sr (uniform over the CTA: it depends on the row tables only)
        while (j < nrows && cy.first <= sr) {
            const int t = sr - cy.first;
            const float beta = ay[t];
#pragma unroll
            for (int c = 0; c < CN; c++) sum[c] = (t == 0) ? fmul(beta, buf[c]) : fadd(sum[c], fmul(beta, buf[c]));
            if (t < cy.n - 1) break;                     // row j continues with the next source row
            T* D = reinterpret_cast<T*>(drow);
#pragma unroll
            for (int c = 0; c < CN; c++) store_out(&D[c], Elem<T>::from_f(sum[c]), a.keep_in_l2);
            drow += a.dst.step;
            j++;
            if (j < nrows) { cy = axy[j]; ay = wy + j * a.kmax_y; }
        }
    }
}

// ------------------------------------------------------------------------------- LINEAR_EXACT ---
// INTER_LINEAR_EXACT, 8U (ufixedpoint16 = 8.8) and 16U (ufixedpoint32 = 16.16): resize_bitExact resize.cpp:895-946,
// interpolationLinear::getCoeffs :789-828, hlineResize :74-109, vlineResize / vlineSet :708-786, fixedpoint.inl.hpp:224-374.
// The reference derives coefficients with softdouble (= IEEE double, op by op): fval = scale*(x+0.5)-0.5, ival = floor,
// c1 = cvRound((fval-ival)*2^shift), c0 = 2^shift - c1.  Columns left of the image copy the first pixel (<<shift), columns at
// or beyond the last source pixel copy the last one; rows likewise (vlineSet: rounded (v + half) >> shift).
struct ExactAxis { int ofs; unsigned c0, c1; int where; };       // where: -1 before the image, 0 inside, +1 at/after the end

__device__ __forceinline__ ExactAxis exact_axis(int val, double scale, int maxsize, int shift)
{
    ExactAxis e; e.ofs = 0; e.c0 = e.c1 = 0;
    const double fval = __dsub_rn(__dmul_rn(scale, __dadd_rn((double)val, 0.5)), 0.5);
    const int ival = floor_d(fval);
    if (ival >= 0 && maxsize > 1) {
        if (ival < maxsize - 1) {
            e.where = 0; e.ofs = ival;
            const double frac = __dsub_rn(fval, (double)ival);
            e.c1 = frac < 0 ? 0u : (unsigned)rne_d(__dmul_rn(frac, (double)(1 << shift)));
            const unsigned one = 1u << shift;
            e.c0 = one > e.c1 ? one - e.c1 : 0u;
        } else { e.where = 1; e.ofs = maxsize - 1; }
    } else e.where = -1;
    return e;
}

struct ExactRow { int r0, r1; unsigned c0, c1; int where, pad0, pad1, pad2; };

template <typename T, int CN, int LT_H>
__global__ void __launch_bounds__(LT_W) resize_linear_exact_kernel(const ResizeArgs a)
{
    constexpr int SHIFT = (sizeof(T) == 1 && !std::is_signed<T>::value) ? 8 : 16;       // 8U: ufixedpoint16; 16U, 8S, 16S: [u]fixedpoint32
    constexpr unsigned long long CAP = sizeof(T) == 1 ? 0xFFFFull : 0xFFFFFFFFull;
    __shared__ ExactRow rows[LT_H];
    const int x = blockIdx.x * LT_W + threadIdx.x, y0 = blockIdx.y * LT_H;
    const int W = a.src.w, H = a.src.h;
    if (threadIdx.x < LT_H) {
        const ExactAxis e = exact_axis(y0 + threadIdx.x, a.scale_y, H, SHIFT);
        const int r0 = e.where < 0 ? 0 : e.where > 0 ? H - 1 : e.ofs;
        rows[threadIdx.x] = ExactRow{r0, e.where == 0 ? r0 + 1 : r0, e.c0, e.c1, e.where, 0, 0, 0};
    }
    __syncthreads();
    if (x >= a.dst.w) return;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* drow = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)x * CN * sizeof(T) + (long long)y0 * a.dst.step;
    const ExactAxis ex = exact_axis(x, a.scale_x, W, SHIFT);
    const int sx = ex.where < 0 ? 0 : ex.where > 0 ? W - 1 : ex.ofs;
    const int o1 = ex.where == 0 ? CN : 0;
    // edge columns: value << shift == c0 * p with c0 = 2^shift, c1 = 0
    const unsigned long long cx0 = ex.where == 0 ? ex.c0 : (1ull << SHIFT), cx1 = ex.where == 0 ? ex.c1 : 0ull;
    const uint8_t* scol = src + (long long)sx * CN * sizeof(T);
    const int ny = min(LT_H, a.dst.h - y0);
    const int sstep = a.src.step, dstep = a.dst.step;
#pragma unroll
    for (int r = 0; r < LT_H; r++) {
        if (r >= ny) break;
        const ExactRow rw = rows[r];
        const T* p0 = reinterpret_cast<const T*>(scol + (long long)rw.r0 * sstep);
        const T* p1 = reinterpret_cast<const T*>(scol + (long long)rw.r1 * sstep);
        T* D = reinterpret_cast<T*>(drow);
        if constexpr (std::is_signed<T>::value) {
            // CV_8S and CV_16S both run in fixedpoint32 (int32_t, 16.16; fixedtype<>, resize.cpp:66-71), vertical products in fixedpoint64
            // (fixedpoint.inl.hpp:169-221): saturating products / sums, the cast back adds half an LSB in the unsigned type of the
            // same width and shifts arithmetically
            constexpr long long FLO = -2147483648LL, FHI = 2147483647LL;
            auto sat = [](long long v, long long lo, long long hi) { return v < lo ? lo : v > hi ? hi : v; };
#pragma unroll
            for (int k = 0; k < CN; k++) {
                const long long h0 = sat(sat((long long)cx0 * p0[k], FLO, FHI) + sat((long long)cx1 * p0[k + o1], FLO, FHI), FLO, FHI);
                const long long h1 = sat(sat((long long)cx0 * p1[k], FLO, FHI) + sat((long long)cx1 * p1[k + o1], FLO, FHI), FLO, FHI);
                long long v;
                if (rw.where != 0) v = (long long)((int)((unsigned)h0 + 32768u) >> 16);
                else {
                    // |h| <= 2^31, c <= 2^16, c0 + c1 = 2^16: the sum of the two products stays inside int64
                    const long long acc = h0 * (long long)rw.c0 + h1 * (long long)rw.c1;
                    v = (long long)((unsigned long long)acc + 0x80000000ull) >> 32;
                }
                D[k] = (T)sat(v, sizeof(T) == 1 ? -128 : -32768, sizeof(T) == 1 ? 127 : 32767);
            }
        } else
#pragma unroll
        for (int k = 0; k < CN; k++) {
            // horizontal pass: saturating products and sum in FT (never saturate for c0+c1 = 2^shift; kept for fidelity)
            unsigned long long a0 = min(cx0 * p0[k], CAP), b0 = min(cx1 * p0[k + o1], CAP);
            unsigned long long a1 = min(cx0 * p1[k], CAP), b1 = min(cx1 * p1[k + o1], CAP);
            const unsigned long long h0 = min(a0 + b0, CAP), h1 = min(a1 + b1, CAP);
            unsigned long long v;
            if (rw.where != 0) v = (h0 + (1ull << (SHIFT - 1))) >> SHIFT;                           // vlineSet
            else v = (h0 * rw.c0 + h1 * rw.c1 + (1ull << (2 * SHIFT - 1))) >> (2 * SHIFT);         // vlineResize
            store_out(&D[k], (T)min(v, (unsigned long long)(sizeof(T) == 1 ? 255u : 65535u)), 0);
        }
        drow += dstep;
    }
}

// INTER_LINEAR_EXACT on CV_32S: fixedpoint64 (32.32) in both passes (fixedpoint.inl.hpp:14-99).  Its product works on magnitudes,
// (|a| * |b| + 2^31) >> 32 over 128 bits, saturated at 2^63 - 1 / -2^63, the sign applied afterwards (rounds half away from zero);
// sums saturate; the cast back is (int64)((uint64)v + 2^31) >> 32 saturated to int.  Coefficients: cvRound64(frac * 2^32).
__device__ __forceinline__ long long fp64_mul(long long a, long long b)
{
    const bool neg = (a < 0) != (b < 0);
    const unsigned long long ua = a < 0 ? 0ull - (unsigned long long)a : (unsigned long long)a;
    const unsigned long long ub = b < 0 ? 0ull - (unsigned long long)b : (unsigned long long)b;
    unsigned long long lo = ua * ub, hi = __umul64hi(ua, ub);
    const unsigned long long lo2 = lo + 0x80000000ull;
    hi += lo2 < lo;                                                    // carry of the rounding term
    if (hi >> 31) return neg ? (long long)0x8000000000000000ull : 0x7FFFFFFFFFFFFFFFll;   // (hi:lo2) >> 32 >= 2^63
    const unsigned long long r = (hi << 32) | (lo2 >> 32);
    return neg ? (long long)(0ull - r) : (long long)r;
}
__device__ __forceinline__ long long fp64_add(long long a, long long b)
{
    const long long res = (long long)((unsigned long long)a + (unsigned long long)b);
    return (((a ^ res) & (b ^ res)) < 0) ? (long long)~((unsigned long long)res & 0x8000000000000000ull) : res;
}
struct ExactAxis64 { int ofs; long long c0, c1; int where; };
__device__ __forceinline__ ExactAxis64 exact_axis64(int val, double scale, int maxsize)
{
    ExactAxis64 e; e.ofs = 0; e.c0 = e.c1 = 0;
    const double fval = __dsub_rn(__dmul_rn(scale, __dadd_rn((double)val, 0.5)), 0.5);
    const int ival = floor_d(fval);
    if (ival >= 0 && maxsize > 1) {
        if (ival < maxsize - 1) {
            e.where = 0; e.ofs = ival;
            e.c1 = __double2ll_rn(__dmul_rn(__dsub_rn(fval, (double)ival), 4294967296.0));
            e.c0 = 4294967296LL - e.c1;
        } else { e.where = 1; e.ofs = maxsize - 1; }
    } else e.where = -1;
    return e;
}

__global__ void __launch_bounds__(LT_W) resize_linear_exact_s32_kernel(const ResizeArgs a, int cn)
{
    const int o = blockIdx.x * LT_W + threadIdx.x, y = blockIdx.y;
    if (o >= a.dst.w * cn) return;
    const int x = o / cn, c = o - x * cn;
    const int W = a.src.w, H = a.src.h;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    int* D = reinterpret_cast<int*>(a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y * a.dst.step) + o;
    const ExactAxis64 ex = exact_axis64(x, a.scale_x, W), ey = exact_axis64(y, a.scale_y, H);
    const int r0 = ey.where < 0 ? 0 : ey.where > 0 ? H - 1 : ey.ofs, r1 = ey.where == 0 ? r0 + 1 : r0;
    const int sx = ex.where < 0 ? 0 : ex.where > 0 ? W - 1 : ex.ofs;
    long long hh[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int* R = reinterpret_cast<const int*>(src + (long long)(k ? r1 : r0) * a.src.step) + sx * cn + c;
        const long long p0 = (long long)((unsigned long long)(long long)R[0] << 32);
        if (ex.where != 0) hh[k] = p0;
        else hh[k] = fp64_add(fp64_mul(ex.c0, p0), fp64_mul(ex.c1, (long long)((unsigned long long)(long long)R[cn] << 32)));
    }
    const long long acc = ey.where != 0 ? hh[0] : fp64_add(fp64_mul(hh[0], ey.c0), fp64_mul(hh[1], ey.c1));
    const long long v = (long long)((unsigned long long)acc + 0x80000000ull) >> 32;
    *D = (int)(v < -2147483648LL ? -2147483648LL : v > 2147483647LL ? 2147483647LL : v);
}

// ------------------------------------------------------------------------------- INTER_AREA, CV_64F ---
// Coverage kernels (one thread per destination element, run-time channel count).  Integer scales: resizeAreaFast_<double, double,
// ResizeAreaFastNoVec> -- the window sum in double, four values at a time, times the FLOAT 1.f/area; a ragged edge divides
// (float)sum by the number of taps (resize.cpp:3010-3050).  Other scales: resizeArea_<double, double> -- float table weights, double
// products and sums, per source row buf = sum_x S*alpha in table order, then sum = beta*buf / sum += beta*buf (:3183-3304).
__global__ void __launch_bounds__(256) resize_area_fast_f64_kernel(const ResizeArgs a, int cn)
{
    const int o = blockIdx.x * 256 + threadIdx.x, y = blockIdx.y;
    if (o >= a.dst.w * cn) return;
    const int x = o / cn, c = o - x * cn;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    double* D = reinterpret_cast<double*>(a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y * a.dst.step) + o;
    const int W = a.src.w, H = a.src.h, kx = a.iscale_x, ky = a.iscale_y, sy0 = y * ky;
    if (sy0 >= H) { *D = 0.0; return; }
    const int full_px = (sy0 + ky <= H) ? W / kx : 0;
    const long long sstep = a.src.step;
    if (x < full_px) {
        const uint8_t* S = src + (long long)sy0 * sstep + ((long long)x * kx * cn + c) * 8;
        const int area = kx * ky;
        auto at = [&](int kk) { return *reinterpret_cast<const double*>(S + (long long)(kk / kx) * sstep + (long long)((kk % kx) * cn) * 8); };
        double sum = 0.0;
        int k = 0;
        for (; k <= area - 4; k += 4) sum = __dadd_rn(sum, __dadd_rn(__dadd_rn(__dadd_rn(at(k), at(k + 1)), at(k + 2)), at(k + 3)));
        for (; k < area; k++) sum = __dadd_rn(sum, at(k));
        *D = __dmul_rn(sum, (double)(1.f / (float)area));
        return;
    }
    const int sx0 = x * kx * cn + c;
    int count = 0;
    double sum = 0.0;
    for (int sy = 0; sy < ky; sy++) {
        if (sy0 + sy >= H) break;
        const double* R = reinterpret_cast<const double*>(src + (long long)(sy0 + sy) * sstep);
        for (int sxo = 0; sxo < kx * cn; sxo += cn) {
            if (sx0 + sxo >= W * cn) break;
            sum = __dadd_rn(sum, R[sx0 + sxo]);
            count++;
        }
    }
    *D = (double)__fdiv_rn(__double2float_rn(sum), (float)count);
}

constexpr int A64_W = 128;
__global__ void __launch_bounds__(A64_W) resize_area_f64_kernel(const ResizeArgs a, int cn)
{
    extern __shared__ float smem_w[];                  // [A64_W][kmax_x] then [kmax_y]
    __shared__ AreaAxis axy;
    const int x = blockIdx.x * A64_W + threadIdx.x, y = blockIdx.y;
    float* wx = smem_w + threadIdx.x * a.kmax_x;
    float* wy = smem_w + A64_W * a.kmax_x;
    AreaAxis cx{0, 0};
    if (x < a.dst.w) area_axis_setup(x, a.src.w, a.scale_x, cx, wx);
    if (threadIdx.x == 0) area_axis_setup(y, a.src.h, a.scale_y, axy, wy);
    __syncthreads();
    if (x >= a.dst.w) return;
    const AreaAxis cy = axy;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    double* D = reinterpret_cast<double*>(a.dst.p + (long long)blockIdx.z * a.dst_frame_stride + (long long)y * a.dst.step) + (long long)x * cn;
    for (int c = 0; c < cn; c++) {
        double sum = 0.0;
        for (int t = 0; t < cy.n; t++) {
            const double* S = reinterpret_cast<const double*>(src + (long long)(cy.first + t) * a.src.step) + (long long)cx.first * cn + c;
            double buf = 0.0;
            for (int i = 0; i < cx.n; i++) buf = __dadd_rn(buf, __dmul_rn(S[i * cn], (double)wx[i]));
            const double beta = (double)wy[t];
            sum = t == 0 ? __dmul_rn(beta, buf) : __dadd_rn(sum, __dmul_rn(beta, buf));
        }
        D[c] = sum;
    }
}

template <typename T>
static cudaError_t launch_exact(const ResizeArgs& a, int cn, int n_frames, cudaStream_t s)
{
    const long long px = (long long)a.dst.w * a.dst.h * n_frames;
    const int lh = px / 8 >= 600000 ? 8 : 2;
    dim3 lgrid((a.dst.w + LT_W - 1) / LT_W, (a.dst.h + lh - 1) / lh, n_frames);
#define B200_EXACT(CNV) (lh == 8 ? resize_linear_exact_kernel<T, CNV, 8><<<lgrid, LT_W, 0, s>>>(a) : resize_linear_exact_kernel<T, CNV, 2><<<lgrid, LT_W, 0, s>>>(a))
    switch (cn) {
    case 1: B200_EXACT(1); break;
    case 2: B200_EXACT(2); break;
    case 3: B200_EXACT(3); break;
    case 4: B200_EXACT(4); break;
    default: return cudaErrorInvalidValue;
    }
#undef B200_EXACT
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------- dispatch -------
// ------------------------------------------------------------------------------- CUBIC ----------
// resize.cpp:3884-3999 (coefficients: interpolateCubic :964-972 in fp32, 8U as int16 * 2^11), HResizeCubic :1993-2041
// (taps sx-1..sx+2, clamped to the row per channel), rows sy-1..sy+2 clamped (:2205-2211), VResizeCubic :2044-2062.
// The reference's vertical pass has an SSE body for the first (width & ~7) elements of a row (width & ~3 for 32F)
// and a scalar tail, and the two add in different orders -- body: S0*b0 + (S1*b1 + (S2*b2 + S3*b3)), for 8U in fp32
// with beta * 2^-22 and round-to-nearest-even (:1410-1443); tail: ((S0*b0 + S1*b1) + S2*b2) + S3*b3, for 8U in int with
// (sum + 2^21) >> 22.  One thread per row element (pixel x channel) so that the split falls on a thread boundary.
__device__ __forceinline__ void cubic_coeffs(float x, float* c)
{
    const float A = -0.75f;
    const float x1 = fadd(x, 1.f), xr = fsub(1.f, x);
    c[0] = fsub(fmul(fadd(fmul(fsub(fmul(A, x1), 5 * A), x1), 8 * A), x1), 4 * A);
    c[1] = fadd(fmul(fmul(fsub(fmul(A + 2, x), A + 3), x), x), 1.f);
    c[2] = fadd(fmul(fmul(fsub(fmul(A + 2, xr), A + 3), xr), xr), 1.f);
    c[3] = fsub(fsub(fsub(1.f, c[0]), c[1]), c[2]);
}

// One thread per destination column (all channels), CR rows per thread: the column terms (source index, four
// coefficients, clamped tap offsets) are computed once per thread, the row terms (clamped source rows, four vertical
// coefficients) once per block into shared memory.  The first version spent one thread per row ELEMENT and re-derived
// both axes in every thread (45 us for 4K -> 1600x900 8UC3).
constexpr int CUB_W = 128, CUB_R = 4;
struct CubRow { int r[4]; float b[4]; };

template <typename T, int CN>
__global__ void __launch_bounds__(CUB_W)
resize_cubic_kernel(const ResizeArgs a)
{
    constexpr bool U8 = std::is_same<T, uint8_t>::value;
    constexpr int LANES = std::is_same<T, float>::value ? 4 : 8;
    __shared__ CubRow rows[CUB_R];
    const int dx = blockIdx.x * CUB_W + threadIdx.x, y0 = blockIdx.y * CUB_R;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride;
    if (threadIdx.x < CUB_R) {
        int sy; float fy;
        lin_coord(y0 + threadIdx.x, a.scale_y, a.inv_scale_y, 0, sy, fy);
        CubRow rw;
        cubic_coeffs(fy, rw.b);
#pragma unroll
        for (int k = 0; k < 4; k++) rw.r[k] = clipi(sy - 1 + k, 0, a.src.h);
        rows[threadIdx.x] = rw;
    }
    __syncthreads();
    if (dx >= a.dst.w) return;
    int sx; float fx;
    lin_coord(dx, a.scale_x, a.inv_scale_x, 0, sx, fx);
    float al[4];
    cubic_coeffs(fx, al);
    int xo[4];                                                       // element offsets of the four taps (channel 0)
#pragma unroll
    for (int j = 0; j < 4; j++) xo[j] = clipi(sx - 1 + j, 0, a.src.w) * CN;   // the +-cn loops of :2017-2023 clamp per channel
    int ia[4];
    if constexpr (U8) {
#pragma unroll
        for (int k = 0; k < 4; k++) ia[k] = sat_s16(rne_f(fmul(al[k], 2048.f)));
    }
    const int dwidth = a.dst.w * CN, nbody = dwidth - dwidth % LANES;
    const int ny = min(CUB_R, a.dst.h - y0);
    for (int r = 0; r < ny; r++) {
        const CubRow rw = rows[r];
        T* D = reinterpret_cast<T*>(dst + (long long)(y0 + r) * a.dst.step) + dx * CN;
        const T* R0 = reinterpret_cast<const T*>(src + (long long)rw.r[0] * a.src.step);
        const T* R1 = reinterpret_cast<const T*>(src + (long long)rw.r[1] * a.src.step);
        const T* R2 = reinterpret_cast<const T*>(src + (long long)rw.r[2] * a.src.step);
        const T* R3 = reinterpret_cast<const T*>(src + (long long)rw.r[3] * a.src.step);
#pragma unroll
        for (int c = 0; c < CN; c++) {
            const bool body = dx * CN + c < nbody;
            if constexpr (U8) {
                int ib[4];
#pragma unroll
                for (int k = 0; k < 4; k++) ib[k] = sat_s16(rne_f(fmul(rw.b[k], 2048.f)));
                auto h = [&](const T* R) { return __ldg(R + xo[0] + c) * ia[0] + __ldg(R + xo[1] + c) * ia[1] + __ldg(R + xo[2] + c) * ia[2] + __ldg(R + xo[3] + c) * ia[3]; };
                const int S0 = h(R0), S1 = h(R1), S2 = h(R2), S3 = h(R3);
                int v;
                if (body) {
                    const float sc = 1.f / (2048.f * 2048.f);
                    const float t = fadd(fmul((float)S0, fmul((float)ib[0], sc)),
                                         fadd(fmul((float)S1, fmul((float)ib[1], sc)),
                                              fadd(fmul((float)S2, fmul((float)ib[2], sc)), fmul((float)S3, fmul((float)ib[3], sc)))));
                    v = sat_u8(sat_s16(rne_f(t)));
                } else
                    v = sat_u8((S0 * ib[0] + S1 * ib[1] + S2 * ib[2] + S3 * ib[3] + (1 << 21)) >> 22);
                store_out(&D[c], (uint8_t)v, a.keep_in_l2);
            } else if constexpr (std::is_same<T, double>::value) {
                // 64F: float coefficients, double products and sums; no vector body in the reference, so both passes add left to right
                auto h = [&](const T* R) {
                    return __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(0.0, __dmul_rn(__ldg(R + xo[0] + c), (double)al[0])), __dmul_rn(__ldg(R + xo[1] + c), (double)al[1])),
                                               __dmul_rn(__ldg(R + xo[2] + c), (double)al[2])), __dmul_rn(__ldg(R + xo[3] + c), (double)al[3]));
                };
                const double S0 = h(R0), S1 = h(R1), S2 = h(R2), S3 = h(R3);
                D[c] = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(S0, (double)rw.b[0]), __dmul_rn(S1, (double)rw.b[1])), __dmul_rn(S2, (double)rw.b[2])), __dmul_rn(S3, (double)rw.b[3]));
            } else {
                // (((0 + v0*a0) + v1*a1) + v2*a2) + v3*a3: the leading 0 only matters for the sign of a zero result
                auto h = [&](const T* R) {
                    return fadd(fadd(fadd(fadd(0.f, fmul(Elem<T>::to_f(__ldg(R + xo[0] + c)), al[0])), fmul(Elem<T>::to_f(__ldg(R + xo[1] + c)), al[1])),
                                     fmul(Elem<T>::to_f(__ldg(R + xo[2] + c)), al[2])), fmul(Elem<T>::to_f(__ldg(R + xo[3] + c)), al[3]));
                };
                const float S0 = h(R0), S1 = h(R1), S2 = h(R2), S3 = h(R3);
                const float t = body ? fadd(fmul(S0, rw.b[0]), fadd(fmul(S1, rw.b[1]), fadd(fmul(S2, rw.b[2]), fmul(S3, rw.b[3]))))
                                     : fadd(fadd(fadd(fmul(S0, rw.b[0]), fmul(S1, rw.b[1])), fmul(S2, rw.b[2])), fmul(S3, rw.b[3]));
                store_out(&D[c], Elem<T>::from_f(t), a.keep_in_l2);
            }
        }
    }
}

template <typename T>
static cudaError_t launch_cubic(const ResizeArgs& a, int cn, int n_frames, cudaStream_t s)
{
    const dim3 grid((a.dst.w + CUB_W - 1) / CUB_W, (a.dst.h + CUB_R - 1) / CUB_R, n_frames);
    if (grid.y > 65535) return cudaErrorInvalidValue;
    switch (cn) {
    case 1: resize_cubic_kernel<T, 1><<<grid, CUB_W, 0, s>>>(a); break;
    case 2: resize_cubic_kernel<T, 2><<<grid, CUB_W, 0, s>>>(a); break;
    case 3: resize_cubic_kernel<T, 3><<<grid, CUB_W, 0, s>>>(a); break;
    case 4: resize_cubic_kernel<T, 4><<<grid, CUB_W, 0, s>>>(a); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------- LANCZOS4 -------
// resize.cpp:2066-2155: 8 x 8 clamped taps (sx-3..sx+4, sy-3..sy+4).  The per-column / per-row coefficients come from
// fp64 sin/cos of the host libm (interpolateLanczos4 :974-1003), which device code cannot reproduce bit for bit, so they
// are built on the host and read from a.ltab_x / a.ltab_y -- only the host-pointer hook has somewhere to put them.
// Vertical pass: 8U int with (sum + 2^21) >> 22; 16U adds left to right everywhere (SSE4.1 body, resize.sse4_1.cpp:189-226);
// 16S (first width & ~7 elements) and 32F (first width & ~3) nest from the right in the vector body (:1562-1626).
template <typename T>
__global__ void __launch_bounds__(256)
resize_lanczos4_kernel(const ResizeArgs a, int cn)
{
    constexpr bool U8 = std::is_same<T, uint8_t>::value;
    constexpr int LANES = std::is_same<T, float>::value ? 4 : std::is_same<T, int16_t>::value ? 8 : 0;
    const int o = blockIdx.x * blockDim.x + threadIdx.x, dy = blockIdx.y;
    const int dwidth = a.dst.w * cn, swidth = a.src.w * cn;
    if (o >= dwidth) return;
    const int dx = o / cn, c = o - dx * cn;
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    uint8_t* dst = a.dst.p + (long long)blockIdx.z * a.dst_frame_stride;
    int sx, sy; float fx, fy;
    lin_coord(dx, a.scale_x, a.inv_scale_x, 0, sx, fx);
    lin_coord(dy, a.scale_y, a.inv_scale_y, 0, sy, fy);
    const float* al = a.ltab_x + (long long)dx * 8;
    const float* be = a.ltab_y + (long long)dy * 8;
    int xi[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        int sxj = (sx - 3 + j) * cn + c;
        if ((unsigned)sxj >= (unsigned)swidth) sxj = sxj < 0 ? c : (a.src.w - 1) * cn + c;
        xi[j] = sxj;
    }
    T* D = reinterpret_cast<T*>(dst + (long long)dy * a.dst.step) + o;
    if constexpr (U8) {
        int ia[8];
#pragma unroll
        for (int k = 0; k < 8; k++) ia[k] = sat_s16(rne_f(fmul(__ldg(al + k), 2048.f)));
        int sum = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const uint8_t* R = src + (long long)clipi(sy - 3 + k, 0, a.src.h) * a.src.step;
            int h = 0;
#pragma unroll
            for (int j = 0; j < 8; j++) h += R[xi[j]] * ia[j];
            sum += h * sat_s16(rne_f(fmul(__ldg(be + k), 2048.f)));
        }
        store_out(D, (uint8_t)sat_u8((sum + (1 << 21)) >> 22), a.keep_in_l2);
    } else if constexpr (std::is_same<T, double>::value) {      // 64F: float coefficients, double products and sums, scalar order
        double S[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const double* R = reinterpret_cast<const double*>(src + (long long)clipi(sy - 3 + k, 0, a.src.h) * a.src.step);
            double h = __dadd_rn(0.0, __dmul_rn(R[xi[0]], (double)__ldg(al)));
#pragma unroll
            for (int j = 1; j < 8; j++) h = __dadd_rn(h, __dmul_rn(R[xi[j]], (double)__ldg(al + j)));
            S[k] = h;
        }
        double t = __dmul_rn(S[0], (double)__ldg(be));
#pragma unroll
        for (int k = 1; k < 8; k++) t = __dadd_rn(t, __dmul_rn(S[k], (double)__ldg(be + k)));
        *D = t;
    } else {
        float av[8], S[8];
#pragma unroll
        for (int k = 0; k < 8; k++) av[k] = __ldg(al + k);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const T* R = reinterpret_cast<const T*>(src + (long long)clipi(sy - 3 + k, 0, a.src.h) * a.src.step);
            float h = fadd(0.f, fmul(Elem<T>::to_f(R[xi[0]]), av[0]));
#pragma unroll
            for (int j = 1; j < 8; j++) h = fadd(h, fmul(Elem<T>::to_f(R[xi[j]]), av[j]));
            S[k] = h;
        }
        const bool body = LANES > 0 && o < dwidth - dwidth % (LANES > 0 ? LANES : 1);
        float t;
        if (body) {
            t = fmul(S[7], __ldg(be + 7));
#pragma unroll
            for (int k = 6; k >= 0; k--) t = fadd(fmul(S[k], __ldg(be + k)), t);
        } else {
            t = fmul(S[0], __ldg(be));
#pragma unroll
            for (int k = 1; k < 8; k++) t = fadd(t, fmul(S[k], __ldg(be + k)));
        }
        store_out(D, Elem<T>::from_f(t), a.keep_in_l2);
    }
}

template <typename T>
static cudaError_t launch_lanczos4(const ResizeArgs& a, int cn, int n_frames, cudaStream_t s)
{
    if (!a.ltab_x || !a.ltab_y || a.dst.h > 65535) return cudaErrorInvalidValue;
    resize_lanczos4_kernel<T><<<dim3((a.dst.w * cn + 255) / 256, a.dst.h, n_frames), 256, 0, s>>>(a, cn);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------- blobFromImages --
// One thread per blob pixel: the CV_8U value cv::resize would produce (generic 2-tap linear with LinOps<uint8_t>, the exact
// 2x2 box, or the source pixel itself when no resize is needed), then the reference's planar fp32 write
// p_blob[ch] = (float(v) - mean[ch]) * scale[ch] with ch the blob channel after swapRB (dnn_utils.cpp:143-241).
template <int CN, int MODE, bool PADDED>   // MODE 0: equal sizes, 1: RS_LINEAR, 2: 2x2 box; PADDED: the blob reaches outside the resized image (letterbox)
__global__ void __launch_bounds__(RT_W * RT_H) blob_u8_kernel(const ResizeArgs a, const BlobArgs b)
{
    const int bx = blockIdx.x * RT_W + threadIdx.x, by = blockIdx.y * RT_H + threadIdx.y;
    if (bx >= b.bw || by >= b.bh) return;
    const int x = bx + b.ox, y = by + b.oy;                               // the pixel of the resized image this blob pixel shows
    const uint8_t* src = a.src.p + (long long)blockIdx.z * a.src_frame_stride;
    const int W = a.src.w, H = a.src.h;
    uint8_t v[CN];
    // (a template flag, not a run-time test: with the test in every instantiation the plain resize form ran 121 us instead of 103)
    if (PADDED && (x < 0 || y < 0 || x >= a.dst.w || y >= a.dst.h)) {     // letterbox border (copyMakeBorder, dnn_utils.cpp:330)
#pragma unroll
        for (int k = 0; k < CN; k++) v[k] = b.pad[k];
    } else if constexpr (MODE == 0) {
        const uint8_t* S = src + (long long)y * a.src.step + x * CN;
#pragma unroll
        for (int k = 0; k < CN; k++) v[k] = S[k];
    } else if constexpr (MODE == 2) {
        const uint8_t* S = src + (long long)(2 * y) * a.src.step + 2 * x * CN;
        const uint8_t* S1 = S + a.src.step;
#pragma unroll
        for (int k = 0; k < CN; k++) v[k] = (uint8_t)(((int)S[k] + (int)S[k + CN] + (int)S1[k] + (int)S1[k + CN] + 2) >> 2);
    } else {
        int sx, sy; float fx, fy;
        lin_coord(x, a.scale_x, a.inv_scale_x, 0, sx, fx);
        lin_coord(y, a.scale_y, a.inv_scale_y, 0, sy, fy);
        if (sx < 0) { fx = 0; sx = 0; }                                   // resize.cpp:3918-3923
        const bool edge = sx + 1 >= W;
        if (sx >= W - 1) { fx = 0; sx = W - 1; }
        const int o1 = edge ? 0 : CN;
        LinOps<uint8_t> ops(fx);
        ops.set_row(fy, LinOps<uint8_t>::pack_row(fy));
        const uint8_t* r0 = src + (long long)clipi(sy, 0, H) * a.src.step + sx * CN;
        const uint8_t* r1 = src + (long long)clipi(sy + 1, 0, H) * a.src.step + sx * CN;
#pragma unroll
        for (int k = 0; k < CN; k++) v[k] = ops(r0[k], r0[k + o1], r1[k], r1[k + o1], edge);
    }
    const long long plane = (long long)b.bw * b.bh;
    float* out = b.blob + (long long)blockIdx.z * CN * plane + (long long)by * b.bw + bx;
#pragma unroll
    for (int k = 0; k < CN; k++) {
        const int ch = (b.swap_rb && CN >= 3 && k < 3) ? 2 - k : k;       // image channel k lands in blob channel ch
        __stcs(out + ch * plane, fmul(fsub((float)v[k], b.mean[ch]), b.scale[ch]));
    }
}

int launch_blob_u8(const ResizeArgs& a, const BlobArgs& b, int mode, int cn, int n_frames, cudaStream_t s)
{
    dim3 block(RT_W, RT_H), grid((b.bw + RT_W - 1) / RT_W, (b.bh + RT_H - 1) / RT_H, n_frames);
    const bool padded = b.ox < 0 || b.oy < 0 || b.ox + b.bw > a.dst.w || b.oy + b.bh > a.dst.h;
#define B200_BLOB_P(CNV, P) do { \
        if (mode == -2) blob_u8_kernel<CNV, 0, P><<<grid, block, 0, s>>>(a, b); \
        else if (mode == RS_LINEAR) blob_u8_kernel<CNV, 1, P><<<grid, block, 0, s>>>(a, b); \
        else blob_u8_kernel<CNV, 2, P><<<grid, block, 0, s>>>(a, b); } while (0)
#define B200_BLOB(CNV) do { if (padded) B200_BLOB_P(CNV, true); else B200_BLOB_P(CNV, false); } while (0)
    if (cn == 1) B200_BLOB(1); else if (cn == 3) B200_BLOB(3); else if (cn == 4) B200_BLOB(4); else return (int)cudaErrorInvalidValue;
#undef B200_BLOB_P
#undef B200_BLOB
    return (int)cudaGetLastError();
}

// largest number of table entries any destination column has (the counting of area_axis_setup / computeResizeAreaTab,
// resize.cpp:3334-3373, in the same fp64 operations): picks the tap bound the kernel is compiled for
static int area_max_taps(int ssize, int dsize, double scale)
{
    static thread_local struct { int ssize, dsize; double scale; int n; } memo = {0, 0, 0.0, 0};
    if (memo.ssize == ssize && memo.dsize == dsize && memo.scale == scale) return memo.n;
    int nmax = 0;
    for (int d = 0; d < dsize; d++) {
        const double fsx1 = d * scale, fsx2 = fsx1 + scale;
        int sx1 = (int)ceil(fsx1), sx2 = (int)floor(fsx2);
        sx2 = std::min(sx2, ssize - 1); sx1 = std::min(sx1, sx2);
        int n = (sx1 - fsx1 > 1e-3) + std::max(sx2 - sx1, 0) + (fsx2 - sx2 > 1e-3);
        nmax = std::max(nmax, n);
    }
    memo = {ssize, dsize, scale, nmax};
    return nmax;
}

int resize_select_mode(ResizeArgs& a, int interpolation)
{
    a.scale_x = 1. / a.inv_scale_x; a.scale_y = 1. / a.inv_scale_y;       // resize.cpp:3764
    a.ifxd = a.scale_x; a.ifyd = a.scale_y;                                 // :1128 computes the same quotients
    a.ifx = a.ifx0 = a.ify = a.ify0 = 0;
    a.area_mode = 0; a.kmax_x = a.kmax_y = 0; a.nmax_x = 0; a.lin_separable = 0;
    a.ltab_x = a.ltab_y = nullptr;
    // saturate_cast<int>(double) == cvRound (x86 cvtsd2si: nearest-even); nearbyint under the default mode
    auto rne = [](double v) { double r = nearbyint(v); return (r >= -2147483648.0 && r <= 2147483647.0) ? (int)r : INT_MIN; };
    a.iscale_x = rne(a.scale_x); a.iscale_y = rne(a.scale_y);               // :3766-3767
    if (interpolation == I_NEAREST) return RS_NEAREST;
    if (interpolation == I_NEAREST_EXACT) {                                 // :1270-1273 (int arithmetic, src dims < 2^15)
        if (a.src.w >= 32768 || a.src.h >= 32768) return -1;
        a.ifx = (int)((((unsigned)a.src.w << 16) + a.dst.w / 2) / a.dst.w); a.ifx0 = a.ifx / 2 - a.src.w % 2;
        a.ify = (int)((((unsigned)a.src.h << 16) + a.dst.h / 2) / a.dst.h); a.ify0 = a.ify / 2 - a.src.h % 2;
        return RS_NEAREST_EXACT;
    }
    const bool is_area_fast = fabs(a.scale_x - a.iscale_x) < DBL_EPSILON && fabs(a.scale_y - a.iscale_y) < DBL_EPSILON; // :3769-3770
    if (interpolation == I_LINEAR_EXACT) {                                  // :3782-3798
        if (is_area_fast && a.iscale_x == 2 && a.iscale_y == 2 && a.cn != 2) interpolation = I_AREA;
        else return RS_LINEAR_EXACT;
    }
    if (interpolation == I_CUBIC) return RS_CUBIC;                          // :3884-3885
    if (interpolation == I_LANCZOS4) return RS_LANCZOS4;                    // :3886-3887 (needs the host-built tables)
    if (interpolation != I_LINEAR && interpolation != I_AREA) return -1;
    if (interpolation == I_LINEAR && is_area_fast && a.iscale_x == 2 && a.iscale_y == 2) interpolation = I_AREA;          // :3817-3818
    if (interpolation == I_AREA && a.scale_x >= 1 && a.scale_y >= 1) {                                                     // :3822
        if (is_area_fast) return RS_AREA_FAST;
        // entries per dst index: optional leading partial + floor(scale)+1 whole + optional trailing partial
        a.kmax_x = (int)floor(a.scale_x) + 3; a.kmax_y = (int)floor(a.scale_y) + 3;
        a.nmax_x = area_max_taps(a.src.w, a.dst.w, a.scale_x);
        return RS_AREA;
    }
    a.area_mode = interpolation == I_AREA;                                  // :3877
    // consecutive destination rows share source rows only when the vertical scale is below 2
    a.lin_separable = a.scale_y < 2.0 && !a.force_generic;
    return RS_LINEAR;
}

template <typename T, int CN>
static cudaError_t launch_mode(const ResizeArgs& a, int mode, dim3 grid, dim3 block, cudaStream_t s)
{
    if (mode == RS_LINEAR) {
        // large images: 8 rows per thread, 4 row groups per CTA (tile 128 x 32); smaller ones shorter threads so that the
        // grid still covers the machine
        const long long px = (long long)a.dst.w * a.dst.h * grid.z;
        typedef typename LinOps<T>::HT HT;
        ResizeArgs b = a;
#define B200_LIN(LH, RGV) do { \
            const size_t sm_ = (size_t)(LH + 2) * CN * LT_W * RGV * sizeof(HT); \
            if (sm_ > 48 * 1024) b.lin_separable = 0;          /* uniform fallback to the direct path */ \
            dim3 lgrid((a.dst.w + LT_W - 1) / LT_W, (a.dst.h + LH * RGV - 1) / (LH * RGV), grid.z); \
            resize_linear_kernel<T, CN, LH, RGV><<<lgrid, dim3(LT_W, RGV), b.lin_separable ? sm_ : 0, s>>>(b); } while (0)
        if (px >= 4800000) B200_LIN(8, 4);
        else if (px >= 2400000) B200_LIN(4, 4);
        else B200_LIN(2, 1);
#undef B200_LIN
    }
    else if (mode == RS_AREA_FAST) resize_area_fast_kernel<T, CN><<<grid, block, 0, s>>>(a);
    else if (mode == RS_AREA) {
        // 4 destination rows per thread (the source rows of a tile are walked once; longer threads measured slower)
        const int rg = 1;
        const size_t sh = sizeof(float) * ((size_t)AT_W * a.kmax_x + (size_t)4 * rg * a.kmax_y);
        if (sh > 40 * 1024) return cudaErrorNotSupported;
        dim3 agrid((a.dst.w + AT_W - 1) / AT_W, (a.dst.h + 4 * rg - 1) / (4 * rg), grid.z);
#define B200_AREA_ROWS(KXV) do { \
            if (rg == 4) resize_area_kernel<T, CN, KXV, 4, 4><<<agrid, dim3(AT_W, 4), sh, s>>>(a); \
            else resize_area_kernel<T, CN, KXV, 4, 1><<<agrid, dim3(AT_W, 1), sh, s>>>(a); } while (0)
        if (a.nmax_x <= 4) B200_AREA_ROWS(4);
        else if (a.nmax_x <= 6) B200_AREA_ROWS(6);
        else B200_AREA_ROWS(0);
#undef B200_AREA_ROWS
    } else return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// 64F: generic linear only (the area / cubic / Lanczos kernels carry fp32 accumulators)
static cudaError_t launch_linear_f64(const ResizeArgs& a, int cn, dim3 grid, cudaStream_t s)
{
    ResizeArgs b = a;
    b.lin_separable = 0;
    dim3 lgrid((a.dst.w + LT_W - 1) / LT_W, (a.dst.h + 7) / 8, grid.z);
    switch (cn) {
    case 1: resize_linear_kernel<double, 1, 2, 4><<<lgrid, dim3(LT_W, 4), 0, s>>>(b); break;
    case 2: resize_linear_kernel<double, 2, 2, 4><<<lgrid, dim3(LT_W, 4), 0, s>>>(b); break;
    case 3: resize_linear_kernel<double, 3, 2, 4><<<lgrid, dim3(LT_W, 4), 0, s>>>(b); break;
    case 4: resize_linear_kernel<double, 4, 2, 4><<<lgrid, dim3(LT_W, 4), 0, s>>>(b); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

template <typename T>
static cudaError_t launch_cn(const ResizeArgs& a, int mode, int cn, dim3 grid, dim3 block, cudaStream_t s)
{
    switch (cn) {
    case 1: return launch_mode<T, 1>(a, mode, grid, block, s);
    case 2: return launch_mode<T, 2>(a, mode, grid, block, s);
    case 3: return launch_mode<T, 3>(a, mode, grid, block, s);
    case 4: return launch_mode<T, 4>(a, mode, grid, block, s);
    default: return cudaErrorInvalidValue;
    }
}

template <typename U, bool EXACT>
static cudaError_t launch_nn(const ResizeArgs& a, int cn, dim3 grid, dim3 block, cudaStream_t s)
{
    switch (cn) {
    case 1: resize_nn_kernel<U, 1, EXACT><<<grid, block, 0, s>>>(a); break;
    case 2: resize_nn_kernel<U, 2, EXACT><<<grid, block, 0, s>>>(a); break;
    case 3: resize_nn_kernel<U, 3, EXACT><<<grid, block, 0, s>>>(a); break;
    case 4: resize_nn_kernel<U, 4, EXACT><<<grid, block, 0, s>>>(a); break;
    default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

int launch_resize(const ResizeArgs& a, int mode, int depth, int cn, int n_frames, cudaStream_t stream)
{
    dim3 block(RT_W, RT_H);
    dim3 grid((a.dst.w + RT_W - 1) / RT_W, (a.dst.h + RT_H - 1) / RT_H, n_frames);
    if (mode == RS_NEAREST || mode == RS_NEAREST_EXACT) {
        const int esz1 = depth <= D_8S ? 1 : depth <= D_16S ? 2 : depth <= D_32F ? 4 : 8;
        if (mode == RS_NEAREST) switch (esz1) {
        case 1: return (int)launch_nn<uint8_t, false>(a, cn, grid, block, stream);
        case 2: return (int)launch_nn<uint16_t, false>(a, cn, grid, block, stream);
        case 4: return (int)launch_nn<uint32_t, false>(a, cn, grid, block, stream);
        default: return (int)launch_nn<unsigned long long, false>(a, cn, grid, block, stream);
        }
        switch (esz1) {
        case 1: return (int)launch_nn<uint8_t, true>(a, cn, grid, block, stream);
        case 2: return (int)launch_nn<uint16_t, true>(a, cn, grid, block, stream);
        case 4: return (int)launch_nn<uint32_t, true>(a, cn, grid, block, stream);
        default: return (int)launch_nn<unsigned long long, true>(a, cn, grid, block, stream);
        }
    }
    if (mode == RS_LINEAR_EXACT) {
        if (depth == D_8U) return (int)launch_exact<uint8_t>(a, cn, n_frames, stream);
        if (depth == D_16U) return (int)launch_exact<uint16_t>(a, cn, n_frames, stream);
        if (depth == D_8S) return (int)launch_exact<int8_t>(a, cn, n_frames, stream);
        if (depth == D_16S) return (int)launch_exact<int16_t>(a, cn, n_frames, stream);
        if (depth == D_32S) {
            resize_linear_exact_s32_kernel<<<dim3((a.dst.w * cn + LT_W - 1) / LT_W, a.dst.h, n_frames), LT_W, 0, stream>>>(a, cn);
            return (int)cudaGetLastError();
        }
        return (int)cudaErrorInvalidValue;
    }
    if (mode == RS_LANCZOS4) {
        switch (depth) {
        case D_8U:  return (int)launch_lanczos4<uint8_t>(a, cn, n_frames, stream);
        case D_16U: return (int)launch_lanczos4<uint16_t>(a, cn, n_frames, stream);
        case D_16S: return (int)launch_lanczos4<int16_t>(a, cn, n_frames, stream);
        case D_32F: return (int)launch_lanczos4<float>(a, cn, n_frames, stream);
        case D_64F: return (int)launch_lanczos4<double>(a, cn, n_frames, stream);
        default: return (int)cudaErrorInvalidValue;
        }
    }
    if (mode == RS_CUBIC) {
        switch (depth) {
        case D_8U:  return (int)launch_cubic<uint8_t>(a, cn, n_frames, stream);
        case D_16U: return (int)launch_cubic<uint16_t>(a, cn, n_frames, stream);
        case D_16S: return (int)launch_cubic<int16_t>(a, cn, n_frames, stream);
        case D_32F: return (int)launch_cubic<float>(a, cn, n_frames, stream);
        case D_64F: return (int)launch_cubic<double>(a, cn, n_frames, stream);
        default: return (int)cudaErrorInvalidValue;
        }
    }
    if (mode == RS_LINEAR && depth == D_32F && cn == 1 && linear_f32c1_x4_applicable(a)) return (int)launch_linear_f32c1_x4(a, n_frames, stream);
    if (mode == RS_AREA_FAST && depth == D_8U && a.iscale_x == 2 && a.iscale_y == 2 && !a.force_generic &&
        box2x2_u8_applicable(a, cn))
        return launch_box2x2_u8(a, cn, n_frames, stream);
    if (mode == RS_AREA_FAST && depth == D_8U && !a.force_generic && area_int_u8_applicable(a, cn))
        return launch_area_int_u8(a, cn, n_frames, stream);
    if (depth == D_64F) {
        if (mode == RS_LINEAR) return (int)launch_linear_f64(a, cn, grid, stream);
        if (mode == RS_AREA_FAST) {
            resize_area_fast_f64_kernel<<<dim3((a.dst.w * cn + 255) / 256, a.dst.h, grid.z), 256, 0, stream>>>(a, cn);
            return (int)cudaGetLastError();
        }
        if (mode == RS_AREA) {
            const size_t sh64 = (size_t)(A64_W * a.kmax_x + a.kmax_y) * sizeof(float);
            if (sh64 > 40 * 1024) return (int)cudaErrorInvalidValue;
            resize_area_f64_kernel<<<dim3((a.dst.w + A64_W - 1) / A64_W, a.dst.h, grid.z), A64_W, sh64, stream>>>(a, cn);
            return (int)cudaGetLastError();
        }
        return (int)cudaErrorInvalidValue;
    }
    switch (depth) {
    case D_8U:  return (int)launch_cn<uint8_t>(a, mode, cn, grid, block, stream);
    case D_16U: return (int)launch_cn<uint16_t>(a, mode, cn, grid, block, stream);
    case D_16S: return (int)launch_cn<int16_t>(a, mode, cn, grid, block, stream);
    case D_32F: return (int)launch_cn<float>(a, mode, cn, grid, block, stream);
    default: return (int)cudaErrorInvalidValue;
    }
}

} // namespace b200
