// TMA / mbarrier primitives and the host-side tensor-map encoder shared by the staged kernels
// (fast_kernels.cu, px4_kernels.cu).  sm_100a only.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstring>
#include <mutex>

namespace b200 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ int ld_acquire(const int* p) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase)
{
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
// tile load: global (tensor map, coordinates innermost first) -> shared, completion on an mbarrier.  The tensor maps used
// with this cover buffers of the launching GPU only.
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// contiguous bytes global -> shared (16-byte aligned on both sides, size a multiple of 16), completion on an mbarrier.  Used on
// buffers of the launching GPU only.
__device__ __forceinline__ void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// the same for data that is read once: its lines are the first to leave L2 again
__device__ __forceinline__ void bulk_load_1d_stream(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar)
{
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* smem_src, int c0, int c1, int c2, bool keep_in_l2)
{
    if (keep_in_l2) {
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                     ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2) : "memory");
    } else {
        unsigned long long pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3, %4}], [%1], %5;"
                     ::"l"(map), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "l"(pol) : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// barrier over the NT threads that work on a tile: the whole CTA (BAR 0) or a named barrier
template <int BAR, int NT> __device__ __forceinline__ void cta_sync()
{
    if constexpr (BAR == 0) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(NT) : "memory");
}

// ---------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn tma_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, []() {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    });
    return fn;
}

// 3-D tensor map (x = elements of a row, y = rows, z = frames) of 1- or 4-byte elements; box (boxw elements, boxh rows, 1
// frame), OOB elements read as zero, stores clipped.  Encoding costs ~1 us on the host; a per-thread cache of the last few
// descriptors makes repeated calls on the same buffers (video loops, benchmarks) skip it.
inline bool make_tensor_map(CUtensorMap* m, int elem_bytes, const void* base, long long row_elems, int rows, int frames,
                            long long step, long long frame_stride, int boxw, int boxh)
{
    EncodeTiledFn enc = tma_encode_fn();
    if (!enc || boxw > 256 || boxh > 256 || boxw < 1 || boxh < 1) return false;
    if (frames == 1 || frame_stride == 0) { frames = 1; frame_stride = step * rows; }
    struct Key { const void* base; long long row_elems, step, frame_stride; int rows, frames, boxw, boxh, elem_bytes, pad; };
    struct Slot { Key k; CUtensorMap map; bool used; };
    static thread_local Slot cache[8];
    static thread_local int next_slot = 0;
    Key key; memset(&key, 0, sizeof(key));               // padding bytes take part in the memcmp
    key.base = base; key.row_elems = row_elems; key.step = step; key.frame_stride = frame_stride;
    key.rows = rows; key.frames = frames; key.boxw = boxw; key.boxh = boxh; key.elem_bytes = elem_bytes;
    for (int i = 0; i < 8; i++)
        if (cache[i].used && memcmp(&cache[i].k, &key, sizeof(Key)) == 0) { *m = cache[i].map; return true; }
    cuuint64_t gdim[3] = {(cuuint64_t)row_elems, (cuuint64_t)rows, (cuuint64_t)frames};
    cuuint64_t gstr[2] = {(cuuint64_t)step, (cuuint64_t)frame_stride};
    cuuint32_t box[3] = {(cuuint32_t)boxw, (cuuint32_t)boxh, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(m, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base),
                     gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return false;
    Slot& sl = cache[next_slot]; next_slot = (next_slot + 1) & 7;
    memset(&sl.k, 0, sizeof(Key)); sl.k = key; sl.map = *m; sl.used = true;
    return true;
}
inline bool make_byte_map(CUtensorMap* m, const void* base, long long row_bytes, int rows, int frames, long long step,
                          long long frame_stride, int boxw, int boxh)
{
    return make_tensor_map(m, 1, base, row_bytes, rows, frames, step, frame_stride, boxw, boxh);
}

// number of SMs and opt-in shared memory per SM of the current device (remembered per device ordinal)
struct DeviceLimits { int sms; int smem_per_sm; int smem_per_block_optin; };
inline DeviceLimits device_limits()
{
    static DeviceLimits lim[64];
    static bool have[64] = {};
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return DeviceLimits{1, 48 * 1024, 48 * 1024};
    std::lock_guard<std::mutex> lk(mu);
    if (!have[dev]) {
        int sms = 0, per_sm = 0, per_blk = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaDeviceGetAttribute(&per_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
        cudaDeviceGetAttribute(&per_blk, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        lim[dev] = DeviceLimits{sms > 0 ? sms : 1, per_sm > 0 ? per_sm : 48 * 1024, per_blk > 0 ? per_blk : 48 * 1024};
        have[dev] = true;
    }
    return lim[dev];
}

} // namespace b200
