// standalone probe: TMA 3D byte-tensor tile load -> smem -> plain global store; then TMA store.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <stdint.h>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void probe(const __grid_constant__ CUtensorMap smap, const __grid_constant__ CUtensorMap dmap, uint8_t* out, int boxw, int boxh, int cx, int cy, int mode)
{
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (mode == 100) { out[threadIdx.x] = 1; return; }
    if (mode == 101) {
        if (threadIdx.x == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar)) : "memory");
        asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP1:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE1;\n\t"
        "bra WAIT_LOOP1;\n\t"
        "DONE1:\n\t}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
        out[threadIdx.x] = 2; return;
    }
    if (mode == 102) {
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(boxw * boxh) : "memory");
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"(smem_u32(smem)), "l"(&dmap), "r"(smem_u32(&bar)), "r"(cx), "r"(cy) : "memory");
        }
        asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP2:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE2;\n\t"
        "bra WAIT_LOOP2;\n\t"
        "DONE2:\n\t}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
        for (int i = threadIdx.x; i < boxw * boxh; i += blockDim.x) out[i] = smem[i];
        return;
    }
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(boxw * boxh) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"(smem_u32(smem)), "l"(&smap), "r"(smem_u32(&bar)), "r"(cx), "r"(cy), "r"(0) : "memory");
    }
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < boxw * boxh; i += blockDim.x) out[i] = smem[i];
    if (mode >= 1) {
        __syncthreads();
        for (int i = threadIdx.x; i < boxw * boxh; i += blockDim.x) smem[i] = (uint8_t)(smem[i] + 1);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                         ::"l"(&dmap), "r"(smem_u32(smem)), "r"(16), "r"(2), "r"(0) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
    }
}
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("ERR %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main(int argc, char** argv)
{
    int mode = argc > 1 ? atoi(argv[1]) : 0;
    int W = argc > 2 ? atoi(argv[2]) : 1440, H = 270, step = (W + 255) & ~255;
    int boxw = argc > 3 ? atoi(argv[3]) : 224, boxh = 34;
    void* p; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    EncodeTiledFn enc = (EncodeTiledFn)p;
    uint8_t *src, *dst, *out;
    CK(cudaMalloc(&src, step * H)); CK(cudaMalloc(&dst, step * H)); CK(cudaMalloc(&out, boxw * boxh));
    std::vector<uint8_t> h(step * H);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)(i * 7 + i / step);
    CK(cudaMemcpy(src, h.data(), h.size(), cudaMemcpyHostToDevice));
    CK(cudaMemset(dst, 0, step * H));
    CUtensorMap sm, dm;
    cuuint64_t gdim[3] = {(cuuint64_t)W, (cuuint64_t)H, 1}, gstr[2] = {(cuuint64_t)step, (cuuint64_t)step * H};
    cuuint32_t box[3] = {(cuuint32_t)boxw, (cuuint32_t)boxh, 1}, es[3] = {1, 1, 1};
    CUresult r = enc(&sm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, src, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode src %d\n", (int)r);
    r = enc(&dm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, dst, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (mode == 102) r = enc(&dm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, src, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode dst %d\n", (int)r);
    int cx = argc > 4 ? atoi(argv[4]) : 5, cy = argc > 5 ? atoi(argv[5]) : 3;
    probe<<<1, 256, boxw * boxh + 128>>>(sm, dm, out, boxw, boxh, cx, cy, mode);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<uint8_t> o(boxw * boxh);
    CK(cudaMemcpy(o.data(), out, o.size(), cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int y = 0; y < boxh; y++) for (int x = 0; x < boxw; x++) {
        int gx = cx + x, gy = cy + y;
        uint8_t e = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? h[(size_t)gy * step + gx] : 0;
        if (o[y * boxw + x] != e) bad++;
    }
    printf("mode %d W %d box %dx%d at (%d,%d): load mismatches %d\n", mode, W, boxw, boxh, cx, cy, bad);
    if (mode >= 1) {
        std::vector<uint8_t> d(step * H);
        CK(cudaMemcpy(d.data(), dst, d.size(), cudaMemcpyDeviceToHost));
        int badst = 0;
        for (int y = 0; y < boxh; y++) for (int x = 0; x < boxw; x++) {
            int gx = 16 + x, gy = 2 + y;
            if (gx < W && gy < H && d[(size_t)gy * step + gx] != (uint8_t)(o[y * boxw + x] + 1)) badst++;
        }
        printf("store mismatches %d\n", badst);
    }
    return 0;
}
