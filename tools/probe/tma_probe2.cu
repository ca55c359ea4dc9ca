// canonical TMA probes: float 2D 32x32 box (programming-guide style), descriptor in param vs global memory
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <stdint.h>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ void body(const CUtensorMap* map, float* out, int bytes, int cx, int cy, int variant)
{
    __shared__ __align__(128) float tile[32 * 32];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(bytes) : "memory");
        if (variant == 0)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(tile)), "l"(map), "r"(smem_u32(&bar)), "r"(cx), "r"(cy) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                         ::"r"(smem_u32(tile)), "l"(map), "r"(smem_u32(&bar)), "r"(cx), "r"(cy) : "memory");
    }
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}" ::"r"(smem_u32(&bar)), "r"(0) : "memory");
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) out[i] = tile[i];
}
__global__ void k_param(const __grid_constant__ CUtensorMap map, float* out, int bytes, int cx, int cy, int variant) { body(&map, out, bytes, cx, cy, variant); }
__global__ void k_global(const CUtensorMap* map, float* out, int bytes, int cx, int cy, int variant) { body(map, out, bytes, cx, cy, variant); }
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("ERR %s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
int main(int argc, char** argv)
{
    int where = argc > 1 ? atoi(argv[1]) : 0, variant = argc > 2 ? atoi(argv[2]) : 0;
    void* p; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    EncodeTiledFn enc = (EncodeTiledFn)p;
    const int W = 256, H = 256;
    float *src, *out;
    CK(cudaMalloc(&src, W * H * 4)); CK(cudaMalloc(&out, 4096));
    std::vector<float> h(W * H);
    for (int i = 0; i < W * H; i++) h[i] = (float)i;
    CK(cudaMemcpy(src, h.data(), W * H * 4, cudaMemcpyHostToDevice));
    CUtensorMap m;
    cuuint64_t gdim[2] = {W, H}, gstr[1] = {W * 4};
    cuuint32_t box[2] = {32, 32}, es[2] = {1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, src, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d where %d variant %d\n", (int)r, where, variant);
    if (where == 0) k_param<<<1, 128>>>(m, out, 4096, 32, 64, variant);
    else {
        CUtensorMap* dm; CK(cudaMalloc(&dm, sizeof(m))); CK(cudaMemcpy(dm, &m, sizeof(m), cudaMemcpyHostToDevice));
        k_global<<<1, 128>>>(dm, out, 4096, 32, 64, variant);
    }
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    std::vector<float> o(1024);
    CK(cudaMemcpy(o.data(), out, 4096, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int y = 0; y < 32; y++) for (int x = 0; x < 32; x++) if (o[y * 32 + x] != h[(64 + y) * W + 32 + x]) bad++;
    printf("mismatches %d\n", bad);
    return 0;
}
