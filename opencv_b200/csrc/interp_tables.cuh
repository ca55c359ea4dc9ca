// INTER_CUBIC / INTER_LANCZOS4 coefficient tables: built on the host, resident in fixed __device__ arrays.
#pragma once
#include <cuda_runtime.h>

namespace b200 {

// device pointers to the tables of `interp` (2 = INTER_CUBIC, 4 = INTER_LANCZOS4) on the current device;
// the first call per device copies them (160 KB, synchronous)
cudaError_t interp_tables(int interp, const float** tab1d, const short** tab2d_i);
// the host copies (tests compare them with the oracle's)
void interp_tables_host(int interp, const float** tab1d, const short** tab2d_i);

// interpolateLanczos4 (imgwarp.cpp:99-127 == resize.cpp:974-1003) on the host: 8 coefficients for fraction x
void lanczos4_coeffs_host(float x, float* c);

} // namespace b200
