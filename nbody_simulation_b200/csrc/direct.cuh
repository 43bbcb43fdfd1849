// direct.cuh — K6: tiled O(N^2) direct sum with the reference pair force (src/main.rs:235-253).
// The device-side accuracy yardstick for the tree walk: fp32 pair terms, fp32 partial sums over
// one 256-source tile, fp64 accumulation across tiles and source chunks.
#pragma once
#include "common.cuh"

namespace nb {

constexpr int kDirectThreads = 256;

template <bool EXACT>
__global__ void __launch_bounds__(kDirectThreads) direct_kernel(const float *__restrict__ x, const float *__restrict__ y,
                                                                const uint32_t *__restrict__ m, uint32_t n,
                                                                const uint32_t *__restrict__ tslot, uint32_t nt,
                                                                double *__restrict__ acc, uint32_t chunk) {
  __shared__ float4 tile[kDirectThreads];
  const uint32_t t = blockIdx.x * kDirectThreads + threadIdx.x;
  const bool valid = t < nt;
  const uint32_t slot = tslot[valid ? t : 0];
  const float px = x[slot], py = y[slot];
  const uint32_t s0 = blockIdx.y * chunk;
  const uint32_t s1 = (s0 + chunk < n) ? s0 + chunk : n;
  double dax = 0.0, day = 0.0;
  for (uint32_t base = s0; base < s1; base += kDirectThreads) {
    const uint32_t j = base + threadIdx.x;
    tile[threadIdx.x] = (j < s1) ? make_float4(x[j], y[j], (float)m[j], 0.f) : make_float4(px, py, 0.f, 0.f);
    __syncthreads();
    const uint32_t cnt = (s1 - base < (uint32_t)kDirectThreads) ? s1 - base : (uint32_t)kDirectThreads;
    float fax = 0.f, fay = 0.f;
    if (cnt == (uint32_t)kDirectThreads) {
#pragma unroll 8
      for (int q = 0; q < kDirectThreads; ++q) {
        const float4 b = tile[q];
        pair_force<EXACT>(px, py, b.x, b.y, b.z, fax, fay);
      }
    } else {
      for (uint32_t q = 0; q < cnt; ++q) {
        const float4 b = tile[q];
        pair_force<EXACT>(px, py, b.x, b.y, b.z, fax, fay);
      }
    }
    dax += (double)fax;
    day += (double)fay;
    __syncthreads();
  }
  if (valid) {
    atomicAdd(&acc[2 * (size_t)t], dax);
    atomicAdd(&acc[2 * (size_t)t + 1], day);
  }
}

__global__ void direct_finish_kernel(const double *__restrict__ acc, uint32_t nt, float *__restrict__ ax,
                                     float *__restrict__ ay) {
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nt) {
    ax[t] = (float)acc[2 * (size_t)t];
    ay[t] = (float)acc[2 * (size_t)t + 1];
  }
}

} // namespace nb
