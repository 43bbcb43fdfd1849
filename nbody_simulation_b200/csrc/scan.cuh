// scan.cuh — device-wide exclusive prefix sum (u32), three launches, no host sync.
#pragma once
#include "common.cuh"

namespace nb {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems; // 2048

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= (unsigned)o) v += t;
  }
  return v;
}

// Exclusive scan of one value per thread across a block of NT threads; returns the exclusive
// prefix and writes the block total to `total` (all threads).
template <int NT> __device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t &total) {
  __shared__ uint32_t wsum[NT / 32];
  __shared__ uint32_t tot;
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint32_t inc = warp_inclusive_scan(v);
  if (lane == 31) wsum[w] = inc;
  __syncthreads();
  if (w == 0) {
    uint32_t s = (lane < NT / 32) ? wsum[lane] : 0;
    uint32_t si = warp_inclusive_scan(s);
    if (lane < NT / 32) wsum[lane] = si - s;
    if (lane == NT / 32 - 1) tot = si;
  }
  __syncthreads();
  total = tot;
  uint32_t r = wsum[w] + inc - v;
  __syncthreads(); // allow reuse of the shared arrays by a following call
  return r;
}

// n is read from *n_ptr when n_ptr != nullptr (device-resident sizes), else n_host.
__device__ __forceinline__ uint32_t scan_n(const uint32_t *n_ptr, uint32_t n_host) {
  return n_ptr ? *n_ptr : n_host;
}

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint32_t *__restrict__ in,
                                                                   const uint32_t *n_ptr, uint32_t n_host,
                                                                   uint32_t *__restrict__ bsum) {
  const uint32_t n = scan_n(n_ptr, n_host);
  const uint32_t base = blockIdx.x * kScanTile;
  uint32_t s = 0;
  if (base < n) {
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
      uint32_t i = base + k * kScanThreads + threadIdx.x;
      if (i < n) s += in[i];
    }
  }
  uint32_t tot;
  block_exclusive_scan<kScanThreads>(s, tot);
  if (threadIdx.x == 0) bsum[blockIdx.x] = tot;
}

// Single block: exclusive scan of the block sums in place; total -> *total_out.
__global__ void __launch_bounds__(1024) scan_bsum_kernel(uint32_t *__restrict__ bsum, uint32_t nb,
                                                         uint32_t *total_out) {
  __shared__ uint32_t carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (uint32_t base = 0; base < nb; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = (i < nb) ? bsum[i] : 0;
    uint32_t tot;
    uint32_t ex = block_exclusive_scan<1024>(v, tot);
    uint32_t carry = carry_s;
    if (i < nb) bsum[i] = carry + ex;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}

__global__ void __launch_bounds__(kScanThreads) scan_apply_kernel(const uint32_t *__restrict__ in,
                                                                  uint32_t *__restrict__ out, const uint32_t *n_ptr,
                                                                  uint32_t n_host, const uint32_t *__restrict__ bsum) {
  const uint32_t n = scan_n(n_ptr, n_host);
  const uint32_t base = blockIdx.x * kScanTile;
  if (base >= n) return;
  // blocked arrangement: thread t owns items [t*8, t*8+8) of the tile
  uint32_t v[kScanItems];
  uint32_t s = 0;
  const uint32_t i0 = base + threadIdx.x * kScanItems;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = (i0 + k < n) ? in[i0 + k] : 0;
    s += v[k];
  }
  uint32_t tot;
  uint32_t ex = block_exclusive_scan<kScanThreads>(s, tot) + bsum[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (i0 + k < n) out[i0 + k] = ex;
    ex += v[k];
  }
}

// Host helper: enqueue the three launches.  `n_max` bounds the grid; the live size is
// *n_ptr (device) or n_max.
inline void exclusive_scan_u32(const uint32_t *in, uint32_t *out, const uint32_t *n_ptr, uint32_t n_max,
                               uint32_t *bsum_scratch, uint32_t *total_out, cudaStream_t st) {
  uint32_t nb = (n_max + kScanTile - 1) / kScanTile;
  if (nb == 0) nb = 1;
  g_launches += 3;
  scan_reduce_kernel<<<nb, kScanThreads, 0, st>>>(in, n_ptr, n_max, bsum_scratch);
  scan_bsum_kernel<<<1, 1024, 0, st>>>(bsum_scratch, nb, total_out);
  scan_apply_kernel<<<nb, kScanThreads, 0, st>>>(in, out, n_ptr, n_max, bsum_scratch);
}

} // namespace nb
