// scan.cuh — building blocks of the device-wide exclusive prefix sums (u32): warp / block scans and the
// single-block scan of per-block totals.  The per-element passes live with their producers (tree_count_kernel
// in tree_build.cuh, the flag scan in bvh_ref.cuh).
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

// Single block: exclusive scan of the block sums in place; total -> *total_out.
// cond (optional): run only if *cond != 0.  stats (optional): when the build is about to be redone from a full sort
// (stats->resort, see radix_sort.cuh), clear the statistics the first tree_count_kernel pass accumulated.
__global__ void __launch_bounds__(1024) scan_bsum_kernel(uint32_t *__restrict__ bsum, uint32_t nb,
                                                         uint32_t *total_out, Meta *stats = nullptr,
                                                         const uint32_t *cond = nullptr) {
  if (cond && *cond == 0u) return;
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
  if (threadIdx.x == 0 && stats && stats->resort) {
    stats->n_leaves = 0;
    stats->overflow_leaves = 0;
    stats->max_depth = 0;
  }
}

// Two-launch version for the tree build (round 2): one block of 1024 is 33 us at 8M bodies (31 250 block totals walked
// by a single CTA).  scan_chunks_kernel scans each 1024-entry chunk in place and leaves the chunk totals;
// scan_bases_kernel adds to every chunk the sum of the totals before it (a few dozen values, summed by each block
// itself) and the last block writes the grand total (+ the statistics reset of a resorted build is gone: round 2).
__global__ void __launch_bounds__(1024) scan_chunks_kernel(uint32_t *__restrict__ bsum, uint32_t nb,
                                                           uint32_t *__restrict__ ctot) {
  const uint32_t i = blockIdx.x * 1024u + threadIdx.x;
  const uint32_t v = (i < nb) ? bsum[i] : 0u;
  uint32_t tot;
  const uint32_t ex = block_exclusive_scan<1024>(v, tot);
  if (i < nb) bsum[i] = ex;
  if (threadIdx.x == 0) ctot[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(1024) scan_bases_kernel(uint32_t *__restrict__ bsum, uint32_t nb,
                                                          const uint32_t *__restrict__ ctot, uint32_t *total_out) {
  __shared__ uint32_t base_s;
  if (threadIdx.x < 32) { // chunks before this one (at most n / 262 144 of them: a handful)
    uint32_t a = 0;
    for (uint32_t j = threadIdx.x; j < blockIdx.x; j += 32) a += ctot[j];
    a = __reduce_add_sync(0xffffffffu, a);
    if (threadIdx.x == 0) base_s = a;
  }
  __syncthreads();
  const uint32_t base = base_s;
  const uint32_t i = blockIdx.x * 1024u + threadIdx.x;
  if (i < nb) bsum[i] += base;
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0 && total_out) *total_out = base + ctot[blockIdx.x];
}

} // namespace nb
