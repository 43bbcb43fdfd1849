// radix_sort.cuh — K2: stable LSD radix sort of (u64 Morton key, u32 body slot) pairs.
//
// Round-1 formulation: per pass  tile histogram -> device-wide scan of the [digit][tile] table ->
// stable scatter.  Ranking inside a tile is warp-synchronous (match_any multi-split) with
// per-warp digit counters in shared memory, so the pass is stable without any sorting network.
// HBM traffic per pass: keys read twice (8+8 B), values once (4 B), pairs written once (12 B).
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace nb {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortRounds = 16;                                // 32-element rounds per warp
constexpr int kSortTile = kSortThreads * kSortRounds;          // 4096 elements per tile

template <int RADIX_BITS>
__global__ void __launch_bounds__(kSortThreads) sort_hist_kernel(const unsigned long long *__restrict__ keys, uint32_t n,
                                                                 int shift, uint32_t nblocks,
                                                                 uint32_t *__restrict__ table) {
  constexpr int RADIX = 1 << RADIX_BITS;
  __shared__ uint32_t hist[RADIX];
  for (int d = threadIdx.x; d < RADIX; d += kSortThreads) hist[d] = 0;
  __syncthreads();
  const uint32_t base = blockIdx.x * kSortTile;
#pragma unroll 4
  for (int k = 0; k < kSortRounds; ++k) {
    uint32_t i = base + k * kSortThreads + threadIdx.x;
    if (i < n) atomicAdd(&hist[(uint32_t)(keys[i] >> shift) & (RADIX - 1)], 1u);
  }
  __syncthreads();
  for (int d = threadIdx.x; d < RADIX; d += kSortThreads) table[(size_t)d * nblocks + blockIdx.x] = hist[d];
}

template <int RADIX_BITS>
__global__ void __launch_bounds__(kSortThreads) sort_scatter_kernel(
    const unsigned long long *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
    unsigned long long *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint32_t n, int shift,
    uint32_t nblocks, const uint32_t *__restrict__ table_scanned) {
  constexpr int RADIX = 1 << RADIX_BITS;
  __shared__ uint32_t cnt[kSortWarps][RADIX];
  __shared__ uint32_t gofs[RADIX];
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  for (int d = threadIdx.x; d < RADIX * kSortWarps; d += kSortThreads) (&cnt[0][0])[d] = 0;
  for (int d = threadIdx.x; d < RADIX; d += kSortThreads) gofs[d] = table_scanned[(size_t)d * nblocks + blockIdx.x];
  __syncthreads();

  // warp w owns the contiguous chunk [base + w*512, base + (w+1)*512) in 16 rounds of 32
  const uint32_t wbase = blockIdx.x * kSortTile + w * (32 * kSortRounds);
  unsigned long long key[kSortRounds];
  uint32_t val[kSortRounds];
  uint32_t rank[kSortRounds]; // digit in the high half, rank in the low half (both < 65536)
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    uint32_t i = wbase + k * 32 + lane;
    bool valid = i < n;
    key[k] = valid ? keys_in[i] : 0ull;
    val[k] = valid ? vals_in[i] : 0u;
  }
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    uint32_t i = wbase + k * 32 + lane;
    bool valid = i < n;
    uint32_t d = (uint32_t)(key[k] >> shift) & (RADIX - 1);
    uint32_t m = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
    uint32_t prev = valid ? cnt[w][d] : 0u;
    __syncwarp();
    uint32_t r = __popc(m & lt);
    if (valid && r == 0) cnt[w][d] = prev + __popc(m);
    __syncwarp();
    rank[k] = (d << 16) | (prev + r);
  }
  __syncthreads();
  // exclusive scan over the warps, per digit
  for (int d = threadIdx.x; d < RADIX; d += kSortThreads) {
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < kSortWarps; ++ww) {
      uint32_t t = cnt[ww][d];
      cnt[ww][d] = run;
      run += t;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    uint32_t i = wbase + k * 32 + lane;
    if (i < n) {
      uint32_t d = rank[k] >> 16;
      uint32_t pos = gofs[d] + cnt[w][d] + (rank[k] & 0xffffu);
      keys_out[pos] = key[k];
      vals_out[pos] = val[k];
    }
  }
}

struct SortScratch {
  uint32_t *table;  // [512 * nblocks_max]
  uint32_t *bsum;   // scan block sums
};

inline size_t sort_table_entries(uint32_t n_max) {
  uint32_t nb = (n_max + kSortTile - 1) / kSortTile;
  if (nb == 0) nb = 1;
  return (size_t)512 * nb;
}

// Sorts `bits` low key bits (48 or 49) in 6 passes: five 8-bit digits, then the remaining 8 or 9.
// Returns 0 if the result is in (keys_a, vals_a), 1 if in (keys_b, vals_b).
inline int radix_sort_pairs(unsigned long long *keys_a, uint32_t *vals_a, unsigned long long *keys_b,
                            uint32_t *vals_b, uint32_t n, int bits, const SortScratch &sc, cudaStream_t st) {
  if (n == 0) return 0;
  const uint32_t nblocks = (n + kSortTile - 1) / kSortTile;
  unsigned long long *kin = keys_a, *kout = keys_b;
  uint32_t *vin = vals_a, *vout = vals_b;
  int which = 0;
  int shift = 0;
  while (shift < bits) {
    int rb = (bits - shift == 9) ? 9 : 8;
    g_launches += 2;
    if (rb == 8) {
      sort_hist_kernel<8><<<nblocks, kSortThreads, 0, st>>>(kin, n, shift, nblocks, sc.table);
      exclusive_scan_u32(sc.table, sc.table, nullptr, 256u * nblocks, sc.bsum, nullptr, st);
      sort_scatter_kernel<8><<<nblocks, kSortThreads, 0, st>>>(kin, vin, kout, vout, n, shift, nblocks, sc.table);
    } else {
      sort_hist_kernel<9><<<nblocks, kSortThreads, 0, st>>>(kin, n, shift, nblocks, sc.table);
      exclusive_scan_u32(sc.table, sc.table, nullptr, 512u * nblocks, sc.bsum, nullptr, st);
      sort_scatter_kernel<9><<<nblocks, kSortThreads, 0, st>>>(kin, vin, kout, vout, n, shift, nblocks, sc.table);
    }
    shift += rb;
    unsigned long long *tk = kin; kin = kout; kout = tk;
    uint32_t *tv = vin; vin = vout; vout = tv;
    which ^= 1;
  }
  return which;
}

} // namespace nb
