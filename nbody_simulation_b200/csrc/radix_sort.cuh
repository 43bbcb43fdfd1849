// radix_sort.cuh — K2: stable LSD radix sort of (u64 Morton key, u32 body slot) pairs, onesweep style.
//
//   1 launch : histograms of ALL passes from one read of the keys (8 B/key)
//   1 launch : exclusive scan of each pass's histogram (tiny)
//   per pass : ONE kernel — each tile ranks its keys (warp match_any multi-split, per-warp digit
//              counters in shared memory => stable), gets its global digit offsets by decoupled
//              look-back over the preceding tiles, stages the tile in shared memory in output order
//              and writes coalesced runs.
// HBM traffic: 8 B + passes x (12 B read + 12 B written) per pair  (6 passes: 152 B/pair).
// Tiles take their index from an atomic ticket, so a tile only ever waits on tiles that already run.
//
// Top-digit (speculative) sort.  The tree only needs the bodies ordered down to the deepest leaf: inside a leaf
// tree_finish_kernel orders them by body id whatever the sort left.  When the previous build's depth is known,
// radix_sort_pairs runs just the top `top_passes` digits (a stable LSD sort of those digits), tree_count_kernel
// checks that no leaf is deeper than the sorted prefix, and if one is it raises meta->resort / error bit 1: the rest
// of the step does nothing and the host redoes it with a full sort (do_steps in nbody_b200.cu).  Round 1 enqueued the
// full sort behind every build as conditional no-op launches instead; that chain (radix_sort_fallback) is gone.
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace nb {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortRoundsMax = 16;                          // 32-element rounds per warp (large inputs)
constexpr int kSortTileMax = kSortThreads * kSortRoundsMax; // 4096 pairs per tile
constexpr int kSortMaxPasses = 7;
constexpr int kSortMaxRadix = 512;

constexpr uint32_t kFlagLocal = 1u << 30;     // status word: tile-local digit count
constexpr uint32_t kFlagInclusive = 2u << 30; // status word: inclusive prefix over tiles 0..t
constexpr uint32_t kStatusValue = (1u << 30) - 1u;

struct SortPlan { // digit layout of one sort, passed by value to the kernels
  int passes;
  int shift[kSortMaxPasses];
  int bits[kSortMaxPasses];
  int hist_off[kSortMaxPasses]; // offset of the pass's bins in the histogram arrays
  int hist_total;
  int out_bin;      // quadtree keys (odd key_bits): first bin of the top digit with the "outside the root" bit set; else -1
  int sorted_bits;  // key bits, from the top, that the passes cover (== key_bits for a full sort)
};

inline void sort_plan_push(SortPlan &p, int shift, int b) {
  p.shift[p.passes] = shift;
  p.bits[p.passes] = b;
  p.hist_off[p.passes] = p.hist_total;
  p.hist_total += 1 << b;
  ++p.passes;
}

inline SortPlan make_sort_plan(int key_bits) { // 8-bit digits; a 9-bit last digit absorbs an odd top bit
  SortPlan p{};
  int shift = 0;
  while (shift < key_bits) {
    int b = (key_bits - shift == 9) ? 9 : 8;
    if (key_bits - shift < 8) b = key_bits - shift;
    sort_plan_push(p, shift, b);
    shift += b;
  }
  p.out_bin = (key_bits & 1) ? 1 << (p.bits[p.passes - 1] - 1) : -1;
  p.sorted_bits = key_bits;
  return p;
}

// Top-digit plan: the fewest passes of at most kSortMaxDigitBits-bit digits that cover the top `top_bits` bits of the
// key (at least 8 bits per digit, so the plan may cover a few more; the wider digits at the top).  Falls back to the
// full plan when that is no shorter.  9 bits at most: 10-bit digits were built and measured in round 2 (counters and
// staging buffer sharing their bytes, four digits per thread in the look-back) and dropped — a 1024-bin pass over
// 4096-pair tiles scatters runs of 4 pairs and costs 1.3x an 8-bit pass: build phase at 8M bodies 1.146 vs 1.166 ms
// (3 passes instead of 4), at 64M Plummer 9.19 vs 8.52 ms (4 instead of 5).
constexpr int kSortMaxDigitBits = 9;
inline SortPlan make_top_plan(int key_bits, int top_bits) {
  const SortPlan full = make_sort_plan(key_bits);
  if (top_bits <= 0) return full;
  const int passes = (top_bits + kSortMaxDigitBits - 1) / kSortMaxDigitBits;
  const int cover = top_bits > 8 * passes ? top_bits : 8 * passes;
  if (passes >= full.passes || cover >= key_bits) return full;
  SortPlan p{};
  const int base = cover / passes, extra = cover % passes;
  int shift = key_bits - cover;
  for (int i = 0; i < passes; ++i) {
    const int b = base + (i >= passes - extra ? 1 : 0);
    sort_plan_push(p, shift, b);
    shift += b;
  }
  p.out_bin = (key_bits & 1) ? 1 << (p.bits[p.passes - 1] - 1) : -1;
  p.sorted_bits = cover;
  return p;
}

// All passes' histograms in one read of the keys.  Four independent loads per trip: with one load in flight per
// thread the kernel was bound by DRAM latency (16 dependent trips per thread at 1M bodies), not by its shared-memory
// atomics — run-length aggregation of the (nearly sorted) digits was tried in round 2 and measured slower.
__global__ void __launch_bounds__(256) sort_histogram_kernel(const unsigned long long *__restrict__ keys, uint32_t n,
                                                             SortPlan plan, uint32_t *__restrict__ ghist,
                                                             int first_pass = 0 /* only the passes that will run */) {
  extern __shared__ uint32_t sh[];
  for (int i = threadIdx.x; i < plan.hist_total; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 4u * stride) {
    unsigned long long k[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const uint32_t j = i + (uint32_t)u * stride;
      ok[u] = j < n && j >= i; // (j >= i: no wrap-around)
      k[u] = ok[u] ? keys[j] : 0ull;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (ok[u])
        for (int p = first_pass; p < plan.passes; ++p)
          atomicAdd(&sh[plan.hist_off[p] + ((uint32_t)(k[u] >> plan.shift[p]) & ((1u << plan.bits[p]) - 1u))], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < plan.hist_total; i += blockDim.x)
    if (sh[i]) atomicAdd(&ghist[i], sh[i]);
}

// Exclusive scan per pass (one block per pass); also resets the tile tickets.  The last pass of a
// 49-bit (quadtree) sort has bit 48 = "outside the root square" as its top digit bit, so its exclusive
// prefix at that bit's first bin (SortPlan::out_bin) IS the number of bodies in the tree: the per-step metadata is
// initialised here.
__global__ void __launch_bounds__(512) sort_scan_hist_kernel(uint32_t *__restrict__ ghist, SortPlan plan,
                                                             uint32_t *__restrict__ tickets, Meta *meta, uint32_t n) {
  const int p = blockIdx.x;
  const int radix = 1 << plan.bits[p];
  uint32_t *h = ghist + plan.hist_off[p];
  const uint32_t v = ((int)threadIdx.x < radix) ? h[threadIdx.x] : 0u;
  uint32_t tot;
  const uint32_t ex = block_exclusive_scan<512>(v, tot);
  if ((int)threadIdx.x < radix) h[threadIdx.x] = ex;
  if (threadIdx.x == 0) {
    tickets[p] = 0;
  }
  if (meta && p == plan.passes - 1) {
    if ((int)threadIdx.x == plan.out_bin) meta->n_in = ex;
    if (threadIdx.x == 0) {
      if (plan.out_bin < 0) meta->n_in = n;
      meta->n = n;
      meta->n_nodes = 0;
      meta->n_leaves = 0;
      meta->max_depth = 0;
      meta->overflow_leaves = 0;
      meta->cog_outside = 0;
      meta->resort = 0;
    }
  }
}

template <int RADIX_BITS, int kSortRounds>
__global__ void __launch_bounds__(kSortThreads) sort_onesweep_kernel(
    const unsigned long long *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
    unsigned long long *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint32_t n, int shift,
    const uint32_t *__restrict__ gbase /* scanned histogram of this pass */, uint32_t *ticket,
    volatile uint32_t *status /* [tiles][RADIX], zeroed */, const uint32_t *cond /* run only if null or *cond != 0 */) {
  if (cond && *cond == 0u) return;
  constexpr int RADIX = 1 << RADIX_BITS;
  constexpr int kSortTile = kSortThreads * kSortRounds;
  constexpr int DPT = RADIX / kSortThreads; // digits per thread in the look-back (1 or 2)
  __shared__ uint16_t cnt[kSortWarps][RADIX]; // per-warp digit counts -> exclusive over warps (< 4096)
  __shared__ uint32_t tile_ofs[RADIX];        // exclusive digit offsets inside the tile
  __shared__ uint32_t goff[RADIX];            // global position of the tile's first element of each digit
  __shared__ unsigned long long stage[kSortTile];
  __shared__ uint32_t tile_s;
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  if (threadIdx.x == 0) tile_s = atomicAdd(ticket, 1u);
  for (int d = threadIdx.x; d < RADIX * kSortWarps; d += kSortThreads) (&cnt[0][0])[d] = 0;
  __syncthreads();
  const uint32_t tile = tile_s;
  const uint32_t tbase = tile * kSortTile;
  const uint32_t tcount = (n - tbase < (uint32_t)kSortTile) ? n - tbase : (uint32_t)kSortTile;

  // warp w owns a contiguous chunk of the tile, in rounds of 32 (load order == rank order)
  const uint32_t wbase = tbase + w * (32 * kSortRounds);
  unsigned long long key[kSortRounds];
  uint32_t rank[kSortRounds]; // digit << 16 | rank among equal digits in this warp's chunk
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    const uint32_t i = wbase + k * 32 + lane;
    key[k] = (i < n) ? keys_in[i] : 0ull;
  }
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    const uint32_t i = wbase + k * 32 + lane;
    const bool valid = i < n;
    const uint32_t d = (uint32_t)(key[k] >> shift) & (RADIX - 1);
    const uint32_t m = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
    const uint32_t prev = valid ? cnt[w][d] : 0u;
    __syncwarp();
    const uint32_t r = __popc(m & lt);
    if (valid && r == 0) cnt[w][d] = (uint16_t)(prev + __popc(m));
    __syncwarp();
    rank[k] = (d << 16) | (prev + r);
  }
  __syncthreads();
  // per digit: exclusive scan over the warps, tile total, publish, look back
  uint32_t dtot[DPT];
#pragma unroll
  for (int q = 0; q < DPT; ++q) {
    const int d = threadIdx.x + q * kSortThreads;
    uint32_t run = 0;
#pragma unroll
    for (int ww = 0; ww < kSortWarps; ++ww) {
      const uint32_t t = cnt[ww][d];
      cnt[ww][d] = (uint16_t)run;
      run += t;
    }
    dtot[q] = run;
    status[(size_t)tile * RADIX + d] = (tile == 0 ? kFlagInclusive : kFlagLocal) | run;
  }
  // exclusive scan of the tile totals over the digits -> tile_ofs (thread t owns digits t, t+256)
  {
    uint32_t tot0;
    const uint32_t e0 = block_exclusive_scan<kSortThreads>(dtot[0], tot0);
    tile_ofs[threadIdx.x] = e0;
    if (DPT == 2) {
      uint32_t tot1;
      const uint32_t e1 = block_exclusive_scan<kSortThreads>(dtot[DPT - 1], tot1);
      tile_ofs[threadIdx.x + (DPT - 1) * kSortThreads] = tot0 + e1;
    }
  }
#pragma unroll
  for (int q = 0; q < DPT; ++q) {
    const int d = threadIdx.x + q * kSortThreads;
    uint32_t prefix = 0;
    if (tile > 0) {
      // look back kLookBatch predecessors per round trip: the loads are independent, only the walk over their
      // values is ordered (tile 0 is always inclusive, so the walk ends at the latest there)
      constexpr int kLookBatch = 4; // measured at 1M / 8M bodies, build phase: 1: 0.2985 / 1.412 ms, 4: 0.2855 / 1.403, 8: 0.2950 / 1.433
      int t = (int)tile - 1;
      bool done = false;
      while (!done) {
        uint32_t sv[kLookBatch];
#pragma unroll
        for (int b = 0; b < kLookBatch; ++b) sv[b] = (t - b >= 0) ? status[(size_t)(t - b) * RADIX + d] : 0u;
#pragma unroll
        for (int b = 0; b < kLookBatch; ++b) {
          if (!done) {
            while ((sv[b] >> 30) == 0u) sv[b] = status[(size_t)(t - b) * RADIX + d];
            prefix += sv[b] & kStatusValue;
            done = (sv[b] >> 30) == 2u;
          }
        }
        t -= kLookBatch;
      }
      status[(size_t)tile * RADIX + d] = kFlagInclusive | (prefix + dtot[q]);
    }
    goff[d] = gbase[d] + prefix;
  }
  __syncthreads();
  // stage the keys in output order, then write coalesced runs
  uint32_t pos[kSortRounds];
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    const uint32_t d = rank[k] >> 16;
    pos[k] = tile_ofs[d] + cnt[w][d] + (rank[k] & 0xffffu);
    if (wbase + k * 32 + lane < n) stage[pos[k]] = key[k];
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < tcount; i += kSortThreads) {
    const unsigned long long kk = stage[i];
    const uint32_t d = (uint32_t)(kk >> shift) & (RADIX - 1);
    keys_out[goff[d] + (i - tile_ofs[d])] = kk;
  }
  __syncthreads();
  uint32_t *vstage = reinterpret_cast<uint32_t *>(stage);
  uint32_t *dstage = vstage + kSortTile; // digit of each staged slot (second half of the buffer)
#pragma unroll
  for (int k = 0; k < kSortRounds; ++k) {
    const uint32_t i = wbase + k * 32 + lane;
    if (i < n) {
      vstage[pos[k]] = vals_in[i];
      dstage[pos[k]] = rank[k] >> 16;
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < tcount; i += kSortThreads) {
    const uint32_t d = dstage[i];
    vals_out[goff[d] + (i - tile_ofs[d])] = vstage[i];
  }
}

struct SortScratch {
  uint32_t *ghist;    // [hist_total] all passes' histograms, scanned in place
  uint32_t *tickets;  // [passes]
  uint32_t *status;   // [passes][tiles][radix]
  size_t status_words;
  uint32_t *bsum;     // scratch for the device-wide scan used by the tree build
};

// tile size: 4096 pairs measured faster than 2048 even at 1M pairs (0.389 vs 0.400 ms build on B200);
// 2048 only for small inputs, where it keeps more SMs busy
inline int sort_rounds_for(uint32_t n) { return n < (1u << 19) ? 8 : 16; }

inline size_t sort_status_words(uint32_t n_max) {
  size_t tiles = ((size_t)n_max + 32 * 8 * kSortWarps - 1) / (32 * 8 * kSortWarps);
  if (tiles == 0) tiles = 1;
  return 2 * tiles * (size_t)(5 * 256 + 512 + 256); // up to 7 passes, one of them 9 bits wide; twice: top-digit passes + fallback
}

// One pass of the plan (p) from (kin, vin) to (kout, vout).
inline void sort_enqueue_pass(const SortPlan &plan, int p, int rounds, uint32_t tiles, const unsigned long long *kin,
                              const uint32_t *vin, unsigned long long *kout, uint32_t *vout, uint32_t n,
                              const uint32_t *gb, uint32_t *ticket, uint32_t *status, const uint32_t *cond,
                              cudaStream_t st) {
#define NB_SORT_PASS(RB, RO)                                                                                   \
  sort_onesweep_kernel<RB, RO><<<tiles, kSortThreads, 0, st>>>(kin, vin, kout, vout, n, plan.shift[p], gb, ticket, \
                                                               status, cond)
  if (plan.bits[p] == 9) {
    if (rounds == 8) NB_SORT_PASS(9, 8);
    else NB_SORT_PASS(9, 16);
  } else {
    if (rounds == 8) NB_SORT_PASS(8, 8);
    else NB_SORT_PASS(8, 16);
  }
#undef NB_SORT_PASS
}

// Sorts by the digits of `plan` (stable LSD): make_sort_plan = the whole key, make_top_plan = its top bits only.
// Returns 0 if the result is in (keys_a, vals_a), 1 if in (keys_b, vals_b).
inline int radix_sort_pairs(unsigned long long *keys_a, uint32_t *vals_a, unsigned long long *keys_b,
                            uint32_t *vals_b, uint32_t n, const SortPlan &plan, const SortScratch &sc, cudaStream_t st,
                            Meta *meta = nullptr) {
  if (n == 0) return 0;
  const int rounds = sort_rounds_for(n);
  const uint32_t tile_elems = (uint32_t)(kSortThreads * rounds);
  const uint32_t tiles = (n + tile_elems - 1) / tile_elems;
  const int first = 0;
  size_t need = 0;
  for (int p = first; p < plan.passes; ++p) need += (size_t)tiles << plan.bits[p];
  cudaMemsetAsync(sc.ghist, 0, plan.hist_total * sizeof(uint32_t), st);
  cudaMemsetAsync(sc.status, 0, need * sizeof(uint32_t), st);
  const uint32_t hb = (n + 2047u) / 2048u; // 8 keys per thread, but no more blocks than fit at once
  const uint32_t hblocks = hb < (uint32_t)kSMs * 8u ? hb : (uint32_t)kSMs * 8u;
  g_launches += 2 + (plan.passes - first);
  sort_histogram_kernel<<<hblocks, 256, plan.hist_total * sizeof(uint32_t), st>>>(keys_a, n, plan, sc.ghist, first);
  sort_scan_hist_kernel<<<plan.passes, 512, 0, st>>>(sc.ghist, plan, sc.tickets, meta, n);
  prof_mark(st, "sort_histogram_kernel");
  unsigned long long *kin = keys_a, *kout = keys_b;
  uint32_t *vin = vals_a, *vout = vals_b;
  int which = 0;
  size_t soff = 0;
  for (int p = first; p < plan.passes; ++p) {
    sort_enqueue_pass(plan, p, rounds, tiles, kin, vin, kout, vout, n, sc.ghist + plan.hist_off[p], sc.tickets + p,
                      sc.status + soff, nullptr, st);
    soff += (size_t)tiles << plan.bits[p];
    unsigned long long *tk = kin; kin = kout; kout = tk;
    uint32_t *tv = vin; vin = vout; vout = tv;
    which ^= 1;
    prof_mark(st, "sort_onesweep_kernel");
  }
  return which;
}

} // namespace nb
