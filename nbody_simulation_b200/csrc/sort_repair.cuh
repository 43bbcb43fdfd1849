// sort_repair.cuh — K2b: sorting a state that is already (almost) in tree order.
//
// integrate_resort_kernel leaves the bodies in the order of the tree it integrated: leaf after leaf in Morton order,
// by body id inside a leaf.  At the next build, keys[sorted_in][j] (the previous build's sorted keys) and leafdepth[j]
// still name the LEAF CELL slot j's body was in: its depth L and the top 2L key bits.  A body whose new key still has
// that prefix is a KEEPER — it did not leave its leaf's cell.  The cells of consecutive leaves are disjoint and
// ascending, so the keepers are already in order leaf by leaf; inside a leaf (at most leaf_cap bodies) each keeper
// ranks itself among its leaf-mates by new key.  Only the MOVERS — bodies that crossed the boundary of their leaf's
// cell during the step — need sorting, and the sorted movers are merged into the keepers by counting:
//   keeper in slot j of the leaf that starts at slot a:
//        position = (a - movers in slots < a) + (its rank among the leaf's keepers) + (movers with a smaller key)
//   mover number r, key t:
//        position = (keepers with a key <= t) + r, where the keepers with key <= t are those of every leaf whose cell
//        ends at or below t (a binary search over the old keys finds where they end) plus, if t falls inside an old
//        leaf's cell, that leaf's keepers with key <= t
// ("key" = the prefix the top-digit sort would cover, radix_sort.cuh).  This replaces the histogram + 3..5 onesweep
// passes (100 us of a 220 us build at 1M bodies, 450 us of 1.16 ms at 8M) by one more key and a depth byte read in the
// Morton kernel, a one-block kernel for the movers, and one streaming pass.  The result is a correct sort for ANY
// motion as long as the old keys and leaf depths describe a tree; motion only decides how many movers there are.
// More than kRepairMax of them (or a leaf too large to rank in place: an over-full leaf at the depth cap), and the
// kernels raise the same flag as a depth guess that was too shallow (error bit 1, fail_seq): the step does nothing and
// the host redoes it with the radix sort.  That is the exception: the Morton kernel counts the movers in front of a
// radix sort too, and the host only switches to the repair after a build that saw at most kRepairMax / 2.
// The order among bodies with equal prefixes differs from the radix sort's; the tree does not depend on it
// (tree_finish_kernel orders every leaf by body id), so the tree and everything after it are bit-identical.
#pragma once
#include "common.cuh"
#include "morton.cuh"

namespace nb {

constexpr int kRepairMax = 2048;     // movers one repair takes (12 B each in the merge kernel's shared memory)
constexpr int kRepairLeafScan = 32;  // leaf-mates looked at on either side of a body

struct RepairScratch {        // (the two counters live in Meta: n_movers, n_out; zeroed before the Morton kernel)
  unsigned long long *mkey;  // [kRepairMax] mover p: new key
  uint32_t *mslot;           //              ... its slot
  uint32_t *s_slot;          // [kRepairMax] the movers' slots, ascending
  unsigned long long *t_sorted; //           the movers' new prefixes, ascending (ties by slot), and in that order:
  unsigned long long *k_sorted; //           ... new key
  uint32_t *slot_sorted;        //           ... slot
};

// The leaf cell of old position i: depth L = ldepth[i], prefix = the top 2L bits (and the outside bit) of prev[i].
__device__ __forceinline__ int rp_shift(uint32_t L) { return kKeyBits - 2 * (int)(L > 24u ? 24u : L); }
__device__ __forceinline__ bool rp_same_leaf(unsigned long long pi, uint32_t Li, unsigned long long pj, uint32_t Lj) {
  return Li == Lj && (pi >> rp_shift(Li)) == (pj >> rp_shift(Lj));
}
__device__ __forceinline__ void rp_fail(Meta *meta, uint32_t step_seq) {
  meta->resort = 1u;
  meta->repair_failed = 1u;
  atomicOr(&meta->error, 2u);
  atomicMin(&meta->fail_seq, step_seq);
}

// morton_kernel + the comparison with the leaf cell the slot's body was in at the previous build.
// Also run in front of a radix sort (vals != nullptr: the slot indices the sort permutes; keys may then be the very
// array prev points to — every thread reads its four old keys before it writes its four new ones): the count of movers
// tells the host whether the next builds can repair (nbody_b200.cu).
template <int LEVELS>
__global__ void __launch_bounds__(256) morton_repair_kernel(const float *__restrict__ x, const float *__restrict__ y,
                                                            Meta *meta, uint32_t n, const unsigned long long *prev,
                                                            const uint8_t *__restrict__ ldepth,
                                                            unsigned long long *keys, uint32_t *vals, RepairScratch R) {
  const float x0 = meta->x0, y0 = meta->y0, h = meta->h;
  const uint32_t i4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u; // (no early exit: the warp counts its movers together)
  float px[4], py[4];
  unsigned long long pk[4];
  uint32_t pl[4];
  const bool full = i4 + 3 < n && i4 + 3 >= i4;
  if (full) {
    const float4 vx = *reinterpret_cast<const float4 *>(x + i4);
    const float4 vy = *reinterpret_cast<const float4 *>(y + i4);
    const ulonglong2 p0 = reinterpret_cast<const ulonglong2 *>(prev + i4)[0], p1 = reinterpret_cast<const ulonglong2 *>(prev + i4)[1];
    const uchar4 l4 = *reinterpret_cast<const uchar4 *>(ldepth + i4);
    px[0] = vx.x; px[1] = vx.y; px[2] = vx.z; px[3] = vx.w;
    py[0] = vy.x; py[1] = vy.y; py[2] = vy.z; py[3] = vy.w;
    pk[0] = p0.x; pk[1] = p0.y; pk[2] = p1.x; pk[3] = p1.y;
    pl[0] = l4.x; pl[1] = l4.y; pl[2] = l4.z; pl[3] = l4.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool ok = i4 < n && i4 + k < n;
      px[k] = ok ? x[i4 + k] : 0.f;
      py[k] = ok ? y[i4 + k] : 0.f;
      pk[k] = ok ? prev[i4 + k] : 0ull;
      pl[k] = ok ? ldepth[i4 + k] : 0u;
    }
  }
  unsigned long long kk[4];
  bool mv[4];
  uint32_t mine = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    kk[k] = morton_chain<LEVELS>(px[k], py[k], x0, y0, h);
    if (!in_domain(px[k], py[k], x0, y0, h)) kk[k] = kKeyOut;
    const bool ok = i4 < n && i4 + k < n;
    // (a body outside the root square makes the repair give up; a handful is all anybody needs to know about)
    if (ok && kk[k] == kKeyOut && *reinterpret_cast<volatile uint32_t *>(&meta->n_out) < 64u) atomicAdd(&meta->n_out, 1u);
    const int sh = rp_shift(pl[k]);
    mv[k] = ok && (kk[k] >> sh) != (pk[k] >> sh);
    mine += mv[k] ? 1u : 0u;
  }
  // One atomic per warp that has movers, and none once the count is past what a repair takes: a scene in which every
  // body changes its cell every step (the two-galaxy scene) must not queue a million atomics on one word.
  const unsigned lane = threadIdx.x & 31;
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= (unsigned)o) inc += v;
  }
  const uint32_t wtot = __shfl_sync(0xffffffffu, inc, 31);
  if (wtot != 0u) {
    uint32_t base = 0;
    if (lane == 31) {
      const uint32_t seen = *reinterpret_cast<volatile uint32_t *>(&meta->n_movers);
      base = seen > (uint32_t)kRepairMax ? seen : atomicAdd(&meta->n_movers, wtot);
    }
    base = __shfl_sync(0xffffffffu, base, 31);
    uint32_t p = base + inc - mine;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (mv[k]) {
        if (p < (uint32_t)kRepairMax) {
          R.mkey[p] = kk[k];
          R.mslot[p] = i4 + k;
        }
        ++p;
      }
    }
  }
  if (full) {
    reinterpret_cast<ulonglong2 *>(keys + i4)[0] = make_ulonglong2(kk[0], kk[1]);
    reinterpret_cast<ulonglong2 *>(keys + i4)[1] = make_ulonglong2(kk[2], kk[3]);
    if (vals) *reinterpret_cast<uint4 *>(vals + i4) = make_uint4(i4, i4 + 1, i4 + 2, i4 + 3);
  } else if (i4 < n) {
    for (int k = 0; k < 4 && i4 + k < n; ++k) {
      keys[i4 + k] = kk[k];
      if (vals) vals[i4 + k] = i4 + k;
    }
  }
}

// Bitonic sort of m (a power of two) records in shared memory, ascending by (a, b); c rides along.
__device__ __forceinline__ void repair_bitonic(unsigned long long *a, uint32_t *b, uint32_t *c, uint32_t m) {
  for (uint32_t size = 2; size <= m; size <<= 1) {
    for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (uint32_t t = threadIdx.x; t < (m >> 1); t += blockDim.x) {
        const uint32_t lo = 2u * t - (t & (stride - 1u)), hi = lo + stride;
        const bool asc = (lo & size) == 0u;
        const unsigned long long al = a[lo], ah = a[hi];
        const uint32_t bl = b[lo], bh = b[hi];
        const bool gt = al > ah || (al == ah && bl > bh);
        if (gt == asc) {
          a[lo] = ah; a[hi] = al;
          b[lo] = bh; b[hi] = bl;
          const uint32_t cl = c[lo];
          c[lo] = c[hi]; c[hi] = cl;
        }
      }
    }
  }
  __syncthreads();
}

// One block: the per-step metadata (what sort_scan_hist_kernel initialises on the radix path), then the movers —
// their slots in ascending order, and themselves in key order.
__global__ void __launch_bounds__(1024) repair_movers_kernel(Meta *meta, uint32_t n, int shift, RepairScratch R,
                                                             uint32_t step_seq) {
  __shared__ unsigned long long a[kRepairMax];
  __shared__ uint32_t b[kRepairMax], c[kRepairMax];
  const uint32_t found = meta->n_movers;
  if (threadIdx.x == 0) {
    meta->n = n;
    meta->n_in = n - meta->n_out;
    meta->n_nodes = 0;
    meta->n_leaves = 0;
    meta->max_depth = 0;
    meta->overflow_leaves = 0;
    meta->cog_outside = 0;
    meta->resort = 0;
    // too many movers, or a body outside the root square (its slot has no leaf): the host redoes the step with the radix
    // sort.  And once a step of this batch has failed, every later one must fail too — the state they were enqueued for
    // never came to be, and keys / leaf depths no longer describe one tree (the radix path repeats its failure by itself)
    if (found > (uint32_t)kRepairMax || meta->n_out != 0u || meta->fail_seq != 0xffffffffu) rp_fail(meta, step_seq);
  }
  if (found == 0 || found > (uint32_t)kRepairMax) return;
  const uint32_t M = found;
  uint32_t m2 = 2;
  while (m2 < M) m2 <<= 1;
  // (1) slots, ascending
  for (uint32_t i = threadIdx.x; i < m2; i += blockDim.x) {
    a[i] = i < M ? (unsigned long long)R.mslot[i] : ~0ull;
    b[i] = 0;
    c[i] = 0;
  }
  repair_bitonic(a, b, c, m2);
  for (uint32_t i = threadIdx.x; i < M; i += blockDim.x) R.s_slot[i] = (uint32_t)a[i];
  __syncthreads();
  // (2) new prefixes, ties by slot; c = which mover
  for (uint32_t i = threadIdx.x; i < m2; i += blockDim.x) {
    a[i] = i < M ? (R.mkey[i] >> shift) : ~0ull;
    b[i] = i < M ? R.mslot[i] : 0xffffffffu;
    c[i] = i;
  }
  repair_bitonic(a, b, c, m2);
  for (uint32_t r = threadIdx.x; r < M; r += blockDim.x) {
    R.t_sorted[r] = a[r];
    R.k_sorted[r] = R.mkey[c[r]];
    R.slot_sorted[r] = b[r];
  }
}

// lower bounds in the movers' tables (global memory: at most 24 KB, hot in L1)
__device__ __forceinline__ uint32_t rp_lower_u32(const uint32_t *__restrict__ v, uint32_t m, uint32_t x) {
  uint32_t lo = 0, hi = m;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (__ldg(v + mid) < x) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ uint32_t rp_lower_u64(const unsigned long long *__restrict__ v, uint32_t m,
                                                 unsigned long long x) {
  uint32_t lo = 0, hi = m;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (__ldg(v + mid) < x) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// The streaming pass, a tile of kRepairTile consecutive slots per block (+ a halo of kRepairLeafScan on either side,
// all staged in shared memory, four slots per thread): every keeper to its merged position — and the movers whose key
// falls between the end of the previous tile's last leaf cell and the end of this tile's to theirs (the last tile
// also takes those beyond).  The two counts a keeper needs from the movers' tables are bracketed once per tile: most
// tiles neither lose nor receive a mover, and their keepers then do no search at all.
constexpr int kRepairTile = 1024, kRepairThreads = 256;
__global__ void __launch_bounds__(kRepairThreads) sort_repair_kernel(Meta *meta, uint32_t n,
                                                                     const unsigned long long *__restrict__ prev,
                                                                     const uint8_t *__restrict__ ldepth,
                                                                     const unsigned long long *__restrict__ knew, int shift,
                                                                     RepairScratch R, unsigned long long *__restrict__ kout,
                                                                     uint32_t *__restrict__ vout, uint32_t step_seq) {
  if (meta->error) return;
  constexpr int H = kRepairLeafScan, W = kRepairTile + 2 * kRepairLeafScan;
  __shared__ unsigned long long sp[W], sk[W];
  __shared__ uint8_t sl[W];
  __shared__ uint32_t s_bounds[4]; // movers: slots < j0 - H, slots < tile end; prefixes below / up to the tile's key range
  const uint32_t M = meta->n_movers;
  const uint32_t j0 = blockIdx.x * (uint32_t)kRepairTile;
  for (int t = threadIdx.x; t < W; t += kRepairThreads) {
    const long long g = (long long)j0 - H + t;
    const bool ok = g >= 0 && g < (long long)n;
    sp[t] = ok ? prev[g] : 0ull;
    sk[t] = ok ? knew[g] : 0ull;
    sl[t] = ok ? ldepth[g] : (uint8_t)255; // never the same leaf as a real slot
  }
  __syncthreads();
  const uint32_t cnt = (n - j0 < (uint32_t)kRepairTile) ? n - j0 : (uint32_t)kRepairTile; // slots of this tile
  auto cell_lo = [&](int i) {
    const int sh = rp_shift(sl[i]);
    return (sp[i] >> sh) << sh;
  };
  auto cell_end = [&](int i) {
    const int sh = rp_shift(sl[i]);
    return ((sp[i] >> sh) + 1ull) << sh;
  };
  const bool last_tile = j0 + cnt >= n;
  const unsigned long long e_prev = j0 == 0 ? 0ull : cell_end(H - 1);
  const unsigned long long e_last = cell_end(H + (int)cnt - 1);
  const unsigned long long round_up = (1ull << shift) - 1ull;
  if (M != 0 && threadIdx.x < 4) {
    // keepers of this tile sit in leaves that start at or after slot j0 - H and hold keys in [cell_lo(first), e_last)
    uint32_t v;
    if (threadIdx.x == 0) v = rp_lower_u32(R.s_slot, M, j0 > (uint32_t)H ? j0 - (uint32_t)H : 0u);
    else if (threadIdx.x == 1) v = rp_lower_u32(R.s_slot, M, j0 + cnt);
    else if (threadIdx.x == 2) v = rp_lower_u64(R.t_sorted, M, cell_lo(H) >> shift);
    else v = rp_lower_u64(R.t_sorted, M, (e_last + round_up) >> shift);
    s_bounds[threadIdx.x] = v;
  }
  __syncthreads();
  const uint32_t bs0 = M ? s_bounds[0] : 0u, bs1 = M ? s_bounds[1] : 0u, bt0 = M ? s_bounds[2] : 0u, bt1 = M ? s_bounds[3] : 0u;
  for (uint32_t q = threadIdx.x; q < cnt; q += kRepairThreads) {
    const int me = (int)q + H;
    const unsigned long long key = sk[me], pj = sp[me];
    const uint32_t Lj = sl[me];
    const int sh = rp_shift(Lj);
    if ((key >> sh) != (pj >> sh)) continue; // a mover: placed by the tile that owns its new key
    const unsigned long long t = key >> shift;
    // rank among the keepers of the leaf, by (prefix, slot), and the leaf's first slot
    uint32_t rank = 0;
    int a = me;
    bool too_long = false;
    for (int s = 0; rp_same_leaf(sp[a - 1], sl[a - 1], pj, Lj); --a, ++s) {
      if (s >= H - 1) { too_long = true; break; }
      const unsigned long long ki = sk[a - 1];
      if ((ki >> sh) == (pj >> sh) && (ki >> shift) <= t) ++rank;
    }
    for (int i = me + 1, s = 0; !too_long && rp_same_leaf(sp[i], sl[i], pj, Lj); ++i, ++s) {
      if (s >= H - 1) { too_long = true; break; }
      const unsigned long long ki = sk[i];
      if ((ki >> sh) == (pj >> sh) && (ki >> shift) < t) ++rank;
    }
    if (too_long) {
      rp_fail(meta, step_seq);
      continue;
    }
    const uint32_t ag = j0 + (uint32_t)a - (uint32_t)H; // (a >= 1, and a < H only when j0 >= H)
    // movers in slots < a, movers with a smaller prefix (keepers go first among equals): inside the tile's brackets
    const uint32_t before = bs0 == bs1 ? bs0 : bs0 + rp_lower_u32(R.s_slot + bs0, bs1 - bs0, ag);
    const uint32_t smaller = bt0 == bt1 ? bt0 : bt0 + rp_lower_u64(R.t_sorted + bt0, bt1 - bt0, t);
    const uint32_t p = (ag - before) + rank + smaller;
    kout[p] = key;
    vout[p] = j0 + q;
  }
  if (M == 0) return;
  // the movers this tile places: end of the previous tile's last cell <= key < end of this tile's last cell
  const uint32_t r_lo = rp_lower_u64(R.t_sorted, M, (e_prev + round_up) >> shift); // first prefix with t << shift >= e_prev
  const uint32_t r_hi = last_tile ? M : bt1;
  for (uint32_t r = r_lo + threadIdx.x; r < r_hi; r += kRepairThreads) {
    const unsigned long long t = __ldg(R.t_sorted + r), tk = t << shift;
    int lo = H, hi = H + (int)cnt; // first slot of the tile whose cell ends above tk (cnt: none, the key goes behind all)
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (cell_end(mid) <= tk) lo = mid + 1;
      else hi = mid;
    }
    uint32_t keepers = 0; // of the leaf whose cell holds tk (if one does): those with a prefix <= t
    bool too_long = false;
    if (lo < H + (int)cnt) {
      const unsigned long long p0 = sp[lo];
      const uint32_t L0 = sl[lo];
      const int sh = rp_shift(L0);
      if (((p0 >> sh) << sh) <= tk) {
        for (int i = lo, s = 0; rp_same_leaf(sp[i], sl[i], p0, L0); ++i, ++s) {
          if (s >= H || i >= W - 1) { too_long = true; break; }
          const unsigned long long ki = sk[i];
          if ((ki >> sh) == (p0 >> sh) && (ki >> shift) <= t) ++keepers;
        }
      }
    }
    if (too_long) {
      rp_fail(meta, step_seq);
      continue;
    }
    const uint32_t log = j0 + (uint32_t)(lo - H);
    const uint32_t p = (log - rp_lower_u32(R.s_slot, M, log)) + keepers + r;
    kout[p] = __ldg(R.k_sorted + r);
    vout[p] = __ldg(R.slot_sorted + r);
  }
}

} // namespace nb
