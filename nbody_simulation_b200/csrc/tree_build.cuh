// tree_build.cuh — K3/K4: fully parallel tree construction from sorted Morton keys
// (common-prefix queries + binary searches, in the manner of Karras 2012) and the atomic
// bottom-up mass / centre-of-gravity / bounds up-sweep.
//
// One formulation serves both trees (template BPL = key bits per tree level):
//   BPL=2, uncompressed  -> region quadtree of reference src/quad_tree.rs (fresh-build topology:
//                           a cell is internal iff it holds > cap bodies, single-child chains kept)
//   BPL=1, compressed    -> binary radix tree (LBVH) with bucket leaves <= cap, the B200 variant of
//                           reference src/bvh_tree.rs (tight AABBs, same up-sweep / traversal)
//
// Definitions over the sorted keys k[0..n_in):
//   cpl(i,j)   common prefix length of k[i], k[j] in levels
//   L(i)       leaf depth of body i = min(D, 1 + (cap-th largest cpl(i,j) over j != i))
//              -> the depth-L(i) cell around i holds <= cap bodies, the depth-(L(i)-1) cell > cap
//   c(i)       cpl(i-1,i)  (c(0) = -1)
//   body i starts the nodes at depths c(i)+1 .. L(i) (it is the first body of those cells); the
//   deepest is a leaf.  Ordering nodes by (first body, depth) IS depth-first pre-order, so an
//   exclusive scan of the per-body node counts yields the pre-order index of every node, and
//   skip(node) = index of the first node starting at the node's end body.
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace nb {

struct TreeArrays {
  float4 *box;        // (x0, y0, x1, y1): the four f32 values Rectangle::contains compares with
  // The two records the traversal reads per node are interleaved (32 B per node, one address):
  //   node[2i]   = cm   (cog.x, cog.y, mass as f32, s2 = max(w,h)^2)
  //   node[2i+1] = link (skip, first body, count | leaf<<31, parent)
  float4 *node;
  __device__ __forceinline__ float4 &cm(uint32_t i) const { return node[2 * (size_t)i]; }
  __device__ __forceinline__ uint4 &link(uint32_t i) const {
    return reinterpret_cast<uint4 *>(node)[2 * (size_t)i + 1];
  }
  uint32_t *mass;     // u32 wrapping mass
  int32_t *depth;
  float2 *bmax;       // LBVH: max corner (for the bounds union)
  uint32_t *arrive;   // up-sweep arrival counters (bodies arrived per internal node)
  uint4 *kids;        // quadtree: the (up to 4) children of an internal node in child order, 0xffffffff = none;
                      // lets the up-sweep fetch all child summaries at once instead of chasing skip links
  uint32_t capacity;  // node pool size
};

constexpr uint32_t kLeafBit = 0x80000000u;
constexpr uint32_t kTwigBit = 0x40000000u; // internal node whose children are all leaves
constexpr uint32_t kCountMask = 0x3fffffffu;

// smallest j in (i, hi) whose depth-d prefix differs from k[i]'s; hi if none.  Gallop + bisect.
template <int BPL>
__device__ __forceinline__ uint32_t find_end(const unsigned long long *__restrict__ skeys, uint32_t i,
                                             unsigned long long ki, int d, uint32_t hi) {
  if (d <= 0) return hi;
  const int sh = kKeyBits - d * BPL;
  const unsigned long long p = ki >> sh;
  uint32_t lo = i + 1, step = 1, probe = lo;
  while (probe < hi && (skeys[probe] >> sh) == p) {
    lo = probe + 1;
    probe = lo + step;
    step <<= 1;
  }
  if (probe < hi) hi = probe;
  while (lo < hi) {
    uint32_t mid = (lo + hi) >> 1;
    if ((skeys[mid] >> sh) == p) lo = mid + 1;
    else hi = mid;
  }
  return lo;
}

// L(i) by sliding window: at least cap other bodies share a depth-d prefix with body i  <=>  some window of
// cap+1 CONSECUTIVE sorted bodies that contains i has a common prefix of >= d levels, and a window's common
// prefix is cpl(first, last).  So (cap-th largest cpl(i,j)) = max over j in [i-cap, i] of cpl(k[j], k[j+cap])
// (windows clipped to [0, n_in); none => -1).  A block stages its keys plus a halo of cap on either side in
// shared memory, forms the window prefixes once and every body takes the maximum over its cap+1 windows.
constexpr int kCapMax = 64; // nb_create bounds leaf_cap by this

// Pre-order index of the first node body i starts: the count kernel leaves the exclusive prefix inside each
// 256-body block in nstart[] and the block totals in btotal[], which one scan_bsum_kernel launch turns into block
// bases (and meta->n_nodes); no separate reduce / apply passes over the bodies.
__device__ __forceinline__ uint32_t node_base(const uint32_t *__restrict__ nstart, const uint32_t *__restrict__ bbase,
                                              uint32_t i) {
  return nstart[i] + bbase[i >> 8];
}

// Pass 1: L(i) and the number of nodes body i starts.
template <int BPL, bool COMPRESS>
__global__ void __launch_bounds__(256) tree_count_kernel(const unsigned long long *__restrict__ skeys, Meta *meta,
                                                         int cap, uint8_t *__restrict__ leafdepth,
                                                         uint32_t *__restrict__ nstart,
                                                         uint32_t *__restrict__ btotal, int sorted_levels = 0,
                                                         uint32_t step_seq = 0) {
  // sorted_levels > 0: the keys are sorted by their top `sorted_levels` levels only (radix_sort.cuh).  Everything
  // below is unaffected as long as every leaf is at most that deep, i.e. no window of cap + 1 consecutive bodies
  // shares a prefix of sorted_levels or more; a window that does raises meta->resort and error bit 1: the rest of
  // the step does nothing (the integrate kernels carry the state over unchanged) and the host, which sees the flag
  // at its next synchronisation, redoes the steps from `fail_seq` on with a full sort.
  const uint32_t n_in = meta->n_in;
  const uint32_t i0 = blockIdx.x * blockDim.x;
  const uint32_t i = i0 + threadIdx.x;
  __shared__ unsigned long long s_k[256 + 2 * kCapMax]; // keys of positions i0 - cap .. i0 + 255 + cap
  __shared__ int8_t s_w[256 + kCapMax];                 // common prefix of the window starting at i0 - cap + t
  for (int t = threadIdx.x; t < 256 + 2 * cap; t += 256) {
    const long long p = (long long)i0 - cap + t;
    s_k[t] = (p >= 0 && p < (long long)n_in) ? skeys[p] : 0ull;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < 256 + cap; t += 256) {
    const long long j = (long long)i0 - cap + t;
    const int8_t wp = (j >= 0 && j + cap < (long long)n_in) ? (int8_t)cpl<BPL>(s_k[t], s_k[t + cap]) : (int8_t)-1;
    s_w[t] = wp;
    if (sorted_levels > 0 && wp >= sorted_levels) {
      meta->resort = 1u;
      atomicOr(&meta->error, 2u);
      atomicMin(&meta->fail_seq, step_seq);
    }
  }
  __syncthreads();
  uint32_t cnt = 0, isleaf = 0, over = 0;
  int L = 0, statdepth = 0;
  if (i < n_in) {
    const unsigned long long ki = s_k[threadIdx.x + cap];
    int cc = -1;
    for (int q = 0; q <= cap; ++q) {
      const int w = s_w[threadIdx.x + q];
      cc = w > cc ? w : cc;
    }
    L = cc + 1 > kKeyBits / BPL ? kKeyBits / BPL : cc + 1;
    const int c = (i > 0) ? cpl<BPL>(s_k[threadIdx.x + cap - 1], ki) : -1;
    leafdepth[i] = (uint8_t)L;
    statdepth = L;
    if (c < L) {
      isleaf = 1;
      if (!COMPRESS) {
        cnt = (uint32_t)(L - c);
        if (L == kKeyBits / BPL) { // depth cap: is the leaf over-full?
          uint32_t e = find_end<BPL>(skeys, i, ki, L, n_in);
          over = (e - i > (uint32_t)cap);
        }
      } else {
        int d = c + 1;
        uint32_t e = find_end<BPL>(skeys, i, ki, d, n_in);
        for (;;) {
          int delta = cpl<BPL>(ki, skeys[e - 1]);
          ++cnt;
          if (delta >= L) {
            over = (e - i > (uint32_t)cap);
            statdepth = delta; // recorded leaf depth = tight common prefix of the bucket
            break;
          }
          d = delta + 1;
          e = find_end<BPL>(skeys, i, ki, d, e);
        }
      }
    }
  }
  { // first half of the device-wide scan: exclusive prefix inside the block, block total for scan_bsum_kernel
    uint32_t tot;
    const uint32_t ex = block_exclusive_scan<256>(cnt, tot);
    if (i < n_in) nstart[i] = ex;
    if (threadIdx.x == 0) btotal[blockIdx.x] = tot;
  }
  // block-aggregated statistics
  uint32_t leaves = __syncthreads_count(isleaf);
  uint32_t overs = __syncthreads_count(over);
  __shared__ int smax;
  if (threadIdx.x == 0) smax = 0;
  __syncthreads();
  if (isleaf) atomicMax(&smax, statdepth);
  __syncthreads();
  if (threadIdx.x == 0) {
    if (leaves) atomicAdd(&meta->n_leaves, leaves);
    if (overs) atomicAdd(&meta->overflow_leaves, overs);
    if (smax) atomicMax(&meta->max_depth, (uint32_t)smax);
  }
}

// Pass 2 (after scan_bsum_kernel: block bases in bbase, total in meta->n_nodes): write the nodes.
template <int BPL, bool COMPRESS>
__global__ void __launch_bounds__(256) tree_emit_kernel(const unsigned long long *__restrict__ skeys, Meta *meta,
                                                        const uint8_t *__restrict__ leafdepth,
                                                        const uint32_t *__restrict__ nstart,
                                                        const uint32_t *__restrict__ bbase, TreeArrays T) {
  const uint32_t n_in = meta->n_in, n_nodes = meta->n_nodes;
  if (n_nodes > T.capacity) {
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(&meta->error, 1u);
    return;
  }
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_in) return;
  const unsigned long long ki = skeys[i];
  const int L = leafdepth[i];
  const int c = (i > 0) ? cpl<BPL>(skeys[i - 1], ki) : -1;
  if (c >= L) return;
  uint32_t idx = node_base(nstart, bbase, i);
  static_assert(COMPRESS, "the uncompressed quadtree is emitted by tree_emit_quad_kernel");
  {
    int d = c + 1;
    uint32_t e = find_end<BPL>(skeys, i, ki, d, n_in);
    for (;;) {
      const int delta = cpl<BPL>(ki, skeys[e - 1]);
      const bool leaf = (delta >= L);
      const uint32_t skip = (e < n_in) ? node_base(nstart, bbase, e) : n_nodes;
      T.depth[idx] = delta;
      T.link(idx) = make_uint4(skip, i, (e - i) | (leaf ? kLeafBit : 0u), 0xffffffffu);
      ++idx;
      if (leaf) break;
      d = delta + 1;
      e = find_end<BPL>(skeys, i, ki, d, e);
    }
  }
}

// Pass 2 for the quadtree, node-parallel.  Only ~30 % of the bodies start a node, so one thread per body
// leaves most lanes idle while the starters walk their levels; here a block compacts the nodes its 1024
// bodies start (block scan of the per-body node counts) and hands them out one per thread.  Each node
// walks the reference's f32 midline chain (quad_tree.rs:172-194) from the root to its own level — the same
// operations in the same order as the per-body walk, so the boxes are bit-identical.
constexpr int kEmitBodies = 1024;
template <bool LINK>
__global__ void __launch_bounds__(256) tree_emit_quad_kernel(const unsigned long long *__restrict__ skeys, Meta *meta,
                                                             const uint8_t *__restrict__ leafdepth,
                                                             const uint32_t *__restrict__ nstart,
                                                             const uint32_t *__restrict__ bbase, TreeArrays T) {
  const uint32_t n_in = meta->n_in, n_nodes = meta->n_nodes;
  if (meta->error & 2u) return; // the sort was too shallow for this tree: the host redoes the step
  if (n_nodes > T.capacity) {
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicOr(&meta->error, 1u);
    return;
  }
  __shared__ uint32_t s_excl[kEmitBodies];
  __shared__ int8_t s_c[kEmitBodies];
  __shared__ uint32_t s_warp[8];
  const uint32_t b0 = blockIdx.x * (uint32_t)kEmitBodies;
  if (b0 >= n_in) return;
  const uint32_t j0 = threadIdx.x * 4u;
  uint32_t cnt[4];
  unsigned long long prev = (b0 + j0 > 0 && b0 + j0 - 1 < n_in) ? skeys[b0 + j0 - 1] : 0ull;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const uint32_t i = b0 + j0 + r;
    cnt[r] = 0;
    int c = -1;
    if (i < n_in) {
      const unsigned long long ki = skeys[i];
      const int L = leafdepth[i];
      c = (i > 0) ? cpl<2>(prev, ki) : -1;
      cnt[r] = (c < L) ? (uint32_t)(L - c) : 0u;
      prev = ki;
    }
    s_c[j0 + r] = (int8_t)c;
  }
  const uint32_t mine = cnt[0] + cnt[1] + cnt[2] + cnt[3];
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
    if ((threadIdx.x & 31) >= (unsigned)o) inc += v;
  }
  if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = inc;
  __syncthreads();
  uint32_t woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    const uint32_t v = s_warp[w];
    if ((unsigned)w < (threadIdx.x >> 5)) woff += v;
    total += v;
  }
  uint32_t ex = woff + inc - mine;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    s_excl[j0 + r] = ex;
    ex += cnt[r];
  }
  __syncthreads();
  const float rx = meta->x0, ry = meta->y0, rh = meta->h;
  for (uint32_t t = threadIdx.x; t < total; t += blockDim.x) {
    // owner: the last body whose exclusive count is <= t (bodies starting no node share their successor's value)
    uint32_t lo = 0, hi = kEmitBodies;
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (s_excl[mid] <= t) lo = mid;
      else hi = mid;
    }
    const uint32_t i = b0 + lo, off = t - s_excl[lo];
    const unsigned long long ki = skeys[i];
    const int level = (int)s_c[lo] + 1 + (int)off;
    const bool leaf = (level == (int)leafdepth[i]);
    float ox = rx, oy = ry, hh = rh;
    for (int l = 0; l < level; ++l) {
      hh = __fmul_rn(hh, 0.5f);
      const float mx = __fadd_rn(ox, hh), my = __fadd_rn(oy, hh);
      const uint32_t dg = (uint32_t)(ki >> (kKeyBits - 2 * (l + 1))) & 3u;
      ox = (dg & 1u) ? mx : ox;
      oy = (dg & 2u) ? my : oy;
    }
    const uint32_t e = find_end<2>(skeys, i, ki, level, n_in);
    const uint32_t skip = (e < n_in) ? node_base(nstart, bbase, e) : n_nodes;
    const uint32_t idx = node_base(nstart, bbase, i) + off;
    // An internal node also links its children: child q starts at the first body whose depth-(level+1) prefix
    // differs from child q-1's, and since that body's predecessor lies in a sibling cell the first node it starts
    // IS the child.  The parent writes the children's parent field, its own children table and its twig bit
    // (all children leaves), so no separate pass chases skip links.
    bool twig = false;
    if (LINK && !leaf) {
      uint32_t kid[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
      uint32_t b = i;
      unsigned long long kb = ki;
      bool all_leaf = true;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (b < e) {
          const uint32_t cidx = (q == 0) ? idx + 1 : node_base(nstart, bbase, b);
          kid[q] = cidx;
          all_leaf = all_leaf && ((int)leafdepth[b] == level + 1);
          reinterpret_cast<uint32_t *>(&T.link(cidx))[3] = idx;
          b = find_end<2>(skeys, b, kb, level + 1, e);
          if (b < e) kb = skeys[b];
        }
      }
      T.kids[idx] = make_uint4(kid[0], kid[1], kid[2], kid[3]);
      twig = all_leaf;
    }
    T.depth[idx] = level;
    uint32_t *lk = reinterpret_cast<uint32_t *>(&T.link(idx)); // [3] (parent) belongs to the parent's thread
    *reinterpret_cast<uint2 *>(lk) = make_uint2(skip, i);
    lk[2] = (e - i) | (leaf ? kLeafBit : 0u) | (twig ? kTwigBit : 0u);
    if (!LINK || idx == 0) lk[3] = 0xffffffffu; // !LINK: tree_link_kernel fills it in (and resets arrive[])
    if (LINK) T.arrive[idx] = 0;
    T.box[idx] = make_float4(ox, oy, __fadd_rn(ox, hh), __fadd_rn(oy, hh));
    T.cm(idx) = make_float4(0.f, 0.f, 0.f, __fmul_rn(hh, hh));
  }
}

// Parent links (from the parents' side) + reset of the arrival counters.
template <bool KIDS>
__global__ void __launch_bounds__(256) tree_link_kernel(const Meta *__restrict__ meta, TreeArrays T) {
  if (meta->error) return;
  const uint32_t n_nodes = meta->n_nodes;
  for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < n_nodes; p += gridDim.x * blockDim.x) {
    T.arrive[p] = 0;
    const uint4 l = T.link(p);
    if (l.z & kLeafBit) continue;
    bool all_leaf = true;
    uint32_t kid[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
    int q = 0;
    for (uint32_t c = p + 1; c < l.x;) {
      uint32_t *lc = reinterpret_cast<uint32_t *>(&T.link(c));
      lc[3] = p;
      all_leaf = all_leaf && (lc[2] & kLeafBit);
      if (q < 4) kid[q] = c;
      ++q;
      c = lc[0];
    }
    if (KIDS) T.kids[p] = make_uint4(kid[0], kid[1], kid[2], kid[3]);
    // a "twig": opening it means summing every body below it, in tree order (main.rs:351-363)
    if (all_leaf) reinterpret_cast<uint32_t *>(&T.link(p))[2] = l.z | kTwigBit;
  }
}

// K3c + K4 fused: leaf fix-up, gather and the atomic bottom-up up-sweep, one thread per leaf.
//
// Leaf fix-up: inside every leaf the bodies are put in ascending ORIGINAL id order — the order
// QuadTree::insert fills the slots (quad_tree.rs:163-164,221-224: insertion order, and subdivide
// re-inserts in slot order), which fixes the f32 summation order of the leaf centre of gravity
// (quad_tree.rs:236-240) and of the leaf interactions (main.rs:355-362).  The tree-order body array
// (x, y, mass as f32, mass bits) is gathered in the same pass and the leaf summary is formed from the
// registers.  Then the thread climbs: at every parent it adds its subtree's body count to the parent's
// arrival counter; the thread that completes the count owns the parent and folds the children IN CHILD
// ORDER (deterministic).
//   leaf:     mass = wrapping u32 sum (bvh_tree.rs:123); cog = UNWEIGHTED mean of positions
//             (bvh_tree.rs:104-107, quad_tree.rs:236-240)
//   internal: mass = sum of child masses; cog = sum(child.cog * (child.mass as f32)) / (mass as f32)
//             (quad_tree.rs:253-266 folds from (0,0); bvh_tree.rs:148-154 is a*m0 + b*m1)
//   LBVH bounds: tight AABB with min seeded f32::MAX and max seeded 0 (bvh_tree.rs:42,59),
//             size = max - min, s2 = max(w,h)^2 as main.rs:371 computes it.
__device__ __forceinline__ uint32_t atomic_add_release(uint32_t *p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.add.release.gpu.global.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
// Acquire side of the hand-off, paid only by the thread that completes a parent: the releasing arrivals of the
// other children synchronise with this fence through the arrival counter (PTX memory model: release pattern on
// the atomic, fence.acq_rel after the read-modify-write that observed them), so the children's summaries read
// next are the ones written before those arrivals.
__device__ __forceinline__ void fence_acquire_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

template <bool LBVH> struct LeafAcc {
  float ax = 0.f, ay = 0.f;
  uint32_t mass = 0;
  float mnx = 3.402823466e+38f, mny = 3.402823466e+38f, mxx = 0.f, mxy = 0.f;
  __device__ __forceinline__ void add(float bx, float by, uint32_t mm) {
    ax = __fadd_rn(ax, bx);
    ay = __fadd_rn(ay, by);
    mass += mm;
    if (LBVH) {
      mnx = mnx < bx ? mnx : bx; // minps
      mny = mny < by ? mny : by;
      mxx = mxx > bx ? mxx : bx; // maxps
      mxy = mxy > by ? mxy : by;
    }
  }
};

// REF (NB_TREE_BVH_REF, bvh_ref.cuh): the leaf slices keep the order `partition` left them in
// (bvh_tree.rs:40-54 borrows the slice as is) and a leaf may be EMPTY (cog = 0/0, box = (MAX, -MAX)), so
// parents are completed by counting arrived children, not bodies.
// WARP (NB_TREE_LBVH, bucket leaves of up to 64 bodies): one WARP per node instead of one thread.  The warp
// sorts the leaf's (id, slot) pairs with a 64-element bitonic network over shuffles (two elements per lane),
// gathers and writes the bodies with all lanes, and lane 0 then folds the leaf IN TREE ORDER from a shared
// staging row (the f32 sum is order dependent) and climbs alone.
template <bool LBVH, int MAXK, bool REF = false, bool WARP = false>
__global__ void __launch_bounds__(128, 12) tree_finish_kernel(Meta *__restrict__ meta, TreeArrays T,
                                                          uint32_t *__restrict__ order, const float *__restrict__ x,
                                                          const float *__restrict__ y, const uint32_t *__restrict__ m,
                                                          const uint32_t *__restrict__ id,
                                                          float4 *__restrict__ sbody) {
  if (meta->error) return;
  const uint32_t n_nodes = meta->n_nodes, n_in = meta->n_in, n = meta->n;
  const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  // bodies outside the root square: targets only
  for (uint32_t j = n_in + tid; j < n; j += nth) {
    const uint32_t sl = order[j];
    const uint32_t mm = m[sl];
    sbody[j] = make_float4(x[sl], y[sl], (float)mm, __uint_as_float(mm));
  }
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const uint32_t unit = WARP ? (tid >> 5) : tid, n_units = WARP ? (nth >> 5) : nth;
  __shared__ float s_bx[WARP ? 4 : 1][64], s_by[WARP ? 4 : 1][64];
  __shared__ uint32_t s_bm[WARP ? 4 : 1][64];
  for (uint32_t node = unit; node < n_nodes; node += n_units) {
    const uint4 l = T.link(node);
    if (!(l.z & kLeafBit)) continue;
    const uint32_t s = l.y, k = l.z & kCountMask;
    LeafAcc<LBVH> A;
    auto place = [&](uint32_t j, uint32_t sl) {
      order[s + j] = sl;
      const uint32_t mm = m[sl];
      const float bx = x[sl], by = y[sl];
      sbody[s + j] = make_float4(bx, by, (float)mm, __uint_as_float(mm));
      A.add(bx, by, mm);
    };
    if (WARP && k <= 64) {
      // element e = r * 32 + lane; padding ids sort to the end
      uint32_t slot[2], bid[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const uint32_t e = r * 32 + lane;
        slot[r] = (e < k) ? order[s + e] : 0u;
        bid[r] = (e < k) ? id[slot[r]] : 0xffffffffu;
      }
#pragma unroll
      for (int size = 2; size <= 64; size <<= 1) {
#pragma unroll
        for (int j = size >> 1; j > 0; j >>= 1) {
          if (j == 32) { // size 64, ascending: the partner is this lane's other element
            if (bid[0] > bid[1]) {
              const uint32_t tk = bid[0], tv = slot[0];
              bid[0] = bid[1]; slot[0] = slot[1];
              bid[1] = tk; slot[1] = tv;
            }
          } else {
#pragma unroll
            for (int r = 0; r < 2; ++r) {
              const bool asc = ((r * 32 + lane) & size) == 0, lower = (lane & j) == 0;
              const uint32_t pk = __shfl_xor_sync(0xffffffffu, bid[r], j), pv = __shfl_xor_sync(0xffffffffu, slot[r], j);
              const bool sw = (asc == lower) ? (pk < bid[r]) : (pk > bid[r]);
              bid[r] = sw ? pk : bid[r];
              slot[r] = sw ? pv : slot[r];
            }
          }
        }
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const uint32_t e = r * 32 + lane;
        if (e < k) {
          const uint32_t sl = slot[r], mm = m[sl];
          const float bx = x[sl], by = y[sl];
          order[s + e] = sl;
          sbody[s + e] = make_float4(bx, by, (float)mm, __uint_as_float(mm));
          s_bx[wib][e] = bx; s_by[wib][e] = by; s_bm[wib][e] = mm;
        }
      }
      __syncwarp();
      if (lane != 0) continue; // the other lanes wait at the next leaf's first shuffle
      for (uint32_t j = 0; j < k; ++j) A.add(s_bx[wib][j], s_by[wib][j], s_bm[wib][j]);
    } else if (WARP && lane != 0) {
      continue;
    } else if (REF) {
      for (uint32_t j = 0; j < k; ++j) place(j, order[s + j]);
    } else if (k <= (uint32_t)MAXK) { // registers: bubble network, padding keys sort to the end
      uint32_t slot[MAXK], bid[MAXK];
#pragma unroll
      for (int j = 0; j < MAXK; ++j) {
        slot[j] = ((uint32_t)j < k) ? order[s + j] : 0u;
        bid[j] = ((uint32_t)j < k) ? id[slot[j]] : 0xffffffffu;
      }
#pragma unroll
      for (int a = 0; a < MAXK - 1; ++a) {
#pragma unroll
        for (int b = 0; b < MAXK - 1 - a; ++b) {
          const bool sw = bid[b] > bid[b + 1];
          const uint32_t t0 = sw ? bid[b + 1] : bid[b], t1 = sw ? bid[b] : bid[b + 1];
          const uint32_t u0 = sw ? slot[b + 1] : slot[b], u1 = sw ? slot[b] : slot[b + 1];
          bid[b] = t0; bid[b + 1] = t1;
          slot[b] = u0; slot[b + 1] = u1;
        }
      }
#pragma unroll
      for (int j = 0; j < MAXK; ++j)
        if ((uint32_t)j < k) place((uint32_t)j, slot[j]);
    } else if (k <= 64) {
      uint32_t slot[64], bid[64];
      for (uint32_t j = 0; j < k; ++j) {
        slot[j] = order[s + j];
        bid[j] = id[slot[j]];
      }
      for (uint32_t a = 1; a < k; ++a) { // insertion sort by id
        uint32_t bs = slot[a], bi = bid[a];
        int b = (int)a - 1;
        while (b >= 0 && bid[b] > bi) {
          slot[b + 1] = slot[b];
          bid[b + 1] = bid[b];
          --b;
        }
        slot[b + 1] = bs;
        bid[b + 1] = bi;
      }
      for (uint32_t j = 0; j < k; ++j) place(j, slot[j]);
    } else { // over-full leaf at the depth cap: in-place insertion sort in global memory
      for (uint32_t a = 1; a < k; ++a) {
        uint32_t bs = order[s + a], bi = id[bs];
        int b = (int)a - 1;
        while (b >= 0 && id[order[s + b]] > bi) {
          order[s + b + 1] = order[s + b];
          --b;
        }
        order[s + b + 1] = bs;
      }
      for (uint32_t j = 0; j < k; ++j) place(j, order[s + j]);
    }
    const float kf = (float)k;
    float cx = __fdiv_rn(A.ax, kf), cy = __fdiv_rn(A.ay, kf);
    float s2;
    if (LBVH) {
      const float w = __fsub_rn(A.mxx, A.mnx), h = __fsub_rn(A.mxy, A.mny);
      const float tx = w > h ? w : h, ty = h > w ? h : w;
      s2 = __fmul_rn(tx, ty);
      T.box[node] = make_float4(A.mnx, A.mny, __fadd_rn(A.mnx, w), __fadd_rn(A.mny, h));
      T.bmax[node] = make_float2(A.mxx, A.mxy);
    } else {
      s2 = T.cm(node).w;
    }
    T.cm(node) = make_float4(cx, cy, (float)A.mass, s2);
    T.mass[node] = A.mass;

    uint32_t cur_count = k, parent = l.w;
    while (parent != 0xffffffffu) {
      const uint4 lp = T.link(parent); // written by the emit kernel: in flight while the fence drains
      uint4 kd = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
      if (!LBVH && !REF) kd = T.kids[parent]; // written by the link kernel
      // release: this subtree's summaries are visible at GPU scope before the arrival is.  atom.release is
      // MEMBAR.ALL.GPU + the atomic; __threadfence() would be the stronger MEMBAR.SC.GPU plus an L1 invalidate
      const uint32_t old = atomic_add_release(&T.arrive[parent], REF ? 1u : cur_count);
      const uint32_t total = lp.z & kCountMask;
      if (REF ? (old != 1u) : (old + cur_count != total)) break;
      fence_acquire_gpu(); // the children's summaries below were written before the arrivals this atomic observed
      uint32_t pm = 0;
      float bx = 0.f, by = 0.f;
      bool first = true;
      float mnx = 3.402823466e+38f, mny = 3.402823466e+38f, mxx = 0.f, mxy = 0.f;
      if (!LBVH && !REF) {
        // quadtree: all child summaries are requested at once (child indices from tree_link_kernel), then folded
        // in child order
        const uint32_t kid[4] = {kd.x, kd.y, kd.z, kd.w};
        uint32_t km[4];
        float4 kc[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (kid[q] != 0xffffffffu) {
            km[q] = __ldcg(&T.mass[kid[q]]);
            kc[q] = __ldcg(&T.cm(kid[q]));
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (kid[q] != 0xffffffffu) {
            const float mf = (float)km[q];
            bx = __fadd_rn(bx, __fmul_rn(kc[q].x, mf));
            by = __fadd_rn(by, __fmul_rn(kc[q].y, mf));
            pm += km[q];
          }
        }
      } else
      for (uint32_t c = parent + 1; c < lp.x;) {
        const uint32_t cm_u = __ldcg(&T.mass[c]);
        const float4 ccm = __ldcg(&T.cm(c));
        const float mf = (float)cm_u;
        const float tx = __fmul_rn(ccm.x, mf), ty = __fmul_rn(ccm.y, mf);
        if (LBVH && first) { bx = tx; by = ty; }
        else { bx = __fadd_rn(bx, tx); by = __fadd_rn(by, ty); }
        first = false;
        pm += cm_u;
        if (LBVH) {
          const float4 cb = __ldcg(&T.box[c]);
          const float2 cmax = __ldcg(&T.bmax[c]);
          mnx = mnx < cb.x ? mnx : cb.x;
          mny = mny < cb.y ? mny : cb.y;
          mxx = mxx > cmax.x ? mxx : cmax.x;
          mxy = mxy > cmax.y ? mxy : cmax.y;
        }
        c = T.link(c).x;
      }
      const float pmf = (float)pm;
      cx = __fdiv_rn(bx, pmf);
      cy = __fdiv_rn(by, pmf);
      if (LBVH) {
        const float w = __fsub_rn(mxx, mnx), h = __fsub_rn(mxy, mny);
        const float tx = w > h ? w : h, ty = h > w ? h : w;
        s2 = __fmul_rn(tx, ty);
        T.box[parent] = make_float4(mnx, mny, __fadd_rn(mnx, w), __fadd_rn(mny, h));
        T.bmax[parent] = make_float2(mxx, mxy);
      } else {
        s2 = T.cm(parent).w;
      }
      T.cm(parent) = make_float4(cx, cy, pmf, s2);
      T.mass[parent] = pm;
      { // the traversal may drop Rectangle::contains only if every cog sits in (or within 1% of) its box
        const float4 pb = T.box[parent];
        const float ex = 0.01f * (pb.z - pb.x), ey = 0.01f * (pb.w - pb.y);
        if (!(cx >= pb.x - ex && cx <= pb.z + ex && cy >= pb.y - ey && cy <= pb.w + ey)) meta->cog_outside = 1u;
      }
      cur_count = total;
      parent = lp.w;
    }
  }
}

} // namespace nb
