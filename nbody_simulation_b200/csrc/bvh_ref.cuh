// bvh_ref.cuh — NB_TREE_BVH_REF: the reference's LIVE tree, BVHTree::from (src/bvh_tree.rs:57-96),
// built level-synchronously on the device with the reference's exact topology and body order.
//
// BVHTree::from is a serial top-down recursion: fold (min, max, sum) over the slice, mean = sum/len,
// count bodies above the mean on each axis, pick the axis whose count is closer to len/2
// (`vert > hori` -> x), `partition` the slice in place (predicate-true side first), recurse while a side
// holds > 64 bodies.  Everything in it is order-independent EXCEPT
//   (1) the f32 `sum` fold (bvh_tree.rs:58-61) — evaluated here bit-exactly in parallel, seqfold_core.h;
//   (2) the in-place `partition` (crate partition 0.1.2, two-pointer swap): the k-th predicate-false body
//       from the left of the true zone is swapped with the k-th predicate-true body from the right of the
//       false zone — ranks from one device-wide prefix sum of the predicate flags per level.
// One level of the tree = all nodes of that depth at once:
//   fold (small: one warp per node, big: one block per node and axis) -> count -> decide (children,
//   next level's lists) -> flag scan -> publish misplaced trues -> swap.
// Nodes live in a breadth-first pool; afterwards subtree sizes (atomic bottom-up) give every node its
// depth-first pre-order index and the tree is emitted into the TreeArrays the up-sweep and the
// traversal already use.
#pragma once
#include "common.cuh"
#include "scan.cuh"
#include "seqfold_core.h"
#include "tree_build.cuh"

namespace nb {

constexpr uint32_t kBvhSmall = 4096;    // a node of at most this many bodies is finished, subtree and all, by one
                                        // block in shared memory (bvh_subtree_kernel)
constexpr int kBvhMaxLevels = 96;       // the reference recurses without bound (stack overflow on
                                        // coincident bodies); here the remaining nodes become leaves
constexpr uint32_t kBNodeLeaf = 0x40000000u;
constexpr uint32_t kBNodeAxisX = 0x80000000u;
constexpr uint32_t kBNodeSub = 0x20000000u; // root of a small subtree: not split by the level-synchronous kernels

struct BNode { // 48 B, breadth-first pool
  uint32_t start, len;
  uint32_t parent;  // pool index, 0xffffffff for the root
  uint32_t child0;  // pool index of the predicate-true child; the other is child0 + 1
  float meanx, meany;
  uint32_t cx, cy;  // bodies with x > mean.x / y > mean.y
  uint32_t T;       // size of the predicate-true side
  uint32_t flags;   // kBNodeAxisX | kBNodeLeaf | level
  uint32_t size;    // nodes in the subtree
  uint32_t arrive;  // children whose size is known
};

struct BvhMeta {
  uint32_t n_nodes;
  uint32_t overflow;                 // leaves forced at the level cap
  uint32_t cnt_big[kBvhMaxLevels + 2];   // big nodes to split, per level
  uint32_t cnt_sub;                      // small-subtree roots (any level)
  uint32_t levels;                   // levels actually built (host bookkeeping)
};

struct BvhWork {
  BNode *nodes;
  uint32_t node_cap;
  uint32_t *list_big[2]; // ping-pong by level parity
  uint32_t *list_sub;    // roots of small subtrees
  float *wx, *wy;     // positions in the current (partially partitioned) order
  uint32_t *order;    // position -> storage slot of the body
  uint32_t *seg;      // position -> pool index of the node that currently owns the body
  uint32_t *F;        // exclusive prefix of predicate-false flags
  uint32_t *listT;    // misplaced predicate-true positions, by rank from the right
  uint32_t *bsum;
  BvhMeta *meta;
};

// ---- init: copy positions into the working arrays, root node --------------------------------------
__global__ void __launch_bounds__(256) bvh_init_kernel(BvhWork W, Meta *meta, const float *__restrict__ x,
                                                       const float *__restrict__ y, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    W.wx[i] = x[i];
    W.wy[i] = y[i];
    W.order[i] = i;
    W.seg[i] = 0;
  }
  if (i == 0) {
    BNode r{};
    r.start = 0;
    r.len = n;
    r.parent = 0xffffffffu;
    r.flags = n > kBvhSmall ? 0u : kBNodeSub; // level 0; always a Root (bvh_tree.rs:90), even with <= 64 bodies
    W.nodes[0] = r;
    BvhMeta *m = W.meta;
    m->n_nodes = 1;
    m->overflow = 0;
    for (int l = 0; l < kBvhMaxLevels + 2; ++l) m->cnt_big[l] = 0;
    m->cnt_sub = 0;
    meta->n = n; // per-step metadata (the Morton pipeline initialises it in its sort)
    meta->n_in = n;
    meta->n_nodes = 0;
    meta->n_leaves = 0;
    meta->max_depth = 0;
    meta->overflow_leaves = 0;
    meta->cog_outside = 0;
    if (n > kBvhSmall) { m->cnt_big[0] = 1; W.list_big[0][0] = 0; }
    else { m->cnt_sub = 1; W.list_sub[0] = 0; }
  }
}

// ---- fold, big nodes -------------------------------------------------------------------------------------
// The exact sequential f32 sum of a long slice (seqfold_core.h), four launches per level, tiles of 256 addends:
//   tile_sum      every tile of every big node: its plain fp64 sum (any order), one warp per tile;
//   tile_prefix   per (node, axis): exclusive fp64 prefix of the tile sums.  It tells a tile, to ~1e-4, what the
//                 f32 accumulator will be when the fold reaches it, i.e. its BINADE;
//   tile_summary  the tile's composed parity function for that binade, one warp per tile (8 addends per lane,
//                 warp scan with compose()) — all tiles of all nodes in parallel;
//   resolve       one warp per (node, axis) walks the tiles in order with the true accumulator: a run of 32 tiles
//                 whose guesses hold is applied by one warp scan; a single tile whose guess holds costs O(1); the
//                 others — tile 0 and the ~1 tile per power of two the sum crosses — are folded plainly (256
//                 dependent adds fed by shuffles).  Nothing depends on a guess being right.
constexpr int kFoldTile = 256;
constexpr uint32_t kTileSlotShift = 7; // slot of a node's tile j = (start >> 7) + j: collision-free for len > 4096

struct TileGuess { // 32 B
  int32_t tot[2], lo[2], hi[2];
  uint32_t key; // biased exponent | sign of the guessed accumulator; 0 = no guess
  uint32_t pad;
};
struct TileArrays {
  double *sum[2];     // [axis][slot]: tile sums, then their exclusive prefix per node
  TileGuess *guess[2];
};

__device__ __forceinline__ sf::Summary shfl_up_summary(const sf::Summary &s, int o) {
  sf::Summary t;
#pragma unroll
  for (int p = 0; p < 2; ++p) {
    t.tot[p] = __shfl_up_sync(0xffffffffu, s.tot[p], o);
    t.lo[p] = __shfl_up_sync(0xffffffffu, s.lo[p], o);
    t.hi[p] = __shfl_up_sync(0xffffffffu, s.hi[p], o);
  }
  return t;
}
__device__ __forceinline__ sf::Summary warp_scan_summary(sf::Summary inc) { // inclusive, lane order = sequence order
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const sf::Summary t = shfl_up_summary(inc, o);
    if (lane >= (unsigned)o) inc = sf::compose(t, inc);
  }
  return inc;
}
__device__ __forceinline__ sf::Summary guess_summary(const TileGuess &g) {
  sf::Summary S;
  S.tot[0] = g.tot[0]; S.tot[1] = g.tot[1];
  S.lo[0] = g.lo[0]; S.lo[1] = g.lo[1];
  S.hi[0] = g.hi[0]; S.hi[1] = g.hi[1];
  return S;
}

// work decomposition shared by the tile kernels: blockIdx.x strides the big-node list, the warps of
// (blockIdx.y, threadIdx.x / 32) stride the node's tiles, blockIdx.z is the axis
struct TileIter {
  uint32_t s, len, ntiles, slot0;
};
__device__ __forceinline__ TileIter tile_iter(const BvhWork &W, uint32_t ni) {
  TileIter t;
  t.s = W.nodes[ni].start;
  t.len = W.nodes[ni].len;
  t.ntiles = (t.len + kFoldTile - 1) / kFoldTile;
  t.slot0 = t.s >> kTileSlotShift;
  return t;
}

__global__ void __launch_bounds__(256) bvh_tile_sum_kernel(BvhWork W, TileArrays TA, int level) {
  const uint32_t cnt = W.meta->cnt_big[level];
  const uint32_t *list = W.list_big[level & 1];
  const int dim = blockIdx.z;
  const float *src = dim ? W.wy : W.wx;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t w0 = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5), nw = gridDim.y * (blockDim.x >> 5);
  for (uint32_t a = blockIdx.x; a < cnt; a += gridDim.x) {
    const TileIter t = tile_iter(W, list[a]);
    for (uint32_t j = w0; j + 1 < t.ntiles; j += nw) { // the last tile's sum is never needed
      const float *x = src + t.s + (size_t)j * kFoldTile;
      double v = 0.0;
#pragma unroll 8
      for (int i = 0; i < kFoldTile / 32; ++i) v += (double)x[i * 32 + lane];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) TA.sum[dim][t.slot0 + j] = v;
    }
  }
}

// one warp per (node, axis): tile sums -> exclusive prefix, in place
__global__ void __launch_bounds__(64) bvh_tile_prefix_kernel(BvhWork W, TileArrays TA, int level) {
  const uint32_t cnt = W.meta->cnt_big[level];
  const uint32_t *list = W.list_big[level & 1];
  const int dim = threadIdx.x >> 5;
  const unsigned lane = threadIdx.x & 31;
  for (uint32_t a = blockIdx.x; a < cnt; a += gridDim.x) {
    const TileIter t = tile_iter(W, list[a]);
    double carry = 0.0;
    auto fetch = [&](uint32_t base) { // the next group's sums are in flight while this one is scanned
      const uint32_t j = base + lane;
      return (j + 1 < t.ntiles) ? TA.sum[dim][t.slot0 + j] : 0.0;
    };
    double v = fetch(0);
    for (uint32_t base = 0; base < t.ntiles; base += 32) {
      const uint32_t j = base + lane;
      const double vnext = fetch(base + 32);
      double inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const double u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += u;
      }
      if (j < t.ntiles) TA.sum[dim][t.slot0 + j] = carry + (inc - v);
      carry += __shfl_sync(0xffffffffu, inc, 31);
      v = vnext;
    }
  }
}

__global__ void __launch_bounds__(256) bvh_tile_summary_kernel(BvhWork W, TileArrays TA, int level) {
  __shared__ float stage[8][kFoldTile + kFoldTile / 32]; // index i at i + (i >> 5): spreads the lane-strided reads
  const uint32_t cnt = W.meta->cnt_big[level];
  const uint32_t *list = W.list_big[level & 1];
  const int dim = blockIdx.z;
  const float *src = dim ? W.wy : W.wx;
  const unsigned lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  float *sx = stage[wib];
  const uint32_t w0 = blockIdx.y * (blockDim.x >> 5) + wib, nw = gridDim.y * (blockDim.x >> 5);
  for (uint32_t a = blockIdx.x; a < cnt; a += gridDim.x) {
    const TileIter t = tile_iter(W, list[a]);
    for (uint32_t j = 1 + w0; j < t.ntiles; j += nw) {
      sf::Binade b;
      TileGuess g{};
      if (sf::binade_of(__double2float_rn(TA.sum[dim][t.slot0 + j]), b)) { // warp-uniform
        const uint32_t tlen = min((uint32_t)kFoldTile, t.len - j * kFoldTile);
        const float *x = src + t.s + (size_t)j * kFoldTile;
        __syncwarp();
        for (uint32_t i = lane; i < tlen; i += 32) sx[i + (i >> 5)] = x[i];
        __syncwarp();
        sf::Summary own = sf::identity();
        bool bad = false;
#pragma unroll
        for (uint32_t k = 0; k < (uint32_t)kFoldTile / 32; ++k) {
          const uint32_t i = lane * (kFoldTile / 32) + k;
          if (i < tlen) {
            const sf::QElem q = sf::quantize(sx[i + (i >> 5)], b);
            bad = bad || q.c == 3;
            sf::push_fast(own, q.n, q.c); // a bad addend (c == 3) poisons the chunk below
          }
        }
        if (bad) sf::poison(own);
        const sf::Summary all = warp_scan_summary(own);
        g.tot[0] = all.tot[0]; g.tot[1] = all.tot[1];
        g.lo[0] = all.lo[0]; g.lo[1] = all.lo[1];
        g.hi[0] = all.hi[0]; g.hi[1] = all.hi[1];
        g.key = b.E | b.sign;
      }
      if (lane == 31) TA.guess[dim][t.slot0 + j] = g;
    }
  }
}

// plain sequential fold of x[0..len), len <= kFoldTile, into acc by one warp: the whole tile is fetched into
// registers first (one memory latency, coalesced), then the dependent chain is fed by shuffles
__device__ __forceinline__ float warp_plain_fold(const float *__restrict__ x, uint32_t len, float acc) {
  const unsigned lane = threadIdx.x & 31;
  float v[kFoldTile / 32];
#pragma unroll
  for (int i = 0; i < kFoldTile / 32; ++i) v[i] = (uint32_t)(i * 32) + lane < len ? x[i * 32 + lane] : 0.f;
  if (len == (uint32_t)kFoldTile) {
#pragma unroll
    for (int i = 0; i < kFoldTile / 32; ++i) {
#pragma unroll
      for (int j = 0; j < 32; ++j) acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, v[i], j));
    }
  } else {
#pragma unroll
    for (int i = 0; i < kFoldTile / 32; ++i) {
      const uint32_t c = len > (uint32_t)(i * 32) ? min(32u, len - i * 32) : 0u;
      for (uint32_t j = 0; j < c; ++j) acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, v[i], j));
    }
  }
  return acc;
}

__global__ void __launch_bounds__(64) bvh_fold_resolve_kernel(BvhWork W, TileArrays TA, int level) {
  const uint32_t cnt = W.meta->cnt_big[level];
  const uint32_t *list = W.list_big[level & 1];
  const int dim = threadIdx.x >> 5;
  const unsigned lane = threadIdx.x & 31;
  const float *src = dim ? W.wy : W.wx;
  for (uint32_t a = blockIdx.x; a < cnt; a += gridDim.x) {
    const uint32_t ni = list[a];
    const TileIter t = tile_iter(W, ni);
    auto fetch = [&](uint32_t base) { // the next group's guesses are in flight while this one is applied
      TileGuess g{};
      if (base + lane < t.ntiles) g = TA.guess[dim][t.slot0 + base + lane];
      return g;
    };
    TileGuess gn = fetch(1);
    float acc = warp_plain_fold(src + t.s, min((uint32_t)kFoldTile, t.len), 0.f); // tile 0: a new binade every few addends
    for (uint32_t base = 1; base < t.ntiles; base += 32) {
      const uint32_t j = base + lane;
      const TileGuess g = gn;
      gn = fetch(base + 32);
      const sf::Summary mine = j < t.ntiles ? guess_summary(g) : sf::identity();
      const uint32_t cntg = min(32u, t.ntiles - base);
      sf::Binade b;
      // the whole group at once: every guess names the accumulator's binade and the composition stays inside it
      bool done = false;
      if (sf::binade_of(acc, b)) {
        const bool match = j >= t.ntiles || g.key == (b.E | b.sign);
        if (__all_sync(0xffffffffu, match)) {
          const sf::Summary inc = warp_scan_summary(mine);
          sf::Summary all;
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            all.tot[p] = __shfl_sync(0xffffffffu, inc.tot[p], 31);
            all.lo[p] = __shfl_sync(0xffffffffu, inc.lo[p], 31);
            all.hi[p] = __shfl_sync(0xffffffffu, inc.hi[p], 31);
          }
          if (sf::valid_from(all, b.A)) {
            acc = sf::binade_value(b, b.A + all.tot[b.A & 1]);
            done = true;
          }
        }
      }
      if (done) continue;
      for (uint32_t k = 0; k < cntg; ++k) { // tile by tile
        const uint32_t key = __shfl_sync(0xffffffffu, g.key, k);
        sf::Summary S;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
          S.tot[p] = __shfl_sync(0xffffffffu, mine.tot[p], k);
          S.lo[p] = __shfl_sync(0xffffffffu, mine.lo[p], k);
          S.hi[p] = __shfl_sync(0xffffffffu, mine.hi[p], k);
        }
        if (sf::binade_of(acc, b) && key == (b.E | b.sign) && sf::valid_from(S, b.A)) {
          acc = sf::binade_value(b, b.A + S.tot[b.A & 1]);
        } else {
          const uint32_t jj = base + k;
          acc = warp_plain_fold(src + t.s + (size_t)jj * kFoldTile, min((uint32_t)kFoldTile, t.len - jj * kFoldTile), acc);
        }
      }
    }
    if (lane == 0) {
      const float mean = __fdiv_rn(acc, __uint2float_rn(t.len)); // `points.len() as f32`, bvh_tree.rs:67
      if (dim) { W.nodes[ni].meany = mean; W.nodes[ni].cy = 0; }
      else { W.nodes[ni].meanx = mean; W.nodes[ni].cx = 0; }
    }
  }
}

// A body's node is split at this level?
__device__ __forceinline__ bool bvh_active(const BNode &nd, int level) {
  return nd.flags == (uint32_t)level || nd.flags == ((uint32_t)level | kBNodeAxisX);
}

// ---- count: bodies above the mean on each axis (bvh_tree.rs:70-72) -----------------------------------
__global__ void __launch_bounds__(256) bvh_count_kernel(BvhWork W, uint32_t n, int level) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31;
  uint32_t ni = 0xffffffffu;
  bool gx = false, gy = false;
  if (i < n) {
    ni = W.seg[i];
    const BNode &nd = W.nodes[ni];
    if (nd.flags == (uint32_t)level) {
      gx = W.wx[i] > nd.meanx;
      gy = W.wy[i] > nd.meany;
    } else ni = 0xffffffffu;
  }
  // bodies of a node are contiguous: almost every warp holds one node
  const uint32_t first = __shfl_sync(0xffffffffu, ni, 0);
  if (__all_sync(0xffffffffu, ni == first)) {
    if (first == 0xffffffffu) return;
    const uint32_t bx = __popc(__ballot_sync(0xffffffffu, gx)), by = __popc(__ballot_sync(0xffffffffu, gy));
    if (lane == 0) {
      if (bx) atomicAdd(&W.nodes[first].cx, bx);
      if (by) atomicAdd(&W.nodes[first].cy, by);
    }
  } else if (ni != 0xffffffffu) {
    if (gx) atomicAdd(&W.nodes[ni].cx, 1u);
    if (gy) atomicAdd(&W.nodes[ni].cy, 1u);
  }
}

// ---- decide: axis, split point, children, next level's lists (bvh_tree.rs:69-88) -------------------------
__global__ void __launch_bounds__(256) bvh_decide_kernel(BvhWork W, int level, uint32_t target) {
  BvhMeta *m = W.meta;
  const uint32_t nb = m->cnt_big[level];
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= nb) return;
  const uint32_t ni = W.list_big[level & 1][a];
  BNode &nd = W.nodes[ni];
  const uint32_t len = nd.len, half = len / 2;
  const uint32_t hori = half > nd.cx ? half - nd.cx : nd.cx - half; // abs_diff, bvh_tree.rs:71
  const uint32_t vert = half > nd.cy ? half - nd.cy : nd.cy - half;
  const bool xaxis = vert > hori;                                   // bvh_tree.rs:73
  const uint32_t T = xaxis ? nd.cx : nd.cy;
  const uint32_t c0 = atomicAdd(&m->n_nodes, 2u);
  nd.T = T;
  nd.flags = (uint32_t)level | (xaxis ? kBNodeAxisX : 0u);
  if (c0 + 2 > W.node_cap) { nd.child0 = 0; return; } // pool exhausted: reported by the emit kernel
  nd.child0 = c0;
  for (int k = 0; k < 2; ++k) {
    BNode ch{};
    ch.start = nd.start + (k ? T : 0u);
    ch.len = k ? len - T : T;
    ch.parent = ni;
    const bool internal = ch.len > target; // TARGET_POINTS, bvh_tree.rs:37,78,84
    if (internal && level + 1 < kBvhMaxLevels) {
      ch.flags = (uint32_t)(level + 1);
      if (ch.len > kBvhSmall) W.list_big[(level + 1) & 1][atomicAdd(&m->cnt_big[level + 1], 1u)] = c0 + k;
      else {
        ch.flags |= kBNodeSub;
        W.list_sub[atomicAdd(&m->cnt_sub, 1u)] = c0 + k;
      }
    } else {
      ch.flags = (uint32_t)(level + 1) | kBNodeLeaf;
      ch.size = 1;
      if (internal) atomicAdd(&m->overflow, 1u);
    }
    W.nodes[c0 + k] = ch;
  }
}

// predicate-false flag of position i at this level (0 for bodies whose node is not being split)
__device__ __forceinline__ uint32_t bvh_false_flag(const BvhWork &W, uint32_t i, int level) {
  const BNode &nd = W.nodes[W.seg[i]];
  if ((nd.flags & ~kBNodeAxisX) != (uint32_t)level) return 0u;
  const bool pred = (nd.flags & kBNodeAxisX) ? (W.wx[i] > nd.meanx) : (W.wy[i] > nd.meany);
  return pred ? 0u : 1u;
}

// ---- device-wide exclusive scan of the predicate-false flags (flags evaluated on the fly) --------------
__global__ void __launch_bounds__(kScanThreads) bvh_flag_reduce_kernel(BvhWork W, uint32_t n, int level) {
  const uint32_t i0 = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
  uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k)
    if (i0 + k < n) s += bvh_false_flag(W, i0 + k, level);
  uint32_t tot;
  block_exclusive_scan<kScanThreads>(s, tot);
  if (threadIdx.x == 0) W.bsum[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(kScanThreads) bvh_flag_apply_kernel(BvhWork W, uint32_t n, int level) {
  const uint32_t i0 = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
  uint32_t v[kScanItems], s = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = (i0 + k < n) ? bvh_false_flag(W, i0 + k, level) : 0u;
    s += v[k];
  }
  uint32_t tot;
  uint32_t ex = block_exclusive_scan<kScanThreads>(s, tot) + W.bsum[blockIdx.x];
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    if (i0 + k < n) W.F[i0 + k] = ex;
    ex += v[k];
  }
}

// ---- partition (crate `partition`, bvh_tree.rs:74-76): publish, then swap -------------------------------
// A predicate-true body at position j >= start+T is misplaced; its rank from the right is the number of
// predicate-true bodies after it.  The two-pointer loop swaps it with the misplaced false of equal rank
// from the left.
__global__ void __launch_bounds__(256) bvh_publish_kernel(BvhWork W, uint32_t n, int level) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const BNode &nd = W.nodes[W.seg[i]];
  if ((nd.flags & ~kBNodeAxisX) != (uint32_t)level) return;
  const uint32_t s = nd.start, T = nd.T;
  if (i < s + T) return;
  const bool pred = (nd.flags & kBNodeAxisX) ? (W.wx[i] > nd.meanx) : (W.wy[i] > nd.meany);
  if (!pred) return;
  const uint32_t fb = W.F[i] - W.F[s];     // falses before i inside the node
  const uint32_t tb = (i - s) - fb;        // trues before i
  W.listT[s + (T - tb - 1u)] = i;
}
__global__ void __launch_bounds__(256) bvh_swap_kernel(BvhWork W, uint32_t n, int level) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t ni = W.seg[i];
  const BNode &nd = W.nodes[ni];
  if ((nd.flags & ~kBNodeAxisX) != (uint32_t)level) return;
  const uint32_t s = nd.start, T = nd.T;
  W.seg[i] = nd.child0 + (i < s + T ? 0u : 1u);
  if (i >= s + T) return;
  const float xi = W.wx[i], yi = W.wy[i];
  const bool pred = (nd.flags & kBNodeAxisX) ? (xi > nd.meanx) : (yi > nd.meany);
  if (pred) return;
  const uint32_t j = W.listT[s + (W.F[i] - W.F[s])]; // partner: misplaced true of the same rank
  const uint32_t oi = W.order[i];
  W.wx[i] = W.wx[j]; W.wy[i] = W.wy[j]; W.order[i] = W.order[j];
  W.wx[j] = xi; W.wy[j] = yi; W.order[j] = oi;
}

// ---- small subtrees: one block finishes a node of <= kBvhSmall bodies, all levels below it, in smem -----------
// Same steps as the level-synchronous kernels above (fold -> count -> decide -> flag scan -> publish -> swap),
// over the bodies of one subtree held in shared memory; nodes are appended to the global pool.
constexpr int kSubThreads = 1024; // few subtree roots (n / 4096 .. n / 2048 blocks): wide blocks keep more lanes on each one
constexpr int kSubItems = kBvhSmall / kSubThreads; // 4
constexpr uint16_t kSubSettled = 0xffffu;

struct SubAct { // an internal node of the current local level
  uint32_t gidx;        // index in the global pool
  uint32_t start, len;  // relative to the subtree's first body
  float meanx, meany;
  uint32_t cx, cy, T, xaxis;
  uint32_t slot[2];     // children: index in the next level's list, or kSubSettled
};
struct SubSmem {
  float x[kBvhSmall], y[kBvhSmall];
  uint32_t ord[kBvhSmall];
  uint16_t seg[kBvhSmall]; // body -> index in the current level's list, kSubSettled once its leaf exists
  uint16_t F[kBvhSmall];   // exclusive prefix of predicate-false flags
  uint16_t lt[kBvhSmall];  // misplaced predicate-true bodies by rank from the right
  SubAct act[2][64];       // > 64 bodies per internal node: at most 63 per level
  uint32_t nact[2];
};

__global__ void __launch_bounds__(kSubThreads) bvh_subtree_kernel(BvhWork W, uint32_t target) {
  extern __shared__ __align__(16) unsigned char sub_raw[];
  SubSmem &sm = *reinterpret_cast<SubSmem *>(sub_raw);
  BvhMeta *m = W.meta;
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t nsub = m->cnt_sub;
  for (uint32_t r = blockIdx.x; r < nsub; r += gridDim.x) {
    const uint32_t groot = W.list_sub[r];
    const uint32_t s0 = W.nodes[groot].start, L = W.nodes[groot].len;
    int level = (int)(W.nodes[groot].flags & 0xffffu);
    __syncthreads();
    for (uint32_t i = tid; i < L; i += kSubThreads) {
      sm.x[i] = W.wx[s0 + i];
      sm.y[i] = W.wy[s0 + i];
      sm.ord[i] = W.order[s0 + i];
      sm.seg[i] = 0;
    }
    if (tid == 0) {
      sm.act[0][0].gidx = groot;
      sm.act[0][0].start = 0;
      sm.act[0][0].len = L;
      sm.nact[0] = 1;
      sm.nact[1] = 0;
    }
    __syncthreads();
    int cur = 0;
    for (;; ++level, cur ^= 1) {
      const uint32_t na = sm.nact[cur];
      if (na == 0) break;
      // fold (bvh_tree.rs:58-61,67): one warp per node, literally sequential
      for (uint32_t a = warp; a < na; a += kSubThreads / 32) {
        SubAct &A = sm.act[cur][a];
        const uint32_t s = A.start, len = A.len;
        float ax = 0.f, ay = 0.f;
        for (uint32_t base = 0; base < len; base += 32) {
          const uint32_t c = min(32u, len - base);
          const float vx = lane < c ? sm.x[s + base + lane] : 0.f;
          const float vy = lane < c ? sm.y[s + base + lane] : 0.f;
          if (c == 32) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              ax = __fadd_rn(ax, __shfl_sync(0xffffffffu, vx, j));
              ay = __fadd_rn(ay, __shfl_sync(0xffffffffu, vy, j));
            }
          } else {
            for (uint32_t j = 0; j < c; ++j) {
              ax = __fadd_rn(ax, __shfl_sync(0xffffffffu, vx, j));
              ay = __fadd_rn(ay, __shfl_sync(0xffffffffu, vy, j));
            }
          }
        }
        if (lane == 0) {
          const float lf = __uint2float_rn(len);
          A.meanx = __fdiv_rn(ax, lf);
          A.meany = __fdiv_rn(ay, lf);
          A.cx = 0;
          A.cy = 0;
        }
      }
      if (tid == 0) sm.nact[cur ^ 1] = 0;
      __syncthreads();
      // count (bvh_tree.rs:70-72)
      for (uint32_t i0 = warp * 32; i0 < L; i0 += kSubThreads) { // a warp's 32 bodies mostly share one node
        const uint32_t i = i0 + lane;
        uint32_t a = i < L ? sm.seg[i] : kSubSettled;
        bool gx = false, gy = false;
        if (a != kSubSettled) {
          gx = sm.x[i] > sm.act[cur][a].meanx;
          gy = sm.y[i] > sm.act[cur][a].meany;
        }
        const uint32_t first = __shfl_sync(0xffffffffu, a, 0);
        if (__all_sync(0xffffffffu, a == first)) {
          const uint32_t bx = __popc(__ballot_sync(0xffffffffu, gx)), by = __popc(__ballot_sync(0xffffffffu, gy));
          if (lane == 0 && first != kSubSettled) {
            if (bx) atomicAdd(&sm.act[cur][first].cx, bx);
            if (by) atomicAdd(&sm.act[cur][first].cy, by);
          }
        } else if (a != kSubSettled) {
          if (gx) atomicAdd(&sm.act[cur][a].cx, 1u);
          if (gy) atomicAdd(&sm.act[cur][a].cy, 1u);
        }
      }
      __syncthreads();
      // decide (bvh_tree.rs:69-88)
      if (tid < na) {
        SubAct &A = sm.act[cur][tid];
        const uint32_t len = A.len, half = len / 2;
        const uint32_t hori = half > A.cx ? half - A.cx : A.cx - half;
        const uint32_t vert = half > A.cy ? half - A.cy : A.cy - half;
        const bool xaxis = vert > hori;
        const uint32_t T = xaxis ? A.cx : A.cy;
        A.T = T;
        A.xaxis = xaxis ? 1u : 0u;
        const uint32_t c0 = atomicAdd(&m->n_nodes, 2u);
        const bool room = c0 + 2 <= W.node_cap; // else: reported by the emit kernel (n_nodes > capacity)
        BNode &nd = W.nodes[A.gidx];
        nd.T = T;
        nd.flags = (uint32_t)level | (xaxis ? kBNodeAxisX : 0u);
        nd.child0 = room ? c0 : 0u;
        for (int k = 0; k < 2; ++k) {
          BNode ch{};
          ch.start = s0 + A.start + (k ? T : 0u);
          ch.len = k ? len - T : T;
          ch.parent = A.gidx;
          const bool internal = ch.len > target;
          uint32_t slot = kSubSettled;
          if (internal && level + 1 < kBvhMaxLevels && room) {
            ch.flags = (uint32_t)(level + 1);
            slot = atomicAdd(&sm.nact[cur ^ 1], 1u);
            SubAct &B = sm.act[cur ^ 1][slot];
            B.gidx = c0 + k;
            B.start = A.start + (k ? T : 0u);
            B.len = ch.len;
          } else {
            ch.flags = (uint32_t)(level + 1) | kBNodeLeaf;
            ch.size = 1;
            if (internal && room) atomicAdd(&m->overflow, 1u);
          }
          A.slot[k] = slot;
          if (room) W.nodes[c0 + k] = ch;
        }
      }
      __syncthreads();
      // exclusive scan of the predicate-false flags over the whole subtree
      {
        const uint32_t i0 = tid * kSubItems;
        uint32_t v[kSubItems], ssum = 0;
#pragma unroll
        for (int k = 0; k < kSubItems; ++k) {
          const uint32_t i = i0 + k;
          uint32_t f = 0;
          if (i < L) {
            const uint32_t a = sm.seg[i];
            if (a != kSubSettled) {
              const SubAct &A = sm.act[cur][a];
              const bool pred = A.xaxis ? (sm.x[i] > A.meanx) : (sm.y[i] > A.meany);
              f = pred ? 0u : 1u;
            }
          }
          v[k] = f;
          ssum += f;
        }
        uint32_t tot;
        uint32_t ex = block_exclusive_scan<kSubThreads>(ssum, tot);
#pragma unroll
        for (int k = 0; k < kSubItems; ++k) {
          if (i0 + k < L) sm.F[i0 + k] = (uint16_t)ex;
          ex += v[k];
        }
      }
      __syncthreads();
      // partition (bvh_tree.rs:74-76): publish the misplaced trues, then swap
      for (uint32_t i = tid; i < L; i += kSubThreads) {
        const uint32_t a = sm.seg[i];
        if (a == kSubSettled) continue;
        const SubAct &A = sm.act[cur][a];
        if (i < A.start + A.T) continue;
        const bool pred = A.xaxis ? (sm.x[i] > A.meanx) : (sm.y[i] > A.meany);
        if (!pred) continue;
        const uint32_t fb = (uint32_t)sm.F[i] - (uint32_t)sm.F[A.start];
        const uint32_t tb = (i - A.start) - fb;
        sm.lt[A.start + (A.T - tb - 1u)] = (uint16_t)i;
      }
      __syncthreads();
      for (uint32_t i = tid; i < L; i += kSubThreads) {
        const uint32_t a = sm.seg[i];
        if (a == kSubSettled) continue;
        const SubAct &A = sm.act[cur][a];
        const bool left = i < A.start + A.T;
        if (left) {
          const float xi = sm.x[i], yi = sm.y[i];
          const bool pred = A.xaxis ? (xi > A.meanx) : (yi > A.meany);
          if (!pred) {
            const uint32_t j = sm.lt[A.start + ((uint32_t)sm.F[i] - (uint32_t)sm.F[A.start])];
            const uint32_t oi = sm.ord[i];
            sm.x[i] = sm.x[j]; sm.y[i] = sm.y[j]; sm.ord[i] = sm.ord[j];
            sm.x[j] = xi; sm.y[j] = yi; sm.ord[j] = oi;
          }
        }
        sm.seg[i] = (uint16_t)A.slot[left ? 0 : 1];
      }
      __syncthreads();
    }
    for (uint32_t i = tid; i < L; i += kSubThreads) W.order[s0 + i] = sm.ord[i];
  }
}

// ---- the host enqueued `levels` levels without looking (as many as the previous build needed, plus one): were they
// enough?  If big nodes remain, the rest of this step does nothing (error bit 1, as for a too-shallow sort guess in
// tree_build.cuh) and the host redoes it with the synchronising loop.
__global__ void bvh_levels_check_kernel(const BvhWork W, Meta *meta, int levels, uint32_t step_seq) {
  if (threadIdx.x == 0 && blockIdx.x == 0 && levels <= kBvhMaxLevels && W.meta->cnt_big[levels] != 0u) {
    atomicOr(&meta->error, 2u);
    atomicMin(&meta->fail_seq, step_seq);
  }
}

// ---- subtree sizes, atomic bottom-up -----------------------------------------------------------------
__global__ void __launch_bounds__(256) bvh_sizes_kernel(BvhWork W) {
  const uint32_t nn = min(W.meta->n_nodes, W.node_cap);
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nn; v += gridDim.x * blockDim.x) {
    if (!(W.nodes[v].flags & kBNodeLeaf)) continue;
    uint32_t p = W.nodes[v].parent;
    while (p != 0xffffffffu) {
      // release: this child's size is visible before its arrival; the owner acquires (same protocol as
      // tree_finish_kernel)
      if (atomic_add_release(&W.nodes[p].arrive, 1u) == 0u) break; // the second child to arrive owns the parent
      fence_acquire_gpu();
      const uint32_t c0 = W.nodes[p].child0;
      const uint32_t sz = 1u + __ldcg(&W.nodes[c0].size) + __ldcg(&W.nodes[c0 + 1].size);
      W.nodes[p].size = sz;
      p = W.nodes[p].parent;
    }
  }
}

// ---- pre-order index by climbing to the root, then emit into the traversal's arrays ------------------
// order[] (position -> storage slot) is what the finish kernel gathers through.
__global__ void __launch_bounds__(256) bvh_emit_kernel(BvhWork W, Meta *meta, TreeArrays T, uint32_t n) {
  if (meta->error & 2u) return; // the blind level count was too small: the host redoes the step
  const uint32_t nn = W.meta->n_nodes;
  if (nn > W.node_cap || nn > T.capacity) {
    if (blockIdx.x == 0 && threadIdx.x == 0) { meta->error |= 1u; meta->n_nodes = nn; }
    return;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    meta->n_nodes = nn;
    meta->n_in = n;
    meta->overflow_leaves = W.meta->overflow;
  }
  uint32_t leaves = 0, maxd = 0;
  for (uint32_t v = blockIdx.x * blockDim.x + threadIdx.x; v < nn; v += gridDim.x * blockDim.x) {
    const BNode nd = W.nodes[v];
    uint32_t pre = 0, c = v, p = nd.parent, parent_pre = 0xffffffffu;
    bool first = true;
    uint32_t up = 0; // offset of v from its parent in pre-order
    while (p != 0xffffffffu) {
      const uint32_t pc0 = W.nodes[p].child0;
      const uint32_t off = 1u + (c == pc0 ? 0u : W.nodes[pc0].size);
      if (first) { up = off; first = false; }
      pre += off;
      c = p;
      p = W.nodes[p].parent;
    }
    if (nd.parent != 0xffffffffu) parent_pre = pre - up;
    const bool leaf = (nd.flags & kBNodeLeaf) != 0;
    uint32_t z = nd.len | (leaf ? kLeafBit : 0u);
    if (!leaf) {
      const uint32_t f0 = W.nodes[nd.child0].flags, f1 = W.nodes[nd.child0 + 1].flags;
      if ((f0 & kBNodeLeaf) && (f1 & kBNodeLeaf)) z |= kTwigBit;
    }
    const uint32_t lvl = nd.flags & 0xffffu;
    T.depth[pre] = (int32_t)lvl;
    T.link(pre) = make_uint4(pre + nd.size, nd.start, z, parent_pre);
    T.arrive[pre] = 0;
    leaves += leaf ? 1u : 0u;
    maxd = max(maxd, lvl);
  }
  leaves = __reduce_add_sync(0xffffffffu, leaves);
  maxd = __reduce_max_sync(0xffffffffu, maxd);
  if ((threadIdx.x & 31) == 0) {
    if (leaves) atomicAdd(&meta->n_leaves, leaves);
    if (maxd) atomicMax(&meta->max_depth, maxd);
  }
}

} // namespace nb
