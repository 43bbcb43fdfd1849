// traverse_group.cuh — K5 (NB_FORCE_FAST, quadtree / LBVH): list-then-evaluate Barnes-Hut traversal.
//
// Reference: World::bvh_sum_gravity (src/main.rs:349-386) per body, opening rule main.rs:370-372, pair force
// main.rs:235-253.  Every body still makes ITS OWN accept / open decision for every node it meets, so each body
// evaluates exactly the interaction set of its private recursive walk (counts equal the scalar walk's); what
// changes against traverse_fast_kernel is who does the deciding and when the forces are evaluated:
//
//  * DECIDE, lane-parallel.  A warp owns 32 tree-order-adjacent bodies and their bounding box.  It pops up to 32
//    (node, lane mask) entries from a shared-memory stack, ONE NODE PER LANE, and classifies each against the box:
//    "every body accepts" when the node passes the opening test for the nearest point of the box, "every body
//    opens" when it fails for the farthest corner.  Both are conservative (1e-5 relative margin against the f32
//    rounding of the per-body test), so they return what each body would have decided.  Only nodes in between
//    ("mixed", ~65 of ~470 per group on the 1M-body disc) take a warp-uniform visit in which every body of the
//    entry's mask runs the reference test itself; the bodies that accept become the item's mask, the others the
//    mask its children inherit.  One classification pass costs about as much as ONE node visit of the
//    depth-first walk and handles 32 nodes.
//  * STAGE.  Accepted nodes and the bodies of opened leaves (an opened twig hands over all its leaves' bodies at
//    once) are copied into two shared-memory lists as (x, y, mass[, mask]) records: F for items every body of
//    the group takes, P for items only some take.
//  * EVALUATE, dense.  F is a plain loop, one broadcast LDS.128 per item and all 32 lanes live.  P is evaluated
//    through per-lane bit words (see eval_partial): a lane runs only over the items of ITS mask.
//
// Summation order: items are added in list order, which is the stack's pop order, a deterministic function of the
// group alone — not the depth-first order of the reference, so results equal the reference's only to rounding
// (tests: <= 1e-5 normwise vs the oracle on the same tree; measured ~1e-7).  NB_FORCE_EXACT keeps the
// depth-first kernel (bit-identical sums), and so does the reference's own tree (NB_TREE_BVH_REF), whose ~2e4
// interactions per body leave no room for a reordering inside 1e-5.
#pragma once
#include "traverse.cuh"

namespace nb {

constexpr int kGrpWarps = 4;    // warps (= groups) per block
constexpr int kGrpStack = 256;  // (node, mask) entries per warp
constexpr int kGrpF = 128;      // full-mask items per warp
constexpr int kGrpP = 128;      // partial-mask items per warp (4 words of 32)
constexpr uint32_t kNoKid = 0xffffffffu;

struct GrpSmem { // per warp: 6.5 KB
  uint2 stack[kGrpStack];
  float4 F[kGrpF];
  float4 P[kGrpP];
  uint32_t words[kGrpP / 32][32]; // words[c][lane]: bit i = lane takes item P[32 c + i]
};

// main.rs:235-253, FAST arithmetic (same instruction sequence as traverse_fast_kernel): np = (-px, -py)
__device__ __forceinline__ void grp_force(const float x, const float y, const float m, const float2 np, float &ax,
                                          float &ay) {
  const float2 d = __fadd2_rn(make_float2(x, y), np);
  const float w = m * rcp_approx(fmaf(fabsf(d.x) + fabsf(d.y), fmaxf(fmaf(d.y, d.y, d.x * d.x), 0.001f), 1e-20f));
  ax = fmaf(d.x, w, ax);
  ay = fmaf(d.y, w, ay);
}

__device__ __forceinline__ void eval_full(const float4 *__restrict__ F, const uint32_t n, const float2 np, float &ax,
                                          float &ay) {
#pragma unroll 4
  for (uint32_t i = 0; i < n; ++i) {
    const float4 it = F[i]; // all lanes read the same address: one broadcast wavefront
    grp_force(it.x, it.y, it.z, np, ax, ay);
  }
}

// P items carry the mask of the lanes that take them in .w: broadcast load, per-lane predicate.
__device__ __forceinline__ void eval_partial(GrpSmem &S, const uint32_t n, const unsigned lane, const float2 np,
                                             float &ax, float &ay) {
  const uint32_t lanebit = 1u << lane;
#pragma unroll 4
  for (uint32_t i = 0; i < n; ++i) {
    const float4 it = S.P[i];
    if (__float_as_uint(it.w) & lanebit) grp_force(it.x, it.y, it.z, np, ax, ay);
  }
}

template <bool COUNT, bool CONTAINS>
__device__ __forceinline__ void traverse_group_body(const float4 *__restrict__ box, const float4 *__restrict__ node,
                                                    const uint4 *__restrict__ kids, Meta *meta,
                                                    const float4 *__restrict__ sbody, float2 *__restrict__ sacc,
                                                    const float theta2, const uint32_t tfirst, const uint32_t tcount,
                                                    GrpSmem &S) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t t0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32u;
  if (t0 >= tcount) return;
  const bool valid = (t0 + lane) < tcount;
  const uint32_t t = tfirst + (valid ? (t0 + lane) : t0);
  const float4 me = sbody[t];
  const float px = me.x, py = me.y;
  const float2 np = make_float2(-px, -py);
  const uint32_t FULL = __ballot_sync(0xffffffffu, valid);
  // bounding box of the group's bodies (fminf / fmaxf drop NaNs: a NaN body decides for itself in the mixed visits
  // or follows the others; its sums are NaN either way, as in the scalar walk)
  float gx0 = px, gx1 = px, gy0 = py, gy1 = py; // invalid lanes duplicate body t0
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    gx0 = fminf(gx0, __shfl_xor_sync(0xffffffffu, gx0, o));
    gx1 = fmaxf(gx1, __shfl_xor_sync(0xffffffffu, gx1, o));
    gy0 = fminf(gy0, __shfl_xor_sync(0xffffffffu, gy0, o));
    gy1 = fmaxf(gy1, __shfl_xor_sync(0xffffffffu, gy1, o));
  }
  const float th_acc = theta2 * 0.99999f, th_open = theta2 * 1.00001f;
  const uint32_t lt = (1u << lane) - 1u;
  float ax = 0.f, ay = 0.f;
  uint32_t n_inter = 0, n_test = 0;
  uint32_t top = 0, fcnt = 0, pcnt = 0;
  if (meta->n_nodes != 0u) {
    if (lane == 0) S.stack[0] = make_uint2(0u, FULL);
    top = 1;
  }
  __syncwarp();
  while (top != 0u) {
    // Pop up to 32 entries, fewer when the stack is nearly full: an entry pushes at most 4 children (net +3), and
    // with one pop per round the walk is depth-first and needs 3 entries per level (<= 24 levels of 2-bit keys,
    // 48 binary levels with net +1): 72 entries of headroom always suffice.
    const int room = (kGrpStack - 72 - (int)top) / 3;
    const uint32_t p = min(min(32u, top), (uint32_t)max(room, 1));
    const bool have = lane < p;
    uint2 e = make_uint2(0u, 0u);
    if (have) e = S.stack[top - 1u - lane];
    top -= p;
    __syncwarp();
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4 l = make_uint4(0u, 0u, 0u, 0u);
    if (have) load_node(node + 2 * (size_t)e.x, c, l);
    const uint32_t mask = e.y;
    uint32_t accm = 0u, openm = 0u, rk = 0u, rmask = 0u;
    bool mixed = false;
    if (have) {
      if (l.z & kLeafBit) { // main.rs:351-363: every body that reaches a leaf sums all its bodies
        rk = l.z & kCountMask;
        rmask = mask;
      } else {
        const float ux = gx0 - c.x, vx = c.x - gx1, uy = gy0 - c.y, vy = c.y - gy1;
        const float nx = fmaxf(fmaxf(ux, vx), 0.f), ny = fmaxf(fmaxf(uy, vy), 0.f); // to the nearest point of the box
        const float fx = fmaxf(-ux, -vx), fy = fmaxf(-uy, -vy);                     // to the farthest corner
        const float dmin2 = fmaf(ny, ny, nx * nx), dmax2 = fmaf(fy, fy, fx * fx);
        bool all_acc = c.w < dmin2 * th_acc;
        if (CONTAINS) { // no body strictly inside the node's box (bvh_tree.rs:15-20) if the boxes do not overlap
          const float4 bx = box[e.x];
          all_acc = all_acc && (bx.z <= gx0 || bx.x >= gx1 || bx.w <= gy0 || bx.y >= gy1);
        }
        const bool all_open = !(c.w < dmax2 * th_open);
        accm = all_acc ? mask : 0u;
        openm = (!all_acc && all_open) ? mask : 0u;
        mixed = !all_acc && !all_open;
        if (COUNT) n_test += __popc(mask);
      }
    }
    // mixed nodes: every body of the entry's mask runs main.rs:370-372 itself (FAST arithmetic)
    uint32_t mb = __ballot_sync(0xffffffffu, mixed);
    while (mb) {
      const int j = __ffs(mb) - 1;
      mb &= mb - 1u;
      const float cx = __shfl_sync(0xffffffffu, c.x, j), cy = __shfl_sync(0xffffffffu, c.y, j);
      const float s2 = __shfl_sync(0xffffffffu, c.w, j);
      const uint32_t mk = __shfl_sync(0xffffffffu, mask, j);
      const float2 d = __fadd2_rn(make_float2(cx, cy), np);
      bool far = s2 < fmaf(d.y, d.y, d.x * d.x) * theta2;
      if (CONTAINS) {
        const float4 bx = box[__shfl_sync(0xffffffffu, e.x, j)];
        far = far && !(py > bx.y && px > bx.x && px < bx.z && py < bx.w);
      }
      const uint32_t b = __ballot_sync(0xffffffffu, far) & mk;
      if ((int)lane == j) {
        accm = b;
        openm = mk & ~b;
      }
    }
    // an opened twig (all children are leaves) hands over every body below it, in tree order
    if (openm != 0u && (l.z & kTwigBit)) {
      rk = l.z & kCountMask;
      rmask = openm;
      openm = 0u;
    }
    if (COUNT) n_inter += __popc(accm) + rk * __popc(rmask);
    { // accepted nodes -> F (every body of the group) or P (some)
      const bool isF = accm == FULL && accm != 0u, isP = accm != 0u && !isF;
      const uint32_t bf = __ballot_sync(0xffffffffu, isF), bp = __ballot_sync(0xffffffffu, isP);
      if (fcnt + 32u > (uint32_t)kGrpF) {
        __syncwarp();
        eval_full(S.F, fcnt, np, ax, ay);
        fcnt = 0;
        __syncwarp();
      }
      if (pcnt + 32u > (uint32_t)kGrpP) {
        __syncwarp();
        eval_partial(S, pcnt, lane, np, ax, ay);
        pcnt = 0;
        __syncwarp();
      }
      if (isF) S.F[fcnt + __popc(bf & lt)] = make_float4(c.x, c.y, c.z, 0.f);
      if (isP) S.P[pcnt + __popc(bp & lt)] = make_float4(c.x, c.y, c.z, __uint_as_float(accm));
      fcnt += __popc(bf);
      pcnt += __popc(bp);
    }
    { // children of opened nodes -> stack, with the mask of the bodies that opened the parent
      uint4 kd = make_uint4(kNoKid, kNoKid, kNoKid, kNoKid);
      if (openm != 0u) kd = kids[e.x];
      const uint32_t kq[4] = {kd.x, kd.y, kd.z, kd.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const bool v = kq[q] != kNoKid;
        const uint32_t b = __ballot_sync(0xffffffffu, v);
        if (v) S.stack[top + __popc(b & lt)] = make_uint2(kq[q], openm);
        top += __popc(b);
      }
    }
    // body ranges (leaves, opened twigs): the warp copies each range, coalesced, into F or P
    uint32_t rb = __ballot_sync(0xffffffffu, rk != 0u);
    while (rb) {
      const int j = __ffs(rb) - 1;
      rb &= rb - 1u;
      const uint32_t first = __shfl_sync(0xffffffffu, l.y, j), k = __shfl_sync(0xffffffffu, rk, j);
      const uint32_t m = __shfl_sync(0xffffffffu, rmask, j);
      for (uint32_t off = 0; off < k; off += 32u) {
        const uint32_t nn = min(32u, k - off);
        if (m == FULL) {
          if (fcnt + 32u > (uint32_t)kGrpF) {
            __syncwarp();
            eval_full(S.F, fcnt, np, ax, ay);
            fcnt = 0;
            __syncwarp();
          }
          if (lane < nn) S.F[fcnt + lane] = sbody[first + off + lane];
          fcnt += nn;
        } else {
          if (pcnt + 32u > (uint32_t)kGrpP) {
            __syncwarp();
            eval_partial(S, pcnt, lane, np, ax, ay);
            pcnt = 0;
            __syncwarp();
          }
          if (lane < nn) {
            float4 b = sbody[first + off + lane];
            b.w = __uint_as_float(m);
            S.P[pcnt + lane] = b;
          }
          pcnt += nn;
        }
      }
    }
    __syncwarp();
  }
  eval_full(S.F, fcnt, np, ax, ay);
  if (pcnt) eval_partial(S, pcnt, lane, np, ax, ay);
  if (valid) sacc[t] = make_float2(ax, ay);
  if (COUNT) {
    unsigned long long wi = n_inter, wt = n_test;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      wi += __shfl_xor_sync(0xffffffffu, wi, o);
      wt += __shfl_xor_sync(0xffffffffu, wt, o);
    }
    if (lane == 0) {
      atomicAdd(&meta->interactions, wi);
      atomicAdd(&meta->node_tests, wt);
    }
  }
}

// ---- hybrid (round 2, second attempt): far field as above, near field by the depth-first walk ---------------------
// The group walk pays for itself in the far field (items every body takes: dense broadcast evaluation, one
// classification per node) and loses in the near field, where the decisions are ragged per body.  So: phase 1 is the
// lane-parallel classification restricted to nodes reached through nodes that EVERY body opens — masks stay full,
// accepted nodes and the bodies of fully opened leaves / twigs go to the F list — and every node on which the bodies
// disagree becomes a ROOT; phase 2 walks each root's subtree with the per-lane depth-first walk of
// traverse_fast_kernel (own decision per lane, `resume` mask).  Per-body decisions, and therefore interaction and
// test counts, are those of the reference walk; sums are in list order, then root by root.
constexpr int kHybStack = 384; // node indices (u32)
constexpr int kHybF = 160;     // F items
constexpr int kHybRoots = 64;  // roots of disagreement awaiting their depth-first walk
struct HybSmem { // per warp: 1.5 + 2.5 + 0.25 KB
  uint32_t stack[kHybStack];
  float4 F[kHybF];
  uint32_t roots[kHybRoots];
};

// the walk of traverse_fast_body over the subtree [cur, end): same instruction sequence, end instead of n_nodes
template <bool COUNT, bool CONTAINS>
__device__ __forceinline__ void hyb_dfs(const float4 *__restrict__ box, const float4 *__restrict__ node,
                                        const float4 *__restrict__ sbody, const float theta2, const float px,
                                        const float py, const float2 np, const bool valid, uint32_t cur,
                                        const uint32_t end, const uint32_t one, float &ax, float &ay, uint32_t &n_inter,
                                        uint32_t &n_test) {
  uint32_t resume = valid ? 0u : 0xffffffffu;
  while (cur < end) {
    float4 c;
    uint4 l;
    load_node(node + 2 * (size_t)cur, c, l);
    const bool active = cur >= resume;
    const float2 dn = __fadd2_rn(make_float2(c.x, c.y), np);
    const float dx = dn.x, dy = dn.y;
    const float d2 = fmaf(dy, dy, dx * dx);
    bool far = c.w < d2 * theta2;
    if (CONTAINS) {
      const float4 bx = box[cur];
      far = far && !(py > bx.y && px > bx.x && px < bx.z && py < bx.w);
    }
    if (l.z & (kLeafBit | kTwigBit)) {
      const bool twig = !(l.z & kLeafBit);
      const bool acc = twig && active && far;
      if (acc) force_from_diff<false>(dx, dy, d2, c.z, ax, ay);
      const bool take = active && !(twig && far);
      const uint32_t k = l.z & kCountMask;
      if (COUNT) {
        n_test += (twig && active) ? 1u : 0u;
        n_inter += (acc ? 1u : 0u) + (take ? k : 0u);
      }
      if (__any_sync(0xffffffffu, take)) {
        uint32_t jb = l.y;
        float lx = ax, ly = ay;
#pragma unroll 1
        for (uint32_t o = k & one; o != 0u; --o) {
          const float4 b0 = sbody[jb++];
          const float2 e0 = __fadd2_rn(make_float2(b0.x, b0.y), np);
          const float dx0 = e0.x, dy0 = e0.y;
          const float w0 = b0.z * rcp_approx(fmaf(fabsf(dx0) + fabsf(dy0), fmaxf(fmaf(dy0, dy0, dx0 * dx0), 0.001f), 1e-20f));
          lx = fmaf(dx0, w0, lx); ly = fmaf(dy0, w0, ly);
        }
        const float4 *bp = sbody + jb;
#pragma unroll 1
        for (uint32_t j = k >> 1; j != 0u; --j, bp += 2) {
          const float4 b0 = bp[0];
          const float4 b1 = bp[1];
          const float2 e0 = __fadd2_rn(make_float2(b0.x, b0.y), np), e1 = __fadd2_rn(make_float2(b1.x, b1.y), np);
          const float dx0 = e0.x, dy0 = e0.y, dx1 = e1.x, dy1 = e1.y;
          const float w0 = b0.z * rcp_approx(fmaf(fabsf(dx0) + fabsf(dy0), fmaxf(fmaf(dy0, dy0, dx0 * dx0), 0.001f), 1e-20f));
          const float w1 = b1.z * rcp_approx(fmaf(fabsf(dx1) + fabsf(dy1), fmaxf(fmaf(dy1, dy1, dx1 * dx1), 0.001f), 1e-20f));
          lx = fmaf(dx0, w0, lx); ly = fmaf(dy0, w0, ly);
          lx = fmaf(dx1, w1, lx); ly = fmaf(dy1, w1, ly);
        }
        ax = take ? lx : ax;
        ay = take ? ly : ay;
      }
      cur = l.x;
    } else {
      const bool acc = active && far;
      if (acc) {
        force_from_diff<false>(dx, dy, d2, c.z, ax, ay);
        resume = l.x;
      }
      if (COUNT) {
        n_test += active ? 1u : 0u;
        n_inter += acc ? 1u : 0u;
      }
      cur = __any_sync(0xffffffffu, active && !far) ? cur + 1 : l.x;
    }
  }
}

template <bool COUNT, bool CONTAINS>
__device__ __forceinline__ void traverse_hybrid_body(const float4 *__restrict__ box, const float4 *__restrict__ node,
                                                     const uint4 *__restrict__ kids, Meta *meta,
                                                     const float4 *__restrict__ sbody, float2 *__restrict__ sacc,
                                                     const float theta2, const uint32_t tfirst, const uint32_t tcount,
                                                     HybSmem &S) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t t0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32u;
  if (t0 >= tcount) return;
  const bool valid = (t0 + lane) < tcount;
  const uint32_t t = tfirst + (valid ? (t0 + lane) : t0);
  const float4 me = sbody[t];
  const float px = me.x, py = me.y;
  const float2 np = make_float2(-px, -py);
  const uint32_t nvalid = __popc(__ballot_sync(0xffffffffu, valid));
  float gx0 = px, gx1 = px, gy0 = py, gy1 = py;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    gx0 = fminf(gx0, __shfl_xor_sync(0xffffffffu, gx0, o));
    gx1 = fmaxf(gx1, __shfl_xor_sync(0xffffffffu, gx1, o));
    gy0 = fminf(gy0, __shfl_xor_sync(0xffffffffu, gy0, o));
    gy1 = fmaxf(gy1, __shfl_xor_sync(0xffffffffu, gy1, o));
  }
  const float th_acc = theta2 * 0.99999f, th_open = theta2 * 1.00001f;
  const uint32_t lt = (1u << lane) - 1u;
  const uint32_t one = 1u | ((meta->one >> 31) << 8);
  float ax = 0.f, ay = 0.f;
  uint32_t n_inter = 0, n_test = 0;   // per lane (phase 2)
  uint32_t u_inter = 0, u_test = 0;   // per group (phase 1): every valid body takes part
  uint32_t top = 0, fcnt = 0, rcnt = 0;
  if (meta->n_nodes != 0u) {
    if (lane == 0) S.stack[0] = 0u;
    top = 1;
  }
  __syncwarp();
  while (top != 0u) {
    const int room = (kHybStack - 72 - (int)top) / 3; // as in traverse_group_body
    const uint32_t p = min(min(32u, top), (uint32_t)max(room, 1));
    const bool have = lane < p;
    uint32_t ni = 0u;
    if (have) ni = S.stack[top - 1u - lane];
    top -= p;
    __syncwarp();
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4 l = make_uint4(0u, 0u, 0u, 0u);
    if (have) load_node(node + 2 * (size_t)ni, c, l);
    bool acc = false, open = false, root = false;
    uint32_t rk = 0u;
    if (have) {
      if (l.z & kLeafBit) {
        rk = l.z & kCountMask;
      } else {
        const float ux = gx0 - c.x, vx = c.x - gx1, uy = gy0 - c.y, vy = c.y - gy1;
        const float nx = fmaxf(fmaxf(ux, vx), 0.f), ny = fmaxf(fmaxf(uy, vy), 0.f);
        const float fx = fmaxf(-ux, -vx), fy = fmaxf(-uy, -vy);
        const float dmin2 = fmaf(ny, ny, nx * nx), dmax2 = fmaf(fy, fy, fx * fx);
        acc = c.w < dmin2 * th_acc;
        if (CONTAINS) {
          const float4 bx = box[ni];
          acc = acc && (bx.z <= gx0 || bx.x >= gx1 || bx.w <= gy0 || bx.y >= gy1);
        }
        open = !acc && !(c.w < dmax2 * th_open);
        root = !acc && !open;
        if (open && (l.z & kTwigBit)) { // every body sums every body below the twig
          rk = l.z & kCountMask;
          open = false;
        }
      }
    }
    if (COUNT) {
      u_test += __popc(__ballot_sync(0xffffffffu, have && !(l.z & kLeafBit) && !root)); // roots are tested in phase 2
      u_inter += __popc(__ballot_sync(0xffffffffu, acc));
    }
    { // accepted by every body -> F
      const uint32_t bf = __ballot_sync(0xffffffffu, acc);
      if (fcnt + 32u > (uint32_t)kHybF) {
        __syncwarp();
        eval_full(S.F, fcnt, np, ax, ay);
        fcnt = 0;
        __syncwarp();
      }
      if (acc) S.F[fcnt + __popc(bf & lt)] = make_float4(c.x, c.y, c.z, 0.f);
      fcnt += __popc(bf);
    }
    { // the bodies disagree -> a root for phase 2
      const uint32_t br = __ballot_sync(0xffffffffu, root);
      if (br) {
        if (rcnt + 32u > (uint32_t)kHybRoots) {
          __syncwarp();
          for (uint32_t r = 0; r < rcnt; ++r) {
            const uint32_t rn = S.roots[r];
            const uint32_t end = __ldg(reinterpret_cast<const uint32_t *>(node + 2 * (size_t)rn + 1));
            hyb_dfs<COUNT, CONTAINS>(box, node, sbody, theta2, px, py, np, valid, rn, end, one, ax, ay, n_inter, n_test);
          }
          rcnt = 0;
          __syncwarp();
        }
        if (root) S.roots[rcnt + __popc(br & lt)] = ni;
        rcnt += __popc(br);
      }
    }
    { // opened by every body -> children
      uint4 kd = make_uint4(kNoKid, kNoKid, kNoKid, kNoKid);
      if (open) kd = kids[ni];
      const uint32_t kq[4] = {kd.x, kd.y, kd.z, kd.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const bool v = kq[q] != kNoKid;
        const uint32_t b = __ballot_sync(0xffffffffu, v);
        if (v) S.stack[top + __popc(b & lt)] = kq[q];
        top += __popc(b);
      }
    }
    uint32_t rb = __ballot_sync(0xffffffffu, rk != 0u);
    while (rb) { // body ranges every body takes -> F (coalesced copies)
      const int j = __ffs(rb) - 1;
      rb &= rb - 1u;
      const uint32_t first = __shfl_sync(0xffffffffu, l.y, j), k = __shfl_sync(0xffffffffu, rk, j);
      if (COUNT) u_inter += k;
      for (uint32_t off = 0; off < k; off += 32u) {
        const uint32_t nn = min(32u, k - off);
        if (fcnt + 32u > (uint32_t)kHybF) {
          __syncwarp();
          eval_full(S.F, fcnt, np, ax, ay);
          fcnt = 0;
          __syncwarp();
        }
        if (lane < nn) S.F[fcnt + lane] = sbody[first + off + lane];
        fcnt += nn;
      }
    }
    __syncwarp();
  }
  eval_full(S.F, fcnt, np, ax, ay);
  for (uint32_t r = 0; r < rcnt; ++r) {
    const uint32_t rn = S.roots[r];
    const uint32_t end = __ldg(reinterpret_cast<const uint32_t *>(node + 2 * (size_t)rn + 1)); // link.x = skip
    hyb_dfs<COUNT, CONTAINS>(box, node, sbody, theta2, px, py, np, valid, rn, end, one, ax, ay, n_inter, n_test);
  }
  if (valid) sacc[t] = make_float2(ax, ay);
  if (COUNT) {
    unsigned long long wi = n_inter, wt = n_test;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      wi += __shfl_xor_sync(0xffffffffu, wi, o);
      wt += __shfl_xor_sync(0xffffffffu, wt, o);
    }
    if (lane == 0) {
      atomicAdd(&meta->interactions, wi + (unsigned long long)u_inter * nvalid);
      atomicAdd(&meta->node_tests, wt + (unsigned long long)u_test * nvalid);
    }
  }
}

template <bool COUNT>
__global__ void __launch_bounds__(kGrpWarps * 32) traverse_hybrid_kernel(const float4 *__restrict__ box,
                                                                         const float4 *__restrict__ node,
                                                                         const uint4 *__restrict__ kids, Meta *meta,
                                                                         const float4 *__restrict__ sbody,
                                                                         float2 *__restrict__ sacc, uint32_t tfirst,
                                                                         uint32_t tcount) {
  if (meta->error) return;
  __shared__ HybSmem smem[kGrpWarps];
  HybSmem &S = smem[threadIdx.x >> 5];
  const float theta2 = meta->theta2;
  node = static_cast<const float4 *>(meta->node_base); // as traverse_fast_kernel: a loaded base stays in its register
  if (!(theta2 <= kThetaSqNoContains) || meta->cog_outside != 0)
    traverse_hybrid_body<COUNT, true>(box, node, kids, meta, sbody, sacc, theta2, tfirst, tcount, S);
  else
    traverse_hybrid_body<COUNT, false>(box, node, kids, meta, sbody, sacc, theta2, tfirst, tcount, S);
}

template <bool COUNT>
__global__ void __launch_bounds__(kGrpWarps * 32) traverse_group_kernel(const float4 *__restrict__ box,
                                                                        const float4 *__restrict__ node,
                                                                        const uint4 *__restrict__ kids, Meta *meta,
                                                                        const float4 *__restrict__ sbody,
                                                                        float2 *__restrict__ sacc, uint32_t tfirst,
                                                                        uint32_t tcount) {
  if (meta->error) return;
  __shared__ GrpSmem smem[kGrpWarps];
  GrpSmem &S = smem[threadIdx.x >> 5];
  const float theta2 = meta->theta2;
  if (!(theta2 <= kThetaSqNoContains) || meta->cog_outside != 0) // see kThetaSqNoContains
    traverse_group_body<COUNT, true>(box, node, kids, meta, sbody, sacc, theta2, tfirst, tcount, S);
  else
    traverse_group_body<COUNT, false>(box, node, kids, meta, sbody, sacc, theta2, tfirst, tcount, S);
}

} // namespace nb
