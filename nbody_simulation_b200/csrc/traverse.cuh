// traverse.cuh — K5: warp-cooperative stackless Barnes-Hut traversal with per-body semantics.
//
// Reference: World::bvh_sum_gravity (src/main.rs:349-386) called per body from the rayon
// par_iter (main.rs:406-416).  The reference decides per body whether to accept a node
// (main.rs:370-372).  Here one warp walks the pre-order node array once for 32 Morton-adjacent
// bodies: the warp descends into a node when ANY lane must open it, while every lane keeps its
// own decision — a lane that accepted node A is masked (resume = skip[A]) until the walk leaves
// A's subtree.  Each lane therefore evaluates exactly the node/leaf sequence its private
// recursive walk would, in the same order, so with EXACT arithmetic the result is bit-identical
// to the scalar recursion.
//
// Instruction economy (the kernel is issue-bound, see profiles/):
//  * the opening test and the accepted interaction share diff and |diff|^2: dist2(p, cog)
//    (main.rs:229-232) squares p - cog, calculate_gravity (main.rs:236,245) squares cog - p — the
//    same bits — so an accepted node costs 7 more instructions, not 11;
//  * FAST mode with theta^2 <= 0.49 (kThetaSqNoContains) drops Rectangle::contains: a body strictly inside the box and a
//    centre of gravity inside the box are at most sqrt(w^2+h^2) <= sqrt(2)*s apart, so
//    s^2 < d^2*theta^2 cannot hold; the up-sweep flags the (pathological, u32-wrapped-mass) case
//    of a centre of gravity outside its box and the kernel then keeps the test;
//  * the interaction / opening-test counters are compiled out unless asked for.
#pragma once
#include "common.cuh"
#include "integrate.cuh"
#include "tree_build.cuh"

namespace nb {

// Largest theta^2 for which FAST mode may drop Rectangle::contains.  The up-sweep raises meta.cog_outside unless
// every centre of gravity lies within 1 % of its box (tree_build.cuh), so for a body strictly inside a w x h box
// |dx| < 1.01 w and |dy| < 1.01 h: d^2 < 1.0201 (w^2 + h^2) <= 2.0402 s^2 with s = max(w, h), and the acceptance
// s^2 < d^2 theta^2 cannot hold while 2.0402 theta^2 <= 1, i.e. theta^2 <= 0.49015.  0.49 keeps a 3e-4 relative
// margin over the f32 rounding of d^2 and the product (~3e-7), and covers theta = 0.7 (0.7f^2 = 0.48999998).
constexpr float kThetaSqNoContains = 0.49f;

// transposed few-taker leaf path of traverse_fast_body: ranges shorter than this stay with the plain loop
#ifndef NB_TRANSPOSE_MIN_BODIES
#define NB_TRANSPOSE_MIN_BODIES 6
#endif
constexpr uint32_t kTransposeMinBodies = NB_TRANSPOSE_MIN_BODIES;
#ifndef NB_TRANSPOSE_MAX_TAKERS
#define NB_TRANSPOSE_MAX_TAKERS 16
#endif
constexpr int kTransposeMaxTakers = NB_TRANSPOSE_MAX_TAKERS; // 8: groups of 4 or 8 lanes; 16 adds groups of 2

// main.rs:235-253 with diff = (dx, dy) and d2 = dx*dx + dy*dy already computed.
template <bool EXACT>
__device__ __forceinline__ void force_from_diff(float dx, float dy, float d2, float m, float &ax, float &ay) {
  if (EXACT) {
    const float sum = __fadd_rn(fabsf(dx), fabsf(dy));
    const uint32_t b = __float_as_uint(sum);
    if (b - 0x00800000u < 0x7f000000u) { // sum.is_normal()
      const float d = d2 < 0.001f ? 0.001f : d2;
      const float den = __fmul_rn(sum, d);
      ax = __fadd_rn(ax, __fdiv_rn(__fmul_rn(dx, m), den));
      ay = __fadd_rn(ay, __fdiv_rn(__fmul_rn(dy, m), den));
    }
  } else {
    const float sum = fabsf(dx) + fabsf(dy);
    const float d = fmaxf(d2, 0.001f);
    const float w = m * rcp_approx(fmaf(sum, d, 1e-20f));
    ax = fmaf(dx, w, ax);
    ay = fmaf(dy, w, ay);
  }
}

// One 256-bit load of the interleaved node record (cm, link): sm_100 has ld.global.v8.
__device__ __forceinline__ void load_node(const float4 *p, float4 &c, uint4 &l) {
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w), "=r"(l.x), "=r"(l.y), "=r"(l.z), "=r"(l.w)
      : "l"(p));
}

template <bool EXACT, bool COUNT>
__global__ void __launch_bounds__(256) traverse_kernel(const float4 *__restrict__ box,
                                                       const float4 *__restrict__ node /* [2i]=cm, [2i+1]=link */,
                                                       Meta *meta,
                                                       const float4 *__restrict__ sbody, float2 *__restrict__ sacc,
                                                       float theta, uint32_t tfirst, uint32_t tcount) {
  if (meta->error) return;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t wg = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t t0 = wg * 32u;
  if (t0 >= tcount) return;
  const bool valid = (t0 + lane) < tcount;
  const uint32_t t = tfirst + (valid ? (t0 + lane) : t0);
  const float4 me = sbody[t];
  const float px = me.x, py = me.y;
  float ax = 0.f, ay = 0.f;
  uint32_t resume = valid ? 0u : 0xffffffffu; // lane is active at node `cur` iff cur >= resume
  uint32_t n_inter = 0, n_test = 0;
  const uint32_t n_nodes = meta->n_nodes;
  const float theta2 = theta * theta;
  // Rectangle::contains is needed unless provably implied (see header)
  const bool need_contains = EXACT || !(theta2 <= kThetaSqNoContains) || meta->cog_outside != 0;
  uint32_t cur = 0;
  while (cur < n_nodes) {
    float4 c;
    uint4 l;
    load_node(node + 2 * (size_t)cur, c, l);
    const bool active = cur >= resume;
    if (l.z & (kLeafBit | kTwigBit)) {
      // terminal visit: a leaf (summed by every active lane, main.rs:351-363), or a "twig" — an internal
      // node whose children are all leaves: lanes that accept it take its centre of gravity, lanes that
      // open it sum every body below it in tree order; nobody descends, so the leaf children are skipped.
      bool take = active;
      if (!(l.z & kLeafBit)) {
        float dx, dy, d2;
        bool far;
        if (EXACT) {
          dx = __fsub_rn(c.x, px);
          dy = __fsub_rn(c.y, py);
          d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
          far = c.w < __fmul_rn(__fmul_rn(d2, theta), theta);
        } else {
          dx = c.x - px;
          dy = c.y - py;
          d2 = fmaf(dy, dy, dx * dx);
          far = c.w < d2 * theta2;
        }
        if (need_contains) {
          const float4 bx = box[cur];
          const bool contains = py > bx.y && px > bx.x && px < bx.z && py < bx.w;
          far = far && !contains;
        }
        if (active && far) force_from_diff<EXACT>(dx, dy, d2, c.z, ax, ay);
        if (COUNT) {
          n_test += active ? 1u : 0u;
          n_inter += (active && far) ? 1u : 0u;
        }
        take = active && !far;
      }
      const uint32_t k = l.z & kCountMask;
      const float4 *bp = sbody + l.y;
      if (take) {
#pragma unroll 4
        for (uint32_t j = 0; j < k; ++j) {
          const float4 b = bp[j];
          pair_force<EXACT>(px, py, b.x, b.y, b.z, ax, ay);
        }
      }
      if (COUNT && take) n_inter += k;
      cur = l.x; // == cur + 1 for a leaf
    } else {
      float dx, dy, d2;
      bool far;
      if (EXACT) {
        dx = __fsub_rn(c.x, px);
        dy = __fsub_rn(c.y, py);
        d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
        far = c.w < __fmul_rn(__fmul_rn(d2, theta), theta); // main.rs:371-372, left to right
      } else {
        dx = c.x - px;
        dy = c.y - py;
        d2 = fmaf(dy, dy, dx * dx);
        far = c.w < d2 * theta2;
      }
      if (need_contains) {
        const float4 bx = box[cur];
        const bool contains = py > bx.y && px > bx.x && px < bx.z && py < bx.w;
        far = far && !contains;
      }
      const bool acc = active && far;
      if (__any_sync(0xffffffffu, acc)) {
        if (acc) {
          force_from_diff<EXACT>(dx, dy, d2, c.z, ax, ay);
          resume = l.x;
        }
      }
      if (COUNT) {
        n_test += active ? 1u : 0u;
        n_inter += acc ? 1u : 0u;
      }
      cur = __any_sync(0xffffffffu, active && !far) ? cur + 1 : l.x;
    }
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
      atomicAdd(&meta->interactions, wi);
      atomicAdd(&meta->node_tests, wt);
    }
  }
}

// FAST-mode kernel (NB_FORCE_FAST): same walk, same per-lane decisions, written for issue slots.
//  * accepted interactions are plain predicated code (no vote guard around them);
//  * a terminal visit runs a warp-uniform body loop, an odd first body and then two bodies per trip; every lane
//    continues its running sum through the leaf in a copy and only the lanes that take the leaf (not masked,
//    did not accept the twig) keep the copy — no divergent branch, no per-body predicate or select;
//  * summation order inside a lane is unchanged (tree order): with ~2e4 interactions per body on the
//    reference's own tree, any reordering shows up as ~1e-5 of rounding noise against the reference.
template <bool COUNT, bool CONTAINS, bool TRANSPOSE = false, bool FUSE = false>
__device__ __forceinline__ void traverse_fast_body(const float4 *__restrict__ box, const float4 *__restrict__ node,
                                                   Meta *meta, const float4 *__restrict__ sbody,
                                                   float2 *__restrict__ sacc, const float theta2, uint32_t tfirst,
                                                   uint32_t tcount, const FuseArgs *fuse = nullptr) {
  const unsigned lane = threadIdx.x & 31;
  __shared__ unsigned char s_src[TRANSPOSE ? 8 : 1][16]; // transposed leaf path: lane of the q-th taker, per warp
  const uint32_t t0 = ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 32u;
  if (t0 >= tcount) return;
  const bool valid = (t0 + lane) < tcount;
  const uint32_t t = tfirst + (valid ? (t0 + lane) : t0);
  const float4 me = sbody[t];
  const float px = me.x, py = me.y;
  float ax = 0.f, ay = 0.f;
  uint32_t resume = valid ? 0u : 0xffffffffu;
  uint32_t n_inter = 0, n_test = 0;
  const uint32_t n_nodes = meta->n_nodes;
  // The odd first body of a leaf must stay behind a (warp-uniform) branch: ptxas if-converts `if (k & 1)` into 16
  // predicated issue slots on every even leaf.  `one` is 1 (Meta::one is 0 or 1: its top bit is clear), but only the
  // host knows: as far as ptxas can tell `k & one` may exceed 1, so the loop over it stays a real loop.
  const uint32_t one = 1u | ((meta->one >> 31) << 8);
  // The same goes for the node array's base (Meta::node_base holds the value of the `node` argument): as a kernel
  // parameter it costs an LDC.64 per node visit.  (Doing it for `sbody` too tips ptxas into running the walk's
  // bookkeeping on the uniform datapath, with R2UR / UR-to-R moves in the leaf loop: measured by instruction count,
  // worse.)
  node = static_cast<const float4 *>(meta->node_base);
  const float2 np = make_float2(-px, -py);
  uint32_t cur = 0;
  while (cur < n_nodes) {
    float4 c;
    uint4 l;
    load_node(node + 2 * (size_t)cur, c, l);
    const bool active = cur >= resume;
    const float2 dn = __fadd2_rn(make_float2(c.x, c.y), np); // one packed FADD2 for both components (same bits as c - p)
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
      const unsigned tkm = __ballot_sync(0xffffffffu, take);
      if (TRANSPOSE && tkm != 0u && __popc(tkm) <= kTransposeMaxTakers && k >= kTransposeMinBodies) {
        // Few takers (39 % of the leaf-loop trips run with <= 8 of 32 lanes taking, 22 % with <= 4): transpose.  Taker
        // number q gets a group of 4 lanes (8 when at most 4 lanes take), each lane sums every 4th (8th) body of the
        // range for it, shuffles reduce the partial sums and one more hands the result back: ceil(k / 4) or
        // ceil(k / 8) trips instead of k / 2 with most lanes idle.  The partial sums re-order the lane's additions:
        // FAST mode on the Morton trees only (the reference's own tree keeps the plain loop, see the kernel's header).
        const unsigned nt = __popc(tkm);
        const unsigned myq = __popc(tkm & ((1u << lane) - 1u));
        if (take) s_src[threadIdx.x >> 5][myq] = (unsigned char)lane;
        __syncwarp();
        const unsigned sh = nt <= 4u ? 3u : (nt <= 8u ? 2u : 1u); // log2(lanes per taker)
        const unsigned q = lane >> sh, r = lane & ((1u << sh) - 1u), step = 1u << sh;
        const unsigned src = q < nt ? s_src[threadIdx.x >> 5][q] : lane;
        const float2 tp = make_float2(__shfl_sync(0xffffffffu, np.x, src), __shfl_sync(0xffffffffu, np.y, src));
        float sx = 0.f, sy = 0.f;
        const float4 *bp = sbody + l.y + r; // a running pointer: one IMAD.WIDE per trip instead of four address instructions
#pragma unroll 1
        for (uint32_t j = r; j < k; j += step, bp += step) {
          const float4 b0 = *bp;
          const float2 e0 = __fadd2_rn(make_float2(b0.x, b0.y), tp);
          const float w0 = b0.z * rcp_approx(fmaf(fabsf(e0.x) + fabsf(e0.y), fmaxf(fmaf(e0.y, e0.y, e0.x * e0.x), 0.001f), 1e-20f));
          sx = fmaf(e0.x, w0, sx); sy = fmaf(e0.y, w0, sy);
        }
        sx += __shfl_xor_sync(0xffffffffu, sx, 1); sy += __shfl_xor_sync(0xffffffffu, sy, 1);
        if (sh >= 2u) { // warp-uniform
          sx += __shfl_xor_sync(0xffffffffu, sx, 2); sy += __shfl_xor_sync(0xffffffffu, sy, 2);
        }
        if (sh == 3u) {
          sx += __shfl_xor_sync(0xffffffffu, sx, 4); sy += __shfl_xor_sync(0xffffffffu, sy, 4);
        }
        const float gx = __shfl_sync(0xffffffffu, sx, (myq << sh) & 31u), gy = __shfl_sync(0xffffffffu, sy, (myq << sh) & 31u);
        ax = take ? ax + gx : ax;
        ay = take ? ay + gy : ay;
        __syncwarp();
      } else
      if (tkm != 0u) {
        // warp-uniform trip count; every lane runs the leaf on a copy of its sum and only the takers keep it.
        // Bodies are summed in index order: an odd first body alone, then pairs.  The loop is kept rolled on
        // purpose: leaves hold 4-5 bodies on average, and the dispatch code of an unrolled loop with remainder
        // loops cost more issue slots per visit than it saved.
        uint32_t jb = l.y;
        float lx = ax, ly = ay; // continue the lane's own running sum: same summation order as the scalar walk
#pragma unroll 1
        for (uint32_t o = k & one; o != 0u; --o) { // at most one trip; see `one`
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
        ax = take ? lx : ax; // selects, not a branch: the loop body of the walk stays convergent
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
  if (FUSE) {
    if (valid) integrate_push_body(*fuse, t, make_float2(ax, ay), false); // multi-GPU: integrate.cuh
  } else if (valid) sacc[t] = make_float2(ax, ay);
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

// theta^2 is read from device memory (Meta::theta2, written by the host) rather than passed as a kernel parameter:
// ptxas re-reads a parameter from the constant bank — or re-multiplies theta * theta — on every node visit, one
// issue slot each time in a kernel bound by issue slots; a value loaded from global memory stays in its register.
template <bool COUNT, bool TRANSPOSE = false>
__global__ void __launch_bounds__(256) traverse_fast_kernel(const float4 *__restrict__ box,
                                                            const float4 *__restrict__ node, Meta *meta,
                                                            const float4 *__restrict__ sbody,
                                                            float2 *__restrict__ sacc, uint32_t tfirst,
                                                            uint32_t tcount) {
  if (meta->error) return;
  const float theta2 = meta->theta2;
  // Rectangle::contains is needed unless provably implied (kThetaSqNoContains)
  if (!(theta2 <= kThetaSqNoContains) || meta->cog_outside != 0)
    traverse_fast_body<COUNT, true, TRANSPOSE>(box, node, meta, sbody, sacc, theta2, tfirst, tcount);
  else
    traverse_fast_body<COUNT, false, TRANSPOSE>(box, node, meta, sbody, sacc, theta2, tfirst, tcount);
}

// Multi-GPU step with the peer mappings: the walk, then integrate + NVLink push of the warp's bodies in the same
// kernel (integrate_push_body).  A failed build still has to carry the state over, as integrate_push_kernel does.
template <bool COUNT>
__global__ void __launch_bounds__(256) traverse_fast_push_kernel(const float4 *__restrict__ box,
                                                                 const float4 *__restrict__ node, Meta *meta,
                                                                 const float4 *__restrict__ sbody, uint32_t tfirst,
                                                                 uint32_t tcount, const __grid_constant__ FuseArgs fuse) {
  if (meta->error) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < tcount) integrate_push_body(fuse, tfirst + k, make_float2(0.f, 0.f), true);
    return;
  }
  const float theta2 = meta->theta2;
  if (!(theta2 <= kThetaSqNoContains) || meta->cog_outside != 0)
    traverse_fast_body<COUNT, true, true, true>(box, node, meta, sbody, nullptr, theta2, tfirst, tcount, &fuse);
  else
    traverse_fast_body<COUNT, false, true, true>(box, node, meta, sbody, nullptr, theta2, tfirst, tcount, &fuse);
}

} // namespace nb
