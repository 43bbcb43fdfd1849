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
#pragma once
#include "common.cuh"
#include "tree_build.cuh"

namespace nb {

template <bool EXACT>
__global__ void __launch_bounds__(256) traverse_kernel(const float4 *__restrict__ box, const float4 *__restrict__ cm,
                                                       const uint4 *__restrict__ link, Meta *meta,
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
  uint32_t cur = 0;
  while (cur < n_nodes) {
    const uint4 l = link[cur];
    const bool active = cur >= resume;
    if (l.z & kLeafBit) {
      const uint32_t s = l.y, k = l.z & ~kLeafBit;
      for (uint32_t j = 0; j < k; ++j) {
        const float4 b = sbody[s + j];
        if (active) pair_force<EXACT>(px, py, b.x, b.y, b.z, ax, ay);
      }
      if (active) n_inter += k;
      cur = cur + 1;
    } else {
      const float4 bx = box[cur];
      const float4 c = cm[cur];
      bool open = false;
      if (active) {
        ++n_test;
        if (accept_node(px, py, bx, c.x, c.y, c.w, theta)) {
          pair_force<EXACT>(px, py, c.x, c.y, c.z, ax, ay);
          resume = l.x;
          ++n_inter;
        } else {
          open = true;
        }
      }
      cur = __any_sync(0xffffffffu, open) ? cur + 1 : l.x;
    }
  }
  if (valid) sacc[t] = make_float2(ax, ay);
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

} // namespace nb
