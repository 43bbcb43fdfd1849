// common.cuh — shared device-side definitions for the sm_100a Barnes-Hut step.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

namespace nb {

constexpr int kSMs = 148;                 // B200

// Kernel launches issued by this library on the calling thread (reported in nb_counting).
static thread_local unsigned long long g_launches = 0;
// nb_profile_step: while g_prof points at a list, every host-side launch site drops a named event behind its
// kernel(s); the differences between consecutive events are the kernels' durations in stream order.
struct ProfMark { const char *name; cudaEvent_t ev; };
static thread_local std::vector<ProfMark> *g_prof = nullptr;
inline void prof_mark(cudaStream_t st, const char *name) {
  if (!g_prof) return;
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, st);
  g_prof->push_back({name, e});
}
constexpr int kKeyBits = 48;              // 24 quadtree levels x 2 bits (y bit above x bit)
constexpr int kQuadLevels = 24;
constexpr unsigned long long kKeyOut = 1ull << 48; // body outside the root square (quadtree mode)

// Device-resident per-step metadata: every kernel of a step reads its sizes from here, so a step
// is enqueued without a single host synchronisation.
struct Meta {
  float x0, y0, h;          // root square (Rectangle::new(offset, height), quad_tree.rs:15-21)
  float theta2;             // THETA * THETA (f32 product), written by the host: nb_create / nb_set_params
  uint32_t n;               // bodies
  uint32_t n_in;            // bodies in the tree (sorted positions [0, n_in))
  uint32_t n_nodes;
  uint32_t n_leaves;
  uint32_t max_depth;
  uint32_t overflow_leaves;
  uint32_t error;           // bit 0: node pool too small; bit 1: this build's top-digit sort was too shallow (the step
                            // does nothing and the host redoes it with a full sort); bit 2: peer exchange time-out
  uint32_t cog_outside;     // != 0: some internal node's centre of gravity lies outside its box
  uint32_t resort;          // != 0: the top-digit sort of this step was too shallow for the tree (radix_sort.cuh)
  uint32_t fail_seq;        // sequence number of the first step whose top-digit sort was too shallow (0xffffffff: none)
  uint32_t one;             // 1, written by the host: a loop bound the compiler cannot see through (traverse.cuh)
  unsigned long long interactions; // evaluations of calculate_gravity (main.rs:235-253)
  unsigned long long node_tests;   // opening tests (main.rs:370-372)
  const void *node_base;           // TreeArrays::node again, for the traversal to LOAD rather than take as a parameter (traverse.cuh)
  uint32_t n_movers;        // sort_repair.cuh: bodies whose key left the cell of the sorted depth since the previous build
  uint32_t n_out;           // ... and bodies outside the root square (both counted by morton_repair_kernel)
  uint32_t repair_failed;   // != 0: a repair found more movers than it takes (the host backs off to the radix sort)
};

// Common prefix length of two sorted 48-bit keys, in tree levels of BPL bits each.
template <int BPL> __device__ __forceinline__ int cpl(unsigned long long a, unsigned long long b) {
  unsigned long long d = a ^ b;
  if (d == 0) return kKeyBits / BPL;
  return (__clzll((long long)d) - (64 - kKeyBits)) / BPL;
}
template <int BPL> __device__ __forceinline__ unsigned long long prefix_of(unsigned long long k, int depth) {
  int sh = kKeyBits - depth * BPL; // depth 0 -> 48
  return k >> sh;                  // keys < 2^48, so depth 0 gives 0
}

__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// Pair force, reference src/main.rs:235-253:
//   diff = p2 - p1; sum = |dx|+|dy|; if !sum.is_normal() return;
//   d = dx*dx + dy*dy; d = max(d, 0.001); accel += (diff*m) / (sum*d)
// EXACT reproduces the reference's rounding sequence (no FMA contraction, two IEEE divides).
// FAST: rcp.approx + FFMA; the 1e-20 seed keeps the self pair (sum == 0) finite and is below
// half an ulp of any real sum*d (>= ~4e-6 for f32 coordinates in the 1e5 domain).
template <bool EXACT>
__device__ __forceinline__ void pair_force(float px, float py, float sx, float sy, float m, float &ax,
                                           float &ay) {
  if (EXACT) {
    float dx = __fsub_rn(sx, px), dy = __fsub_rn(sy, py);
    float sum = __fadd_rn(fabsf(dx), fabsf(dy));
    uint32_t b = __float_as_uint(sum);
    if (b - 0x00800000u < 0x7f000000u) { // is_normal (sum >= 0 here)
      float d = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
      if (d < 0.001f) d = 0.001f;
      float den = __fmul_rn(sum, d);
      ax = __fadd_rn(ax, __fdiv_rn(__fmul_rn(dx, m), den));
      ay = __fadd_rn(ay, __fdiv_rn(__fmul_rn(dy, m), den));
    }
  } else {
    float dx = sx - px, dy = sy - py;
    float sum = fabsf(dx) + fabsf(dy);
    float d = fmaf(dy, dy, dx * dx);
    d = fmaxf(d, 0.001f);
    float w = m * rcp_approx(fmaf(sum, d, 1e-20f));
    ax = fmaf(dx, w, ax);
    ay = fmaf(dy, w, ay);
  }
}

// Opening test, reference src/main.rs:370-372 (+ Rectangle::contains, bvh_tree.rs:15-20 /
// quad_tree.rs:23-28).  box = (x0, y0, x1, y1) holds the four f32 values `contains` compares
// against; the products are evaluated left to right, unfused, so the decision is bit-identical.
__device__ __forceinline__ bool accept_node(float px, float py, const float4 &box, float cx, float cy,
                                            float s2, float theta) {
  bool contains = py > box.y && px > box.x && px < box.z && py < box.w;
  float ddx = __fsub_rn(px, cx), ddy = __fsub_rn(py, cy);
  float d2 = __fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy));
  return !contains && s2 < __fmul_rn(__fmul_rn(d2, theta), theta);
}

} // namespace nb
