// integrate.cuh — K7: fused kick-drift integrate + physical re-order, and the AoS<->SoA /
// original-order helpers of the C ABI.
#pragma once
#include "common.cuh"
#include "scan.cuh"

namespace nb {

struct BodyArrays {
  float *x, *y, *vx, *vy;
  uint32_t *m, *id; // integer mass; original body index of the slot
};

// A step whose build failed (Meta::error: node pool overflow, peer time-out) must not advance the state with
// stale accelerations and a stale order: the integrate kernels then copy every body to the same slot of the
// other buffer, so the state stays that of the last good step.
__device__ __forceinline__ void copy_body(const BodyArrays &in, const BodyArrays &out, uint32_t j) {
  out.x[j] = in.x[j];
  out.y[j] = in.y[j];
  out.vx[j] = in.vx[j];
  out.vy[j] = in.vy[j];
  out.m[j] = in.m[j];
  out.id[j] = in.id[j];
}

// Reference src/main.rs:419-423:  v += a*dt;  x += v*dt  (separate mul and add: Rust does not
// contract).  Body j of the tree order is read from its old storage slot and written to slot j of
// the other buffer, so the next step's storage is already Morton-ordered (coalesced writes, near-
// coalesced gathers).  Acceleration follows the body (identity semantics; see DESIGN.md on the
// reference's index-mismatch bug).
__global__ void __launch_bounds__(256) integrate_resort_kernel(const uint32_t *__restrict__ order,
                                                               const float2 *__restrict__ sacc, BodyArrays in,
                                                               BodyArrays out, uint32_t jfirst, uint32_t jcount,
                                                               float dt, const Meta *__restrict__ meta) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= jcount) return;
  const uint32_t j = jfirst + k;
  if (meta->error) { // the build of this step failed: the state is carried over unchanged, slot for slot
    copy_body(in, out, j);
    return;
  }
  const uint32_t s = order[j];
  const float2 a = sacc[j];
  const float vx = __fadd_rn(in.vx[s], __fmul_rn(a.x, dt));
  const float vy = __fadd_rn(in.vy[s], __fmul_rn(a.y, dt));
  out.vx[j] = vx;
  out.vy[j] = vy;
  out.x[j] = __fadd_rn(in.x[s], __fmul_rn(vx, dt));
  out.y[j] = __fadd_rn(in.y[s], __fmul_rn(vy, dt));
  out.m[j] = in.m[s];
  out.id[j] = in.id[s];
}

// main.rs:398-423 AS SHIPPED (cfg.faithful_index_bug): `acceleration[idx]` was computed for the body at index idx
// of the Vec BEFORE BVHTree::from permuted it, and is applied to whatever body sits at idx AFTER.  With
// pos_of_slot the inverse of `order`: the body at tree position j gets the acceleration of the body whose
// pre-build slot was j.
__global__ void __launch_bounds__(256) invert_order_kernel(const uint32_t *__restrict__ order, uint32_t n,
                                                           uint32_t *__restrict__ pos_of_slot) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) pos_of_slot[order[j]] = j;
}
__global__ void __launch_bounds__(256) integrate_resort_bug_kernel(const uint32_t *__restrict__ order,
                                                                   const uint32_t *__restrict__ pos_of_slot,
                                                                   const float2 *__restrict__ sacc, BodyArrays in,
                                                                   BodyArrays out, uint32_t n, float dt,
                                                                   const Meta *__restrict__ meta) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  if (meta->error) {
    copy_body(in, out, j);
    return;
  }
  const uint32_t s = order[j];
  const float2 a = sacc[pos_of_slot[j]];
  const float vx = __fadd_rn(in.vx[s], __fmul_rn(a.x, dt));
  const float vy = __fadd_rn(in.vy[s], __fmul_rn(a.y, dt));
  out.vx[j] = vx;
  out.vy[j] = vy;
  out.x[j] = __fadd_rn(in.x[s], __fmul_rn(vx, dt));
  out.y[j] = __fadd_rn(in.y[s], __fmul_rn(vy, dt));
  out.m[j] = in.m[s];
  out.id[j] = in.id[s];
}

// Permute only mass/id for tree positions outside [jfirst, jfirst+jcount) (multi-GPU: the
// other ranks' bodies arrive by all-gather, mass and id are replicated and permuted locally).
__global__ void __launch_bounds__(256) permute_mid_kernel(const uint32_t *__restrict__ order, BodyArrays in,
                                                          BodyArrays out, uint32_t n, uint32_t jfirst,
                                                          uint32_t jcount, const Meta *__restrict__ meta) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n || (j >= jfirst && j < jfirst + jcount)) return;
  const uint32_t s = meta->error ? j : order[j];
  out.m[j] = in.m[s];
  out.id[j] = in.id[s];
}

// ---- multi-GPU peer exchange (NVLink stores instead of ncclAllGather) -------------------------------------
// integrate_push_kernel is integrate_resort_kernel whose position / velocity stores go to EVERY rank's copy of
// the output arrays (CUDA IPC mappings of the peers' buffers; index r == own rank is the local array), so the
// slice travels over NVLink while the kernel is still integrating.  Ranks then hand-shake through epoch flags:
// peer_signal_kernel runs after the push in stream order (its stores are complete at the kernel boundary),
// fences at system scope and writes this rank's epoch into every peer's flag word; peer_wait_kernel spins until
// every rank's flag reached the epoch.  A peer may start overwriting a buffer only after that hand-shake, i.e.
// after every rank finished reading it.
constexpr int kMaxPeers = 8;
struct PeerOut { float *x[kMaxPeers], *y[kMaxPeers], *vx[kMaxPeers], *vy[kMaxPeers]; };
struct PeerFlags { uint32_t *flags[kMaxPeers]; }; // flags[r] = rank r's flag array, indexed by the writer's rank

// Round 2: only POSITIONS travel.  A velocity is read once per step, by the rank that integrates the body, so the
// velocity of slot s stays on the rank that owns position s (owner = s / per, nb_shard_range) and the integrating
// rank fetches it from there: a local read unless the body crossed a shard boundary in the sort (peer load over
// NVLink through the same IPC mappings).  Halves the bytes pushed (8 B instead of 16 B per body and peer).
// Readers of the whole state call pull_velocity_kernel first (ensure_velocities in nbody_b200.cu).
__global__ void __launch_bounds__(256) integrate_push_kernel(const uint32_t *__restrict__ order,
                                                             const float2 *__restrict__ sacc, BodyArrays in,
                                                             BodyArrays out, PeerOut pin, PeerOut po, int n_ranks,
                                                             uint32_t per, uint32_t jfirst, uint32_t jcount, float dt,
                                                             const Meta *__restrict__ meta) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= jcount) return;
  const uint32_t j = jfirst + k;
  const bool bad = meta->error != 0u; // failed build (the same on every rank): carry the state over unchanged
  const uint32_t s = bad ? j : order[j];
  const float2 a = bad ? make_float2(0.f, 0.f) : sacc[j];
  const int o = min((int)(s / per), n_ranks - 1); // the rank that holds slot s's velocity
  const float vx0 = pin.vx[o][s], vy0 = pin.vy[o][s];
  const float vx = bad ? vx0 : __fadd_rn(vx0, __fmul_rn(a.x, dt));
  const float vy = bad ? vy0 : __fadd_rn(vy0, __fmul_rn(a.y, dt));
  const float x = bad ? in.x[s] : __fadd_rn(in.x[s], __fmul_rn(vx, dt));
  const float y = bad ? in.y[s] : __fadd_rn(in.y[s], __fmul_rn(vy, dt));
  for (int r = 0; r < n_ranks; ++r) {
    po.x[r][j] = x;
    po.y[r][j] = y;
  }
  out.vx[j] = vx;
  out.vy[j] = vy;
  out.m[j] = in.m[s];
  out.id[j] = in.id[s];
}

// The same integrate + push as the tail of the traversal kernel (traverse.cuh, FUSE): a warp integrates its 32 bodies
// the moment their walk ends and stores the new positions into every rank's arrays, so the slice crosses NVLink while
// the other warps are still walking the tree — the exchange hides behind the traversal instead of following it.
// Same arithmetic, same stores as integrate_push_kernel; `bad` = failed build (carry the state over unchanged).
struct FuseArgs {
  const uint32_t *order;
  BodyArrays in, out;
  PeerOut pin, po;
  int n_ranks;
  uint32_t per;
  float dt;
};
__device__ __forceinline__ void integrate_push_body(const FuseArgs &f, uint32_t j, float2 a, bool bad) {
  const uint32_t s = bad ? j : f.order[j];
  const int o = min((int)(s / f.per), f.n_ranks - 1);
  const float vx0 = f.pin.vx[o][s], vy0 = f.pin.vy[o][s];
  const float vx = bad ? vx0 : __fadd_rn(vx0, __fmul_rn(a.x, f.dt));
  const float vy = bad ? vy0 : __fadd_rn(vy0, __fmul_rn(a.y, f.dt));
  const float x = bad ? f.in.x[s] : __fadd_rn(f.in.x[s], __fmul_rn(vx, f.dt));
  const float y = bad ? f.in.y[s] : __fadd_rn(f.in.y[s], __fmul_rn(vy, f.dt));
  for (int r = 0; r < f.n_ranks; ++r) {
    f.po.x[r][j] = x;
    f.po.y[r][j] = y;
  }
  f.out.vx[j] = vx;
  f.out.vy[j] = vy;
  f.out.m[j] = f.in.m[s];
  f.out.id[j] = f.in.id[s];
}

// Make the velocities of ALL slots valid on this rank again (after steps whose exchange pushed positions only).
__global__ void __launch_bounds__(256) pull_velocity_kernel(BodyArrays b, PeerOut pcur, int n_ranks, int rank,
                                                            uint32_t per, uint32_t n) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int o = min((int)(s / per), n_ranks - 1);
  if (o == rank) return;
  b.vx[s] = pcur.vx[o][s];
  b.vy[s] = pcur.vy[o][s];
}

// nb_world_update_shard: rank r holds bodies [first, first + count) of the caller's Vec.  One kernel splits the AoS
// records and stores the slice into EVERY rank's state arrays (slot == original index); the u32 weights travel as
// raw bits through the VX array of the other ping-pong buffer and are moved into m[] by unpack_shard_kernel after the
// hand-shake.  vx, not x: the step's integrate kernel pushes positions into the peers' other buffer, and nothing orders a
// slow rank's unpack before a fast rank's push; velocities are never pushed (integrate_push_kernel), so vx is only
// written by its owner, after its own unpack in stream order.
__global__ void __launch_bounds__(256) push_shard_kernel(const float *__restrict__ aos, PeerOut cur, PeerOut other,
                                                         int n_ranks, uint32_t first, uint32_t count) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= count) return;
  const uint32_t i = first + k;
  const float2 p = *reinterpret_cast<const float2 *>(aos + 6 * (size_t)i);
  const float2 v = *reinterpret_cast<const float2 *>(aos + 6 * (size_t)i + 2);
  const float w = aos[6 * (size_t)i + 4];
  for (int r = 0; r < n_ranks; ++r) {
    cur.x[r][i] = p.x;
    cur.y[r][i] = p.y;
    cur.vx[r][i] = v.x;
    cur.vy[r][i] = v.y;
    other.vx[r][i] = w;
  }
}
__global__ void __launch_bounds__(256) unpack_shard_kernel(BodyArrays b, const float *__restrict__ wbits, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  b.m[i] = __float_as_uint(wbits[i]);
  b.id[i] = i;
}
// Slot order -> ORIGINAL order AoS, only the records whose original index lies in [first, first + count)
// (velocities from the rank that holds them, see integrate_push_kernel)
__global__ void __launch_bounds__(256) soa_to_aos_range_kernel(BodyArrays b, uint32_t n, float *__restrict__ aos,
                                                               uint32_t first, uint32_t count, PeerOut pcur,
                                                               int n_ranks, uint32_t per) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const uint32_t id = b.id[s];
  if (id - first >= count) return;
  const size_t o = 6 * (size_t)id;
  const int ow = min((int)(s / per), n_ranks - 1);
  *reinterpret_cast<float2 *>(aos + o) = make_float2(b.x[s], b.y[s]);
  *reinterpret_cast<float2 *>(aos + o + 2) = make_float2(pcur.vx[ow][s], pcur.vy[ow][s]);
  *reinterpret_cast<uint2 *>(aos + o + 4) = make_uint2(b.m[s], 0u);
}

__global__ void peer_signal_kernel(PeerFlags pf, int n_ranks, int rank, uint32_t epoch) {
  if ((int)threadIdx.x < n_ranks) {
    __threadfence_system();
    *reinterpret_cast<volatile uint32_t *>(pf.flags[threadIdx.x] + rank) = epoch;
  }
}

__global__ void peer_wait_kernel(const uint32_t *my_flags, int n_ranks, uint32_t epoch, Meta *meta) {
  if ((int)threadIdx.x < n_ranks) {
    const volatile uint32_t *f = my_flags + threadIdx.x;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while ((int32_t)(*f - epoch) < 0) {
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 10000000000ull) { // 10 s: a peer is gone; do not hang the device
        meta->error |= 4u;
        break;
      }
    }
    __threadfence_system();
  }
}

// id_from_pad: the record's pad word holds the body's original index (slot-order snapshots of NB_TREE_BVH_REF)
__global__ void __launch_bounds__(256) aos_to_soa_kernel(const float *__restrict__ aos /*6 words per body*/,
                                                         BodyArrays b, uint32_t n, bool id_from_pad = false) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float2 p = *reinterpret_cast<const float2 *>(aos + 6 * (size_t)i);
  const float2 v = *reinterpret_cast<const float2 *>(aos + 6 * (size_t)i + 2);
  const uint2 w = *reinterpret_cast<const uint2 *>(aos + 6 * (size_t)i + 4);
  b.x[i] = p.x; b.y[i] = p.y; b.vx[i] = v.x; b.vy[i] = v.y; b.m[i] = w.x; b.id[i] = id_from_pad ? w.y : i;
}

// After an upload the state sits in CALLER order, and everything behind the sort would gather it through `order`
// four bytes at a time — a 32-byte sector per word: at 64M bodies tree_finish_kernel and integrate_resort_kernel
// moved 14x their algorithmic bytes that way (profiles/r01e_64m_kernels.md).  While the uploaded AoS copy is still
// in HBM, ONE pass gathers whole 24-byte records in tree order into the other SoA buffer and makes `order` the
// identity, so every later kernel of the step reads coalesced.
__global__ void __launch_bounds__(256) permute_from_aos_kernel(const float *__restrict__ aos, uint32_t *__restrict__ order,
                                                               BodyArrays out, uint32_t n) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t s = order[j];
  const float2 p = *reinterpret_cast<const float2 *>(aos + 6 * (size_t)s);
  const float2 v = *reinterpret_cast<const float2 *>(aos + 6 * (size_t)s + 2);
  const uint32_t w = __float_as_uint(aos[6 * (size_t)s + 4]);
  out.x[j] = p.x; out.y[j] = p.y; out.vx[j] = v.x; out.vy[j] = v.y; out.m[j] = w; out.id[j] = s;
  order[j] = j;
}

// Slot order -> ORIGINAL body order AoS.
__global__ void __launch_bounds__(256) soa_to_aos_kernel(BodyArrays b, uint32_t n, float *__restrict__ aos) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const size_t o = 6 * (size_t)b.id[s];
  *reinterpret_cast<float2 *>(aos + o) = make_float2(b.x[s], b.y[s]);
  *reinterpret_cast<float2 *>(aos + o + 2) = make_float2(b.vx[s], b.vy[s]);
  *reinterpret_cast<uint2 *>(aos + o + 4) = make_uint2(b.m[s], 0u);
}

// Device-side scene generation for the two big-N bench scenes: the same counter-based draws
// as csrc/scene.cpp (splitmix64 of (seed, stream, index)), one body per thread, straight into the SoA state —
// no host buffer, no upload.  fp64 sin/cos/sqrt as on the host; after rounding to f32 the two agree except
// where a libm result sits within ~1e-16 of an f32 rounding boundary.
__device__ __forceinline__ unsigned long long scene_mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ float scene_u01(unsigned long long seed, uint32_t stream, unsigned long long idx) {
  const unsigned long long r =
      scene_mix64(scene_mix64(seed ^ (0xD1B54A32D192ED03ull * (stream + 1))) + idx * 0x9E3779B97F4A7C15ull);
  return (float)(r >> 40) * (1.0f / 16777216.0f);
}
// scene 1: uniform disc (r = 45000 sqrt(u) about (50000, 50000)); scene 2: projected Plummer a = 2000, clipped
__global__ void __launch_bounds__(256) scene_kernel(int scene, unsigned long long seed, BodyArrays b, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t stream = scene == 1 ? 3u : 4u;
  const double a = 6.283185307179586 * (double)scene_u01(seed, stream, 2ull * i);
  const double u = (double)scene_u01(seed, stream, 2ull * i + 1);
  double r;
  if (scene == 1) r = 45000.0 * sqrt(u);
  else {
    const double v = 0.999 * u;
    r = 2000.0 * sqrt(v / (1.0 - v));
  }
  float x = (float)(50000.0 + r * cos(a)), y = (float)(50000.0 + r * sin(a));
  if (scene != 1) {
    if (!(x >= 0.f && x < 100000.f)) x = x < 0.f ? 0.f : 99999.99f;
    if (!(y >= 0.f && y < 100000.f)) y = y < 0.f ? 0.f : 99999.99f;
  }
  b.x[i] = x; b.y[i] = y; b.vx[i] = 0.f; b.vy[i] = 0.f; b.m[i] = 1u; b.id[i] = i;
}

// ---- the two scenes whose host generators draw sequentially (csrc/scene.cpp): default scene (World::new,
// main.rs:276-346: the grid-disc acceptance is independent per grid point) and the two-galaxy scene (rejection
// sampling: body i is the i-th ACCEPTED candidate of a counter-based candidate stream).  Both become
// flag -> scan -> compact over the candidate stream, which reproduces the sequential push order.
// kind 0: default-scene grid (candidate k = (x - lo) * ny + (y - lo), x outer / y inner as main.rs:317-318)
// kind 1 / 2: galaxy 0 / 1 of the two-galaxy scene (candidate k = the host loop's `ctr`)
struct SceneCand { bool ok; float x, y, vx, vy; };
__device__ __forceinline__ SceneCand scene_candidate(int kind, unsigned long long seed, unsigned long long k,
                                                     uint32_t lo, uint32_t ny) {
  SceneCand r{false, 0.f, 0.f, 0.f, 0.f};
  if (kind == 0) {
    const uint32_t gx = lo + (uint32_t)(k / ny), gy = lo + (uint32_t)(k % ny);
    const float px = __fmul_rn((float)gx, 14.0f), py = __fmul_rn((float)gy, 14.0f);
    const float dx = __fsub_rn(px, 60000.0f), dy = __fsub_rn(py, 60000.0f);
    const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    if (d2 < 15000000.0f && d2 > 500000.0f) {
      const float hi_r = __fadd_rn(__fsub_rn(15000000.0f, d2), 1.0f);
      const float g = __fmul_rn(scene_u01(seed, 0u, (unsigned long long)gx * 8192ull + gy), hi_r);
      if (g > 6000000.0f) {
        const float kk = __fsqrt_rn(__fdiv_rn(__fsqrt_rn(750000.0f), d2));
        r = SceneCand{true, px, py, __fmul_rn(dy, kk), __fmul_rn(-dx, kk)};
      }
    }
  } else {
    const int g = kind - 1;
    const float cx = g ? 60000.0f : 35000.0f, cvx = g ? 0.0f : 200.0f, cvy = g ? 0.0f : 250.0f;
    const float cmf = g ? 750000.0f : 75000000.0f;
    const double d2s = 500000.0 + (9000000.0 - 500000.0) * (double)scene_u01(seed, 5u + g, 3ull * k);
    const double keep = 1.0 - 6000000.0 / (15000000.0 - d2s + 1.0);
    const double t = (double)scene_u01(seed, 5u + g, 3ull * k + 1);
    if (t < keep) {
      const double a = 6.283185307179586 * (double)scene_u01(seed, 5u + g, 3ull * k + 2);
      const double rr = 4.0 * sqrt(d2s);
      const float px = (float)((double)cx + rr * cos(a)), py = (float)((double)cx + rr * sin(a));
      const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cx);
      const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
      const float kk = __fsqrt_rn(__fdiv_rn(__fsqrt_rn(cmf), d2));
      r = SceneCand{true, px, py, __fadd_rn(__fmul_rn(dy, kk), cvx), __fadd_rn(__fmul_rn(-dx, kk), cvy)};
    }
  }
  return r;
}
// pass 1 over candidates [k0, k0 + cnt): accept flags -> exclusive prefix inside each 256-block + block totals
__global__ void __launch_bounds__(256) scene_flag_kernel(int kind, unsigned long long seed, unsigned long long k0,
                                                         uint32_t cnt, uint32_t lo, uint32_t ny,
                                                         uint32_t *__restrict__ pre, uint32_t *__restrict__ btotal) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t f = (i < cnt && scene_candidate(kind, seed, k0 + i, lo, ny).ok) ? 1u : 0u;
  uint32_t tot;
  const uint32_t ex = block_exclusive_scan<256>(f, tot);
  if (i < cnt) pre[i] = ex | (f << 31);
  if (threadIdx.x == 0) btotal[blockIdx.x] = tot;
}
// pass 2 (after scan_bsum_kernel turned btotal into block bases): the accepted candidates, in candidate order, become
// bodies slot0 + emitted + rank while rank stays below `want`; *emitted advances by the chunk's accepted count
__global__ void __launch_bounds__(256) scene_emit_kernel(int kind, unsigned long long seed, unsigned long long k0,
                                                         uint32_t cnt, uint32_t lo, uint32_t ny,
                                                         const uint32_t *__restrict__ pre,
                                                         const uint32_t *__restrict__ bbase, BodyArrays b, uint32_t slot0,
                                                         uint32_t want, const uint32_t *__restrict__ emitted) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt) return;
  const uint32_t pv = pre[i];
  if (!(pv >> 31)) return;
  const uint32_t rank = *emitted + bbase[blockIdx.x] + (pv & 0x7fffffffu);
  if (rank >= want) return;
  const SceneCand c = scene_candidate(kind, seed, k0 + i, lo, ny);
  const uint32_t s = slot0 + rank;
  b.x[s] = c.x; b.y[s] = c.y; b.vx[s] = c.vx; b.vy[s] = c.vy; b.m[s] = 1u; b.id[s] = s;
}
__global__ void scene_advance_kernel(uint32_t *emitted, const uint32_t *chunk_total) { *emitted += *chunk_total; }
// the bodies that need no compaction: the heavy cores, and the default scene's 100 000 rand_body draws (main.rs:334-335)
__global__ void __launch_bounds__(256) scene_fixed_kernel(int scene, unsigned long long seed, BodyArrays b,
                                                          uint32_t slot0, uint32_t cnt) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i == 0) { // main.rs:282-291 (both scenes share the two cores)
    b.x[0] = 35000.0f; b.y[0] = 35000.0f; b.vx[0] = 200.0f; b.vy[0] = 250.0f; b.m[0] = 75000000u; b.id[0] = 0u;
    b.x[1] = 60000.0f; b.y[1] = 60000.0f; b.vx[1] = 0.0f; b.vy[1] = 0.0f; b.m[1] = 750000u; b.id[1] = 1u;
  }
  if (scene != 0 || i >= cnt) return;
  // rand_disc (main.rs:255-258) in f32 on the host (cosf / sinf); here fp64 cos / sin rounded to f32
  const float TAU = 6.28318530717958647692f;
  float pv[4];
#pragma unroll
  for (int st = 0; st < 2; ++st) {
    const float theta = __fmul_rn(scene_u01(seed, 1u + st, 2ull * i), TAU);
    const float r = scene_u01(seed, 1u + st, 2ull * i + 1);
    pv[2 * st] = __fmul_rn((float)cos((double)theta), r);
    pv[2 * st + 1] = __fmul_rn((float)sin((double)theta), r);
  }
  const uint32_t s = slot0 + i;
  b.x[s] = __fadd_rn(__fmul_rn(pv[0], 25000.0f), 50000.0f);
  b.y[s] = __fadd_rn(__fmul_rn(pv[1], 25000.0f), 50000.0f);
  b.vx[s] = pv[2]; b.vy[s] = pv[3]; b.m[s] = 1u; b.id[s] = s;
}

// Slot order -> AoS in the SAME order: what the reference's Vec looks like after BVHTree::from permuted it.
__global__ void __launch_bounds__(256) soa_to_aos_slot_kernel(BodyArrays b, uint32_t n, float *__restrict__ aos,
                                                              bool id_in_pad = false) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const size_t o = 6 * (size_t)s;
  *reinterpret_cast<float2 *>(aos + o) = make_float2(b.x[s], b.y[s]);
  *reinterpret_cast<float2 *>(aos + o + 2) = make_float2(b.vx[s], b.vy[s]);
  *reinterpret_cast<uint2 *>(aos + o + 4) = make_uint2(b.m[s], id_in_pad ? b.id[s] : 0u);
}

__global__ void __launch_bounds__(256) set_identity_kernel(uint32_t *id, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) id[i] = i;
}

// Slot order -> original order, one u32/f32 word per body.
__global__ void __launch_bounds__(256) unpermute_word_kernel(const uint32_t *__restrict__ src,
                                                             const uint32_t *__restrict__ id, uint32_t n,
                                                             uint32_t *__restrict__ dst) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) dst[id[s]] = src[s];
}
__global__ void __launch_bounds__(256) unpermute_key_kernel(const unsigned long long *__restrict__ src,
                                                            const uint32_t *__restrict__ id, uint32_t n,
                                                            unsigned long long *__restrict__ dst) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) dst[id[s]] = src[s];
}

// Tree-order accelerations -> ORIGINAL body order (the `acceleration` Vec of main.rs:406-416).
__global__ void __launch_bounds__(256) accel_to_original_kernel(const float2 *__restrict__ sacc,
                                                                const uint32_t *__restrict__ order,
                                                                const uint32_t *__restrict__ id, uint32_t n,
                                                                float *__restrict__ ax, float *__restrict__ ay) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t o = id[order[j]];
  const float2 a = sacc[j];
  ax[o] = a.x;
  ay[o] = a.y;
}

// tree position -> original id
__global__ void __launch_bounds__(256) order_to_original_kernel(const uint32_t *__restrict__ order,
                                                                const uint32_t *__restrict__ id, uint32_t n,
                                                                uint32_t *__restrict__ out) {
  const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < n) out[j] = id[order[j]];
}

// original id -> slot
__global__ void __launch_bounds__(256) slot_of_id_kernel(const uint32_t *__restrict__ id, uint32_t n,
                                                         uint32_t *__restrict__ slot_of) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) slot_of[id[s]] = s;
}

// `draw` (reference src/main.rs:41-72) as two passes over a headless RGBA8 frame:
//   pass 1 per body: heavy bodies (weight > 10) mark their pixel; light bodies count hits and
//          record the LAST body (largest original index) hitting the pixel — the reference's
//          sequential loop leaves the colour of the last light body drawn before alpha hit 0xff.
//   pass 2 per pixel: compose exactly what the sequential loop would have left.
// The sequential semantics depend on array order only through (a) the last light writer while
// alpha != 0xff and (b) heavy bodies overriding; see nbody_b200.cu for the composition rule.
struct DrawScratch {
  uint32_t *hits;   // light-body hits per pixel
  uint32_t *last;   // max original id + 1 of a light body on the pixel
  uint32_t *heavy;  // max original id + 1 of a heavy body on the pixel
};

// by_slot (NB_TREE_BVH_REF): the reference draws its Vec in the order BVHTree::from left it (main.rs:46, :400), i.e.
// slot order, so "last" is the largest SLOT; otherwise the Vec order is the original body order.
__global__ void __launch_bounds__(256) draw_splat_kernel(BodyArrays b, uint32_t n, uint32_t render_h, float world_h,
                                                         uint32_t cell, DrawScratch d, bool by_slot) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const float px = b.x[s], py = b.y[s];
  if (!(py < world_h && px < world_h && py >= 0.f && px >= 0.f)) return; // within_bounds, main.rs:223-226
  const uint32_t pix = ((uint32_t)py / cell) * render_h + ((uint32_t)px / cell);
  const uint32_t oid = (by_slot ? s : b.id[s]) + 1u;
  if (b.m[s] > 10u) atomicMax(&d.heavy[pix], oid);
  else {
    atomicAdd(&d.hits[pix], 1u);
    atomicMax(&d.last[pix], oid);
  }
}

} // namespace nb
