// nbody_b200.cu — context + C ABI (include/nbody_b200.h) of the B200-native Barnes-Hut step.
//
// Host-side mirror of the reference's `World` (src/main.rs:37-39) and `World::update`
// (src/main.rs:388-425): build tree -> up-sweep -> traverse -> integrate, three timed phases
// accumulated into a Counting-like struct.  All state lives in HBM; a step is enqueued on one
// stream with no host synchronisation inside it.  There is no CPU fallback.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h> // types only; libnccl is dlopen'ed in nb_comm_init

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/nbody_b200.h"
#include "common.cuh"
#include "direct.cuh"
#include "integrate.cuh"
#include "morton.cuh"
#include "radix_sort.cuh"
#include "sort_repair.cuh"
#include "scan.cuh"
#include "traverse.cuh"
#include "traverse_group.cuh"
#include "ubench.cuh"
#include "tree_build.cuh"
#include "bvh_ref.cuh"

using namespace nb;

static_assert(sizeof(nb_particle) == 24, "nb_particle mirrors the 24-byte Particle (src/main.rs:193-198)");
static_assert(sizeof(nb_config) == 88 && sizeof(nb_counting) == 96 && sizeof(nb_tree_info) == 72,
              "C ABI struct layout (tests/test_abi.py and the ctypes binding rely on it)");

namespace {

thread_local std::string g_create_error;

struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool load(std::string &err) {
    if (lib) return true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *nm : names) {
      lib = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (lib) break;
    }
    if (!lib) {
      err = std::string("cannot dlopen libnccl: ") + dlerror();
      return false;
    }
#define NB_SYM(field, name)                                      \
  *(void **)(&field) = dlsym(lib, name);                         \
  if (!field) {                                                  \
    err = std::string("libnccl lacks symbol ") + name;           \
    return false;                                                \
  }
    NB_SYM(GetUniqueId, "ncclGetUniqueId");
    NB_SYM(CommInitRank, "ncclCommInitRank");
    NB_SYM(CommDestroy, "ncclCommDestroy");
    NB_SYM(AllGather, "ncclAllGather");
    NB_SYM(GroupStart, "ncclGroupStart");
    NB_SYM(GroupEnd, "ncclGroupEnd");
    NB_SYM(GetErrorString, "ncclGetErrorString");
#undef NB_SYM
    return true;
  }
};
NcclApi g_nccl;

} // namespace

struct nb_ctx {
  nb_config cfg{};
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t side = nullptr;           // multi-GPU: the mass / id re-order of the other ranks' slots runs beside the traversal
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  uint32_t n = 0;       // bodies set
  uint32_t n_alloc = 0; // array capacity (padded)
  int leaf_cap = 8;
  bool quad = true;
  bool bvh_ref = false;   // NB_TREE_BVH_REF: level-synchronous BVHTree::from (bvh_ref.cuh)
  BvhWork bw{};
  TileArrays tiles{};
  BvhMeta *h_bmeta = nullptr; // pinned mirror of bw.meta
  uint32_t bvh_levels = 0;    // levels of the last build
  bool bmeta_in_flight = false;  // a copy of BvhMeta into h_bmeta is enqueued: read it after the next synchronisation
  bool bvh_levels_known = false; // ... from a completed build of this context: the next build is enqueued blind

  BodyArrays body[2]{};
  int cur = 0; // body[cur] holds the current state (slot order)
  unsigned long long *keys[2]{};
  unsigned long long *keys_new = nullptr; // sort_repair.cuh: this build's keys in slot order, beside the old sorted keys
  uint32_t *vals[2]{};
  int sorted_in = 0; // keys[sorted_in] / vals[sorted_in] hold the sorted keys / tree order
  uint8_t *leafdepth = nullptr;
  uint32_t *nstart = nullptr; // per-body exclusive node-count prefix inside its 256-body block (+ sort.bsum block bases = pre-order index)
  SortScratch sort{};
  TreeArrays T{};
  float4 *sbody = nullptr; // bodies in tree order: (x, y, mass f32, mass bits)
  float2 *sacc = nullptr;  // accelerations in tree order
  float4 *bounds_partial = nullptr;
  Meta *meta = nullptr; // device
  Meta *h_meta = nullptr; // pinned host mirror
  float *aos = nullptr;  // 6 words per body staging
  void *flush_buf = nullptr;
  size_t flush_bytes = 0;
  bool tree_valid = false;
  bool depth_known = false;    // last_max_depth comes from a completed quadtree build of the current body set
  uint32_t last_max_depth = 0; // deepest leaf of the last checked build: sizes the next step's top-digit sort
  uint32_t resorts = 0;        // builds (as seen at the checks) whose top-digit sort had to be redone in full
  int last_top_passes = 0;     // top digits the last build sorted before checking (0 = full sort)
  int last_top_levels = 0;     // ... and the quadtree levels they cover
  bool counting = true; // interaction / opening-test counters in the traversal
  unsigned long long seen_interactions = 0, seen_tests = 0;
  uint32_t step_seq = 0;       // sequence number of the build being enqueued (tree_count_kernel -> Meta::fail_seq)
  bool resort_hit = false;     // the last synchronisation found a build whose top-digit sort was too shallow
  uint32_t resort_at = 0;      // ... and this is its sequence number
  bool redo_next = false, last_redone = false; // the next / the last build is the full-sort redo of a failed guess
  int redo_guess = 0, last_failed_guess = 0;   // the guess that failed
  int redo_guess_levels = 0, last_failed_levels = 0;

  std::vector<cudaEvent_t> ev; // 5 per timed step, drained in batches
  size_t ev_used = 0;

  // multi-GPU
  int rank = 0, n_ranks = 1;
  ncclComm_t comm = nullptr;
  // peer-store exchange (nb_comm_peer_import): IPC mappings of every rank's x/y/vx/vy (both ping-pong parities)
  bool peer = false;
  PeerOut peer_out[2]{};          // by output parity
  PeerFlags peer_flags_of{};      // every rank's flag array
  uint32_t *peer_flags = nullptr; // this rank's flag array: [r] = epoch of rank r's last completed push
  uint32_t peer_epoch = 0;
  bool aos_fresh = false;         // c->aos holds the uploaded records of the CURRENT state (caller order == slot order)
  uint32_t aos_permute_min = 1u << 21; // bodies from which the first build after an upload re-orders through the AoS copy
  // sort_repair.cuh: the state is in the order of the tree it was last integrated with, and keys[sorted_in] holds that
  // tree's sorted keys (sorted on their top keys_sorted_bits bits, keys_sorted_n of them)
  bool order_coherent = false;
  int keys_sorted_bits = 0;
  uint32_t keys_sorted_n = 0;
  bool movers_valid = false;     // movers_seen is the count of the last build (of this state), read at the last synchronisation
  uint32_t movers_seen = 0;
  bool last_counted = false;     // the last build enqueued counted its movers
  uint32_t repair_off = 0;       // builds to stay away from the repair after one failed
  uint32_t batch_steps = 1;      // steps the current do_steps batch enqueues before the host looks again
  bool no_repair = false;        // NB_B200_NO_REPAIR=1
  bool last_repaired = false, redo_guess_repaired = false, last_failed_repaired = false;
  RepairScratch repair{};
  int sort_min_levels = 8;        // NB_B200_SORT_MIN_LEVELS: least depth a top-digit sort covers (tests: reach every digit width)
  bool no_fuse = false;           // NB_B200_NO_FUSE=1: keep integrate + push a kernel of its own (A/B runs)
  bool v_partial = false;         // peer exchange: only the velocities of the owned slots are valid here (integrate.cuh)
  std::vector<void *> peer_mapped; // pointers to close in nb_destroy

  std::string err;
};

namespace {

int fail(nb_ctx *c, int code, const std::string &msg) {
  if (c) c->err = msg;
  else g_create_error = msg;
  return code;
}

#define NB_CUDA(c, call)                                                                           \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return fail(c, NB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));            \
  } while (0)

template <class Tp> cudaError_t dmalloc(Tp **p, size_t count) { return cudaMalloc((void **)p, count * sizeof(Tp)); }

inline uint32_t cdiv(uint32_t a, uint32_t b) { return (a + b - 1) / b; }

// THETA * THETA as one f32 product (rounded once): the bits a device FMUL of the same operands gives.
inline float theta_squared(float theta) {
  const volatile float t = theta;
  const volatile float t2 = t * t;
  return t2;
}

int alloc_bodies(nb_ctx *c, BodyArrays &b, size_t n) {
  NB_CUDA(c, dmalloc(&b.x, n));
  NB_CUDA(c, dmalloc(&b.y, n));
  NB_CUDA(c, dmalloc(&b.vx, n));
  NB_CUDA(c, dmalloc(&b.vy, n));
  NB_CUDA(c, dmalloc(&b.m, n));
  NB_CUDA(c, dmalloc(&b.id, n));
  return NB_OK;
}
void free_bodies(BodyArrays &b) {
  cudaFree(b.x); cudaFree(b.y); cudaFree(b.vx); cudaFree(b.vy); cudaFree(b.m); cudaFree(b.id);
  b = BodyArrays{};
}

// slots per rank of nb_shard_range (also: which rank holds a slot's velocity, integrate.cuh)
uint32_t shard_per(const nb_ctx *c) {
  const uint32_t g = (uint32_t)std::max(1, c->n_ranks);
  return std::max(32u, ((c->n + g - 1) / g + 31u) / 32u * 32u);
}

// Peer exchange pushes positions only: before anything reads the whole state, fetch the other ranks' velocities.
int ensure_velocities(nb_ctx *c) {
  if (!c->v_partial) return NB_OK;
  c->v_partial = false;
  if (c->n == 0 || c->n_ranks < 2 || !c->peer) return NB_OK;
  pull_velocity_kernel<<<(c->n + 255) / 256, 256, 0, c->stream>>>(c->body[c->cur], c->peer_out[c->cur], c->n_ranks, c->rank,
                                                                 shard_per(c), c->n);
  g_launches += 1;
  NB_CUDA(c, cudaGetLastError());
  return NB_OK;
}

// owned range of tree positions for this rank
void owned_range(const nb_ctx *c, uint32_t &first, uint32_t &count) {
  uint64_t f, n;
  nb_shard_range(c->n, c->rank, c->n_ranks, &f, &n);
  first = (uint32_t)f;
  count = (uint32_t)n;
}

// ---- one build: morton -> sort -> topology -> up-sweep (BVHTree::from + calculate_gravity,
// bvh_tree.rs:57-158 / QuadTree::insert + calculate_gravity, quad_tree.rs:153-270) --------------
// NB_TREE_BVH_REF: BVHTree::from (bvh_tree.rs:57-96) level by level, then the shared up-sweep.
int enqueue_build_bvh_ref(nb_ctx *c) {
  const uint32_t n = c->n;
  cudaStream_t st = c->stream;
  BodyArrays &b = c->body[c->cur];
  BvhWork &W = c->bw;
  bvh_init_kernel<<<cdiv(n, 256), 256, 0, st>>>(W, c->meta, b.x, b.y, n);
  g_launches += 1;
  const uint32_t scan_blocks = cdiv(n, (uint32_t)kScanTile);
  // Level-synchronous part: nodes of more than kBvhSmall bodies.  Every kernel reads its work list from device
  // memory, so a level can be enqueued blind; the host only has to know WHEN the big nodes run out.  It enqueues
  // as many levels as the previous build needed without looking, then checks once (one sync per build).
  auto enqueue_level = [&](int level, uint32_t nbig_bound) {
    const uint32_t gx = std::max(1u, std::min(nbig_bound, (uint32_t)kSMs * 8u));
    // the warps of gy blocks share one node's tiles; with many nodes one block per node suffices
    const uint32_t gy = std::max(1u, std::min(128u, cdiv(cdiv(n, (uint32_t)kFoldTile * 8u), gx)));
    bvh_tile_sum_kernel<<<dim3(gx, gy, 2), 256, 0, st>>>(W, c->tiles, level);
    bvh_tile_prefix_kernel<<<gx, 64, 0, st>>>(W, c->tiles, level);
    bvh_tile_summary_kernel<<<dim3(gx, gy, 2), 256, 0, st>>>(W, c->tiles, level);
    bvh_fold_resolve_kernel<<<gx, 64, 0, st>>>(W, c->tiles, level);
    bvh_count_kernel<<<cdiv(n, 256), 256, 0, st>>>(W, n, level);
    bvh_decide_kernel<<<cdiv(nbig_bound, 256), 256, 0, st>>>(W, level, (uint32_t)c->leaf_cap);
    bvh_flag_reduce_kernel<<<scan_blocks, kScanThreads, 0, st>>>(W, n, level);
    scan_bsum_kernel<<<1, 1024, 0, st>>>(W.bsum, scan_blocks, nullptr);
    bvh_flag_apply_kernel<<<scan_blocks, kScanThreads, 0, st>>>(W, n, level);
    bvh_publish_kernel<<<cdiv(n, 256), 256, 0, st>>>(W, n, level);
    bvh_swap_kernel<<<cdiv(n, 256), 256, 0, st>>>(W, n, level);
    g_launches += 11;
  };
  // big nodes at a level: at most 2^level, and each holds more than kBvhSmall bodies
  auto bound = [&](int level) {
    const uint32_t by_size = n / kBvhSmall + 1u;
    return level >= 31 ? by_size : std::min(by_size, 1u << level);
  };
  int level = 0;
  if (n > kBvhSmall && c->bvh_levels_known) {
    // No host synchronisation: as many levels as the previous build needed plus one, then a one-thread kernel
    // checks that no big node is left.  If one is, the rest of the step is a no-op and the host — which sees the flag
    // at its next look at Meta — redoes the step through the synchronising loop below (do_steps, nb_build_tree).
    const int blind = std::min((int)c->bvh_levels + 1, kBvhMaxLevels);
    for (; level < blind; ++level) enqueue_level(level, bound(level));
    bvh_levels_check_kernel<<<1, 32, 0, st>>>(W, c->meta, blind, c->step_seq);
    g_launches += 1;
    // the level count of THIS build reaches the host with the next copy of BvhMeta (sync_and_check)
    NB_CUDA(c, cudaMemcpyAsync(c->h_bmeta, W.meta, sizeof(BvhMeta), cudaMemcpyDeviceToHost, st));
    c->bmeta_in_flight = true;
  } else if (n > kBvhSmall) {
    // first build of a context (or the redo of a failed guess): look after every level
    for (; level < (int)c->bvh_levels && level < kBvhMaxLevels; ++level) enqueue_level(level, bound(level));
    for (;; ++level) {
      NB_CUDA(c, cudaMemcpyAsync(c->h_bmeta, W.meta, sizeof(BvhMeta), cudaMemcpyDeviceToHost, st));
      NB_CUDA(c, cudaStreamSynchronize(st));
      if (c->h_bmeta->n_nodes > W.node_cap) return fail(c, NB_ERR_OVERFLOW, "BVH node pool too small");
      if (level >= kBvhMaxLevels || c->h_bmeta->cnt_big[level] == 0) break;
      enqueue_level(level, c->h_bmeta->cnt_big[level]);
    }
    // levels that really held big nodes (the blind part may have overshot)
    int used = 0;
    while (used < kBvhMaxLevels && c->h_bmeta->cnt_big[used] != 0) ++used;
    c->bvh_levels = (uint32_t)used;
    c->bvh_levels_known = true;
  }
  // everything below: one block per small subtree, in shared memory
  bvh_subtree_kernel<<<kSMs * 2, kSubThreads, sizeof(SubSmem), st>>>(W, (uint32_t)c->leaf_cap);
  bvh_sizes_kernel<<<kSMs * 8, 256, 0, st>>>(W);
  bvh_emit_kernel<<<kSMs * 8, 256, 0, st>>>(W, c->meta, c->T, n);
  tree_finish_kernel<true, 8, true><<<std::max((uint32_t)kSMs * 16u, cdiv(n / 16u, 128u)), 128, 0, st>>>(c->meta, c->T, W.order, b.x, b.y, b.m, b.id, c->sbody);
  g_launches += 4;
  NB_CUDA(c, cudaGetLastError());
  c->sorted_in = 0; // the tree order lives in vals[0]
  c->tree_valid = true;
  ++c->step_seq;
  return NB_OK;
}

inline uint32_t gb_of(uint32_t n) { return (n + 255u) / 256u; }

int enqueue_build(nb_ctx *c) {
  const uint32_t n = c->n;
  cudaStream_t st = c->stream;
  if (n == 0) return NB_OK;
  // a failed build is not sticky: every build starts with a clean error word (the failed step itself left the
  // state untouched, see copy_body in integrate.cuh)
  NB_CUDA(c, cudaMemsetAsync(&c->meta->error, 0, sizeof(uint32_t), st));
  if (c->bvh_ref) {
    c->aos_fresh = false;
    return enqueue_build_bvh_ref(c);
  }
  BodyArrays &b0 = c->body[c->cur]; // the state as it is stored now (caller order right after an upload)
  if (!c->quad && c->cfg.domain_auto) {
    const int np = kSMs * 4;
    g_launches += 2;
    bounds_partial_kernel<<<np, 256, 0, st>>>(b0.x, b0.y, n, c->bounds_partial);
    bounds_final_kernel<<<1, 32, 0, st>>>(c->bounds_partial, np, c->meta);
  }
  g_launches += 5; // morton, count, block-base scan, emit, finish (+ the sort, counted there; + link, counted below)
  // Quadtree: sort only as many top key bits as the previous build's depth (+1 level) needs, in as few passes as
  // digits of up to 9 bits allow; tree_count_kernel verifies, and the host redoes the step with a full sort if the
  // guess was too shallow (radix_sort.cuh).  The top bit of the key is the outside bit, two bits per level below it.
  SortPlan plan = make_sort_plan(c->quad ? 49 : 48);
  int top_passes = 0, sorted_levels = 0;
  if (c->quad && c->depth_known && !c->cfg.full_sort) {
    const int levels = std::max((int)c->last_max_depth + 1, c->sort_min_levels);
    const SortPlan top = make_top_plan(49, 2 * levels + 1);
    if (top.passes < plan.passes) {
      plan = top;
      top_passes = top.passes;
      sorted_levels = (top.sorted_bits - 1) / 2;
    }
  }
  c->last_top_passes = top_passes;
  c->last_top_levels = sorted_levels;
  // The state is still in the order of the tree it was integrated with: repair that order instead of sorting
  // (sort_repair.cuh).  Needs the old keys sorted at least as deep as this build sorts.
  const bool aligned = top_passes > 0 && c->order_coherent && !c->no_repair && c->repair_off == 0 && c->keys_sorted_n == n &&
                       c->keys_sorted_bits >= plan.sorted_bits && n >= 4096u && c->h_meta->n_in == n &&
                       c->leaf_cap < kRepairLeafScan && c->h_meta->overflow_leaves == 0; // leaves it can rank in place
  if (top_passes > 0 && c->order_coherent && c->repair_off) --c->repair_off;
  // ... and only after a build that counted few enough movers; the others count them in front of their radix sort
  // (a batch of steps enqueued blind cannot react to a count that grows: stay further below the limit there)
  const bool repair = aligned && c->movers_valid && c->movers_seen <= (uint32_t)kRepairMax / (c->batch_steps > 8u ? 8u : 2u);
  c->last_counted = aligned;
  c->order_coherent = false;
  c->last_repaired = repair;
  c->last_redone = c->redo_next;
  c->last_failed_guess = c->redo_guess;
  c->last_failed_levels = c->redo_guess_levels;
  c->last_failed_repaired = c->redo_guess_repaired;
  c->redo_next = false;
  int which;
  { // only the levels the top-digit sort covers (the check in tree_count_kernel catches a deeper leaf)
    const uint32_t mg = cdiv(cdiv(n, 4), 256);
    const int q = c->quad ? 1 : 0;
    const int lv = (sorted_levels == 0 || sorted_levels > 20) ? kQuadLevels : sorted_levels <= 8 ? 8 : sorted_levels <= 10 ? 10
                 : sorted_levels <= 12 ? 12 : sorted_levels <= 13 ? 13 : sorted_levels <= 14 ? 14 : sorted_levels <= 16 ? 16 : 20;
    if (aligned) {
      const int p = c->sorted_in, shift = plan.shift[0];
      const unsigned long long *prev = c->keys[p];
      unsigned long long *knew = repair ? c->keys_new : c->keys[0];
      uint32_t *vnew = repair ? nullptr : c->vals[0];
      NB_CUDA(c, cudaMemsetAsync(&c->meta->n_movers, 0, 2 * sizeof(uint32_t), st)); // n_movers, n_out
#define NB_MORTON(L) morton_repair_kernel<L><<<mg, 256, 0, st>>>(b0.x, b0.y, c->meta, n, prev, c->leafdepth, knew, vnew, c->repair)
      switch (lv) {
      case 8: NB_MORTON(8); break;
      case 10: NB_MORTON(10); break;
      case 12: NB_MORTON(12); break;
      case 13: NB_MORTON(13); break;
      case 14: NB_MORTON(14); break;
      case 16: NB_MORTON(16); break;
      case 20: NB_MORTON(20); break;
      default: NB_MORTON(kQuadLevels); break;
      }
#undef NB_MORTON
      prof_mark(st, "morton_repair_kernel");
      if (repair) {
        repair_movers_kernel<<<1, 1024, 0, st>>>(c->meta, n, shift, c->repair, c->step_seq);
        prof_mark(st, "repair_movers_kernel");
        sort_repair_kernel<<<cdiv(n, (uint32_t)kRepairTile), kRepairThreads, 0, st>>>(c->meta, n, prev, c->leafdepth, knew, shift, c->repair, c->keys[p ^ 1],
                                                         c->vals[p ^ 1], c->step_seq);
        prof_mark(st, "sort_repair_kernel");
        g_launches += 2;
        which = p ^ 1;
      } else {
        which = radix_sort_pairs(c->keys[0], c->vals[0], c->keys[1], c->vals[1], n, plan, c->sort, st, c->meta);
      }
    } else {
#define NB_MORTON(L) morton_kernel<L><<<mg, 256, 0, st>>>(b0.x, b0.y, c->meta, n, q, c->keys[0], c->vals[0])
      switch (lv) {
      case 8: NB_MORTON(8); break;
      case 10: NB_MORTON(10); break;
      case 12: NB_MORTON(12); break;
      case 13: NB_MORTON(13); break;
      case 14: NB_MORTON(14); break;
      case 16: NB_MORTON(16); break;
      case 20: NB_MORTON(20); break;
      default: NB_MORTON(kQuadLevels); break;
      }
#undef NB_MORTON
      prof_mark(st, "morton_kernel");
      which = radix_sort_pairs(c->keys[0], c->vals[0], c->keys[1], c->vals[1], n, plan, c->sort, st, c->meta);
    }
  }
  c->keys_sorted_bits = plan.sorted_bits;
  c->keys_sorted_n = n;
  c->sorted_in = which;
  const unsigned long long *skeys = c->keys[which];
  uint32_t *order = c->vals[which];
  // Not with the peer exchange: the re-order flips the ping-pong parity, so this step's integrate kernel would push
  // into the peers' ORIGINAL buffer — which a slower peer may still be reading for its own Morton keys (no hand-shake
  // separates an upload from the first step).  There the first step pays the caller-order gathers instead.
  const bool reorder = c->aos_fresh && n >= c->aos_permute_min && !(c->n_ranks > 1 && c->peer);
  c->aos_fresh = false;
  if (reorder) { // first build after an upload: bring the state into tree order once, through the AoS copy
    permute_from_aos_kernel<<<gb_of(n), 256, 0, st>>>(c->aos, order, c->body[c->cur ^ 1], n);
    c->cur ^= 1;
    g_launches += 1;
    prof_mark(st, "permute_from_aos_kernel");
  }
  BodyArrays &b = c->body[c->cur];
  const uint32_t gb = cdiv(n, 256);
  uint32_t *bbase = c->sort.bsum; // per-256-body block: node count, then (scan_bsum_kernel) the block's first node index
  if (c->quad)
    tree_count_kernel<2, false><<<gb, 256, 0, st>>>(skeys, c->meta, c->leaf_cap, c->leafdepth, c->nstart, bbase, sorted_levels, c->step_seq);
  else tree_count_kernel<1, true><<<gb, 256, 0, st>>>(skeys, c->meta, c->leaf_cap, c->leafdepth, c->nstart, bbase);
  prof_mark(st, "tree_count_kernel");
  { // block bases + meta->n_nodes: chunks of 1024 block totals, then their bases (scan.cuh)
    const uint32_t chunks = cdiv(gb, 1024u);
    uint32_t *ctot = bbase + gb; // bsum holds n_alloc / 256 + n_alloc / 262144 + 64 words: chunk totals behind the block totals
    scan_chunks_kernel<<<chunks, 1024, 0, st>>>(bbase, gb, ctot);
    scan_bases_kernel<<<chunks, 1024, 0, st>>>(bbase, gb, ctot, &c->meta->n_nodes);
    g_launches += 1;
  }
  prof_mark(st, "scan_bsum_kernel");
  // A depth guess that was too shallow (tree_count_kernel raised error bit 1) turns the rest of this step, and of
  // every step enqueued behind it, into no-ops that leave the state untouched; the host sees it at its next
  // synchronisation and redoes those steps with a full sort (do_steps, build_tree_checked).  Round 1 enqueued the
  // full sort + second count behind every build as conditional no-ops: 8 empty launches, ~30 us of a 250 us build.
  // Parent / children / twig links: from the parents' side, either inside the emit kernel (extra searches on the
  // ~30 % of lanes that hold internal nodes) or by tree_link_kernel (chases skip links).  Measured build phase,
  // link kernel vs emit: 1M 0.298 / 0.313 ms, 2M 0.515 / 0.517, 4M 0.813 / 0.807, 8M 1.586 / 1.467 — once the node
  // arrays outgrow L2 the separate pass pays for re-reading them.
  const bool emit_links = n >= 4000000u;
  if (c->quad && emit_links) tree_emit_quad_kernel<true><<<cdiv(n, (uint32_t)kEmitBodies), 256, 0, st>>>(skeys, c->meta, c->leafdepth, c->nstart, bbase, c->T);
  else if (c->quad) tree_emit_quad_kernel<false><<<cdiv(n, (uint32_t)kEmitBodies), 256, 0, st>>>(skeys, c->meta, c->leafdepth, c->nstart, bbase, c->T);
  else tree_emit_kernel<1, true><<<gb, 256, 0, st>>>(skeys, c->meta, c->leafdepth, c->nstart, bbase, c->T);
  prof_mark(st, "tree_emit_kernel");
  const uint32_t link_blocks = std::max((uint32_t)kSMs * 8u, cdiv(n / 2u, 256u));
  if (!c->quad || !emit_links) g_launches += 1;
  if (!c->quad) tree_link_kernel<true><<<link_blocks, 256, 0, st>>>(c->meta, c->T);
  else if (!emit_links) tree_link_kernel<true><<<link_blocks, 256, 0, st>>>(c->meta, c->T);
  if (!c->quad || !emit_links) prof_mark(st, "tree_link_kernel");
  // about one node per thread (the grid-stride loop covers the rest): a thread that completes a parent climbs on
  // while its warp-mates idle, so short-lived warps keep the machine full where a long loop would serialise tails
  const uint32_t fin_blocks = std::max((uint32_t)kSMs * 16u, cdiv(n / 2u, 128u)); // measured: 0.38 -> 0.36 ms build at 1M, 2.32 -> 1.91 at 8M
  if (c->quad) tree_finish_kernel<false, 8><<<fin_blocks, 128, 0, st>>>(c->meta, c->T, order, b.x, b.y, b.m, b.id, c->sbody);
  else // one warp per node: about n / 16 warps (an LBVH of 64-body buckets has ~n / 20 nodes)
    tree_finish_kernel<true, 8, false, true><<<std::max((uint32_t)kSMs * 16u, cdiv(n / 16u, 4u)), 128, 0, st>>>(c->meta, c->T, order, b.x, b.y, b.m, b.id, c->sbody);
  prof_mark(st, "tree_finish_kernel");
  NB_CUDA(c, cudaGetLastError());
  c->tree_valid = true;
  ++c->step_seq;
  return NB_OK;
}

// World::bvh_sum_gravity over the par_iter (main.rs:349-386, 406-416) for tree positions
// [first, first+count).
// True if this step's traversal kernel may integrate and push its bodies itself (traverse_fast_push_kernel): the
// multi-GPU step with the peer mappings, on the default walk of NB_FORCE_FAST.
bool step_can_fuse(const nb_ctx *c) {
  return c->n_ranks > 1 && c->peer && c->cfg.force_mode != NB_FORCE_EXACT && !c->bvh_ref &&
         c->cfg.fast_walk_group == 0 && !c->cfg.faithful_index_bug && !c->no_fuse;
}

FuseArgs make_fuse_args(nb_ctx *c, float dt) {
  FuseArgs f{};
  f.order = c->vals[c->sorted_in];
  f.in = c->body[c->cur];
  f.out = c->body[c->cur ^ 1];
  // Where the velocities are read from: the owners' arrays — except right after an upload / pull, when every
  // rank holds all of them itself and a peer may not even have finished ITS upload yet (no hand-shake between
  // an upload and the first step): then, and only then, the local arrays.
  f.pin = c->peer_out[c->cur];
  if (!c->v_partial)
    for (int r = 0; r < kMaxPeers; ++r) { f.pin.vx[r] = f.in.vx; f.pin.vy[r] = f.in.vy; }
  f.po = c->peer_out[c->cur ^ 1];
  f.n_ranks = c->n_ranks;
  f.per = shard_per(c);
  f.dt = dt;
  return f;
}

int enqueue_traverse(nb_ctx *c, uint32_t first, uint32_t count, const FuseArgs *fuse = nullptr) {
  if (count == 0) return NB_OK;
  const int tpb = 256; // 64 / 128 measured slower (0.84 / 0.83 vs 0.80 ms at 1M); persistent warps with a
                       // global work ticket also measured slower (0.84 ms: consecutive groups lose their shared L1)
  const uint32_t blocks = cdiv(cdiv(count, 32), (uint32_t)tpb / 32u);
  g_launches += 1;
#define NB_TRAVERSE(EX, CNT)                                                                                  \
  traverse_kernel<EX, CNT><<<blocks, tpb, 0, c->stream>>>(c->T.box, c->T.node, c->meta, c->sbody, c->sacc,          \
                                                          c->cfg.theta, first, count)
#define NB_TRAVERSE_FAST(CNT)                                                                                     \
  do {                                                                                                            \
    if (transpose)                                                                                                \
      traverse_fast_kernel<CNT, true><<<blocks, tpb, 0, c->stream>>>(c->T.box, c->T.node, c->meta, c->sbody,      \
                                                                     c->sacc, first, count);                      \
    else                                                                                                          \
      traverse_fast_kernel<CNT, false><<<blocks, tpb, 0, c->stream>>>(c->T.box, c->T.node, c->meta, c->sbody,     \
                                                                      c->sacc, first, count);                     \
  } while (0)
  const bool ex = c->cfg.force_mode == NB_FORCE_EXACT;
  // transposed few-taker leaf path (traverse.cuh): re-orders sums, so not on the reference's own tree; fast_walk_group
  // == 3 switches it off for A/B runs
  const bool transpose = !c->bvh_ref && c->cfg.fast_walk_group != 3;
  if (!ex && !c->bvh_ref && c->cfg.fast_walk_group == 2) {
    // hybrid: far field by lane-parallel classification + dense lists, near field by the depth-first walk
    const uint32_t gblocks = cdiv(cdiv(count, 32), (uint32_t)kGrpWarps);
    if (c->counting)
      traverse_hybrid_kernel<true><<<gblocks, kGrpWarps * 32, 0, c->stream>>>(c->T.box, c->T.node, c->T.kids, c->meta, c->sbody, c->sacc, first, count);
    else
      traverse_hybrid_kernel<false><<<gblocks, kGrpWarps * 32, 0, c->stream>>>(c->T.box, c->T.node, c->T.kids, c->meta, c->sbody, c->sacc, first, count);
  } else
  if (!ex && !c->bvh_ref && c->cfg.fast_walk_group == 1) {
    // list-then-evaluate group walk (traverse_group.cuh): quadtree and LBVH in NB_FORCE_FAST
    const uint32_t gblocks = cdiv(cdiv(count, 32), (uint32_t)kGrpWarps);
    if (c->counting)
      traverse_group_kernel<true><<<gblocks, kGrpWarps * 32, 0, c->stream>>>(c->T.box, c->T.node, c->T.kids, c->meta, c->sbody, c->sacc, first, count);
    else
      traverse_group_kernel<false><<<gblocks, kGrpWarps * 32, 0, c->stream>>>(c->T.box, c->T.node, c->T.kids, c->meta, c->sbody, c->sacc, first, count);
  } else
  if (fuse) {
    if (c->counting)
      traverse_fast_push_kernel<true><<<blocks, tpb, 0, c->stream>>>(c->T.box, c->T.node, c->meta, c->sbody, first, count, *fuse);
    else
      traverse_fast_push_kernel<false><<<blocks, tpb, 0, c->stream>>>(c->T.box, c->T.node, c->meta, c->sbody, first, count, *fuse);
  } else
  if (ex && c->counting) NB_TRAVERSE(true, true);
  else if (ex) NB_TRAVERSE(true, false);
  else if (c->counting) NB_TRAVERSE_FAST(true);
  else NB_TRAVERSE_FAST(false);
#undef NB_TRAVERSE
#undef NB_TRAVERSE_FAST
  prof_mark(c->stream, "traverse");
  NB_CUDA(c, cudaGetLastError());
  return NB_OK;
}

int nccl_fail(nb_ctx *c, ncclResult_t r, const char *what) {
  return fail(c, NB_ERR_NCCL, std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"));
}

// integrate (main.rs:419-423) fused with the physical re-order; multi-GPU: the slices are pushed to the peers
// by the integrate kernel itself (nb_comm_peer_import) or all-gathered by NCCL afterwards.
int enqueue_integrate(nb_ctx *c, float dt, cudaEvent_t ev_after_integrate, bool fused = false, bool permuted = false) {
  const uint32_t n = c->n;
  if (n == 0) return NB_OK;
  uint32_t first, count;
  owned_range(c, first, count);
  BodyArrays &in = c->body[c->cur], &out = c->body[c->cur ^ 1];
  const uint32_t *order = c->vals[c->sorted_in];
  g_launches += (count && !fused ? 1 : 0) + (c->n_ranks > 1 ? 1 : 0);
  if (c->cfg.faithful_index_bug && count) { // single GPU, NB_TREE_BVH_REF (checked in nb_create / nb_comm_init)
    uint32_t *pos_of_slot = c->vals[c->sorted_in ^ 1];
    invert_order_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(order, n, pos_of_slot);
    integrate_resort_bug_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(order, pos_of_slot, c->sacc, in, out, n, dt, c->meta);
    g_launches += 1;
  } else if (count && !(c->n_ranks > 1 && c->peer))
    integrate_resort_kernel<<<cdiv(count, 256), 256, 0, c->stream>>>(order, c->sacc, in, out, first, count, dt, c->meta);
  if (c->n_ranks > 1 && c->peer) {
    // integrate + exchange in one kernel: every rank's copy of the output slice is stored from here (unless the
    // traversal kernel did it already: traverse_fast_push_kernel)
    if (count && !fused) {
      const FuseArgs f = make_fuse_args(c, dt);
      integrate_push_kernel<<<cdiv(count, 256), 256, 0, c->stream>>>(order, c->sacc, in, out, f.pin, f.po, c->n_ranks,
                                                                    f.per, first, count, dt, c->meta);
    }
    c->v_partial = true;
    if (permuted) NB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    else permute_mid_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(order, in, out, n, first, count, c->meta);
    if (ev_after_integrate) NB_CUDA(c, cudaEventRecord(ev_after_integrate, c->stream));
    ++c->peer_epoch;
    g_launches += 2;
    peer_signal_kernel<<<1, 32, 0, c->stream>>>(c->peer_flags_of, c->n_ranks, c->rank, c->peer_epoch);
    peer_wait_kernel<<<1, 32, 0, c->stream>>>(c->peer_flags, c->n_ranks, c->peer_epoch, c->meta);
  } else if (c->n_ranks > 1) {
    if (permuted) NB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    else permute_mid_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(order, in, out, n, first, count, c->meta);
    if (ev_after_integrate) NB_CUDA(c, cudaEventRecord(ev_after_integrate, c->stream));
    const size_t per = (size_t)cdiv(cdiv(n, (uint32_t)c->n_ranks), 32u) * 32u; // == nb_shard_range
    ncclResult_t r;
    if ((r = g_nccl.GroupStart()) != ncclSuccess) return nccl_fail(c, r, "ncclGroupStart");
    float *arrs[4] = {out.x, out.y, out.vx, out.vy};
    for (float *a : arrs) {
      r = g_nccl.AllGather(a + per * (size_t)c->rank, a, per, ncclFloat32, c->comm, c->stream);
      if (r != ncclSuccess) {
        g_nccl.GroupEnd(); // do not leave the group open behind the error
        return nccl_fail(c, r, "ncclAllGather");
      }
    }
    if ((r = g_nccl.GroupEnd()) != ncclSuccess) return nccl_fail(c, r, "ncclGroupEnd");
  } else if (ev_after_integrate) {
    NB_CUDA(c, cudaEventRecord(ev_after_integrate, c->stream));
  }
  prof_mark(c->stream, c->n_ranks > 1 ? "integrate + exchange" : "integrate_resort_kernel");
  NB_CUDA(c, cudaGetLastError());
  c->cur ^= 1;
  c->tree_valid = false;
  c->order_coherent = c->quad; // the new state is in the order of this step's tree (sort_repair.cuh)
  return NB_OK;
}

int sync_and_check(nb_ctx *c) {
  NB_CUDA(c, cudaMemcpyAsync(c->h_meta, c->meta, sizeof(Meta), cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  c->resort_hit = false;
  if (c->bmeta_in_flight) { // NB_TREE_BVH_REF, blind build: how many levels the last build enqueued really needed
    c->bmeta_in_flight = false;
    int used = 0;
    while (used < kBvhMaxLevels && c->h_bmeta->cnt_big[used] != 0) ++used;
    c->bvh_levels = (uint32_t)used;
  }
  if (c->h_meta->fail_seq != 0xffffffffu) { // some build since the last check guessed the depth too shallow
    c->resort_hit = true;
    c->resort_at = c->h_meta->fail_seq;
    c->depth_known = false; // the next build sorts every digit
    c->bvh_levels_known = false; // ... and NB_TREE_BVH_REF looks after every level again
    c->redo_next = true;
    c->redo_guess = c->last_top_passes;
    c->redo_guess_levels = c->last_top_levels;
    c->redo_guess_repaired = c->h_meta->repair_failed != 0;
    c->order_coherent = false; // the failed steps carried the state over: not in the order of the last build's keys
    c->movers_valid = false; // count again before the next repair
    if (c->h_meta->repair_failed) { // a burst of movers, a body left the root square, an over-full leaf: radix sort for a while
      c->repair_off = 256;
      NB_CUDA(c, cudaMemsetAsync(&c->meta->repair_failed, 0, sizeof(uint32_t), c->stream));
    }
    ++c->resorts;
    NB_CUDA(c, cudaMemsetAsync(&c->meta->fail_seq, 0xff, sizeof(uint32_t), c->stream));
    NB_CUDA(c, cudaMemsetAsync(&c->meta->error, 0, sizeof(uint32_t), c->stream));
    c->tree_valid = false;
  } else if (c->quad && c->n && c->h_meta->n_nodes) { // the mirror describes the last build enqueued
    c->last_max_depth = c->h_meta->max_depth;
    c->depth_known = true;
    if (c->last_counted) { // sort_repair.cuh: few enough bodies changed cells for the next builds to repair the order?
      c->movers_seen = c->h_meta->n_movers;
      c->movers_valid = true;
    }
  }
  if (c->h_meta->error & 4u) return fail(c, NB_ERR_NCCL, "peer exchange timed out: a rank did not signal its push within 10 s");
  if (c->h_meta->error & 1u) {
    return fail(c, NB_ERR_OVERFLOW,
                "node pool too small: " + std::to_string(c->h_meta->n_nodes) + " nodes > capacity " +
                    std::to_string(c->T.capacity));
  }
  return NB_OK;
}

void absorb_counters(nb_ctx *c, nb_counting *out) {
  const unsigned long long ti = c->h_meta->interactions, tt = c->h_meta->node_tests;
  if (out) {
    out->interactions += ti - c->seen_interactions;
    out->node_tests += tt - c->seen_tests;
  }
  c->seen_interactions = ti;
  c->seen_tests = tt;
}

cudaEvent_t next_event(nb_ctx *c) {
  if (c->ev_used == c->ev.size()) {
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return nullptr;
    c->ev.push_back(e);
  }
  return c->ev[c->ev_used++];
}

// events are recorded 5 per step: start, after build, after traverse, after integrate, end
void drain_events(nb_ctx *c, nb_counting *out, size_t max_steps = (size_t)-1) {
  const size_t steps = std::min(c->ev_used / 5, max_steps);
  for (size_t s = 0; s < steps; ++s) {
    cudaEvent_t *e = &c->ev[5 * s];
    float ms = 0.f;
    if (out) {
      if (cudaEventElapsedTime(&ms, e[0], e[1]) == cudaSuccess) out->build_s += ms * 1e-3;
      if (cudaEventElapsedTime(&ms, e[1], e[2]) == cudaSuccess) out->traverse_s += ms * 1e-3;
      if (cudaEventElapsedTime(&ms, e[2], e[3]) == cudaSuccess) out->integrate_s += ms * 1e-3;
      if (cudaEventElapsedTime(&ms, e[3], e[4]) == cudaSuccess) out->exchange_s += ms * 1e-3;
    }
  }
  if (out && steps) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev[0], c->ev[5 * (steps - 1) + 4]) == cudaSuccess) out->total_s += ms * 1e-3;
  }
  c->ev_used = 0;
}

int do_steps(nb_ctx *c, uint32_t n_steps, float dt, nb_counting *out) {
  int rc;
  const unsigned long long launches0 = g_launches;
  uint32_t remaining = n_steps;
  while (remaining) {
    // enqueue a batch blind (no host synchronisation inside a step), then look once
    const uint32_t batch = std::min(remaining, 256u); // bounds the number of live events
    const uint32_t seq0 = c->step_seq;
    c->batch_steps = batch;
    for (uint32_t s = 0; s < batch; ++s) {
      cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr, e3 = nullptr, e4 = nullptr;
      if (out) {
        e0 = next_event(c); e1 = next_event(c); e2 = next_event(c); e3 = next_event(c); e4 = next_event(c);
        if (!e4) return fail(c, NB_ERR_CUDA, "cudaEventCreate failed");
        NB_CUDA(c, cudaEventRecord(e0, c->stream));
      }
      if ((rc = enqueue_build(c)) != NB_OK) return rc;
      if (out) NB_CUDA(c, cudaEventRecord(e1, c->stream));
      uint32_t first, count;
      owned_range(c, first, count);
      // multi-GPU: the masses / ids of the slots other ranks integrate are re-ordered here (every rank knows the order),
      // on a second stream beside the traversal — memory-bound work next to an issue-bound kernel
      const bool permuted = c->n_ranks > 1 && !c->cfg.faithful_index_bug && !c->no_fuse;
      if (permuted) {
        NB_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));
        NB_CUDA(c, cudaStreamWaitEvent(c->side, c->ev_fork, 0));
        permute_mid_kernel<<<cdiv(c->n, 256), 256, 0, c->side>>>(c->vals[c->sorted_in], c->body[c->cur], c->body[c->cur ^ 1], c->n,
                                                                 first, count, c->meta);
        NB_CUDA(c, cudaEventRecord(c->ev_join, c->side));
      }
      const bool fused = count != 0 && step_can_fuse(c);
      FuseArgs fa{};
      if (fused) fa = make_fuse_args(c, dt);
      if ((rc = enqueue_traverse(c, first, count, fused ? &fa : nullptr)) != NB_OK) return rc;
      if (out) NB_CUDA(c, cudaEventRecord(e2, c->stream));
      if ((rc = enqueue_integrate(c, dt, e3, fused, permuted)) != NB_OK) return rc;
      if (out) NB_CUDA(c, cudaEventRecord(e4, c->stream));
    }
    if ((rc = sync_and_check(c)) != NB_OK) return rc;
    // steps from the first too-shallow depth guess on did nothing (state carried over): redo them, full sort first
    const uint32_t done = c->resort_hit ? std::min(batch, c->resort_at - seq0) : batch;
    drain_events(c, out, done);
    remaining -= done;
  }
  absorb_counters(c, out);
  if (out) {
    out->steps += n_steps;
    out->launches += g_launches - launches0;
  }
  return NB_OK;
}

__global__ void draw_compose_kernel(DrawScratch d, const uint32_t *__restrict__ slot_of, BodyArrays b,
                                    uint32_t npix, uint8_t *__restrict__ frame) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npix) return;
  uchar4 px = make_uchar4(0, 0, 0, 0); // frame zeroed first, main.rs:43
  if (d.heavy[p]) {
    px = make_uchar4(0x00, 0xff, 0x00, 0xff); // main.rs:55-59
  } else if (d.hits[p]) {
    const uint32_t s = slot_of ? slot_of[d.last[p] - 1u] : d.last[p] - 1u; // null: `last` is a slot already
    // main.rs:61-63: 0x10 + (((|vx|+|vy|)*10.0) as u8).min(0xef); `as u8` saturates, NaN -> 0
    const float sp = __fmul_rn(__fadd_rn(fabsf(b.vx[s]), fabsf(b.vy[s])), 10.0f);
    uint32_t q = (sp != sp) ? 0u : (sp <= 0.f ? 0u : (sp >= 255.f ? 255u : (uint32_t)sp));
    q = q < 0xefu ? q : 0xefu;
    const uint32_t vel = 0x10u + q;
    const uint32_t hits = d.hits[p];
    const uint32_t alpha = hits >= 25u ? 250u : hits * 10u; // += 10 while <= 240, main.rs:67-69
    px = make_uchar4(0xff, (unsigned char)(0xffu - vel), (unsigned char)(0xffu - vel), (unsigned char)alpha);
  }
  reinterpret_cast<uchar4 *>(frame)[p] = px;
}

// QuadTree::empty / prune (quad_tree.rs:66-137) as statistics of the freshly built tree: what the two methods
// would RETURN if the reference ran them on a tree holding exactly this step's bodies after calculate_gravity.
//   empty(): 1 per node visited; a Root whose total_mass is 0 returns at once (:78-80), so a node is visited iff
//            every proper ancestor has a non-zero (u32, wrapping) mass.
//   prune(): 1 per child it drops: Leaf children with count 0 (none exist in a fresh build: children are created
//            by the first insert into them) and Root children with total_mass 0 (:112-119), found while recursing
//            only through Roots of non-zero mass below the top node (:120-121).
__global__ void __launch_bounds__(256) quad_stats_kernel(const Meta *__restrict__ meta, TreeArrays T,
                                                         unsigned long long *__restrict__ out /* [2] */) {
  const uint32_t n_nodes = meta->n_nodes;
  uint32_t e = 0, p = 0;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_nodes; i += gridDim.x * blockDim.x) {
    const uint4 l = T.link(i);
    bool anc_all = true, anc_below_top = true; // all proper ancestors / all except the top node have mass != 0
    for (uint32_t a = l.w; a != 0xffffffffu;) {
      const uint32_t pa = T.link(a).w;
      const bool nz = T.mass[a] != 0u;
      anc_all = anc_all && nz;
      if (pa != 0xffffffffu) anc_below_top = anc_below_top && nz;
      a = pa;
    }
    e += anc_all ? 1u : 0u;
    if (i != 0u && !(l.z & kLeafBit) && T.mass[i] == 0u && anc_below_top) p += 1u;
  }
  e = __reduce_add_sync(0xffffffffu, e);
  p = __reduce_add_sync(0xffffffffu, p);
  if ((threadIdx.x & 31) == 0) {
    if (e) atomicAdd(&out[0], (unsigned long long)e);
    if (p) atomicAdd(&out[1], (unsigned long long)p);
  }
}

int quad_stats(nb_ctx *c, uint32_t *emptied, uint32_t *pruned) {
  if (!c->quad) return fail(c, NB_ERR_INVALID, "QuadTree::empty / prune: the context was not created with NB_TREE_QUAD");
  if (!c->tree_valid) return fail(c, NB_ERR_INVALID, "no valid tree: call nb_build_tree or nb_accel first");
  NB_CUDA(c, cudaSetDevice(c->device));
  unsigned long long *d = nullptr, h[2] = {0, 0};
  NB_CUDA(c, cudaMalloc(&d, 16));
  NB_CUDA(c, cudaMemsetAsync(d, 0, 16, c->stream));
  if (c->n) quad_stats_kernel<<<kSMs * 4, 256, 0, c->stream>>>(c->meta, c->T, d);
  NB_CUDA(c, cudaMemcpyAsync(h, d, 16, cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  cudaFree(d);
  if (c->n && c->h_meta->n_nodes == 0) h[0] = 1; // all bodies outside the square: the empty top Leaf (:68-72)
  if (!c->n) h[0] = 1;
  if (emptied) *emptied = (uint32_t)h[0];
  if (pruned) *pruned = (uint32_t)h[1];
  return NB_OK;
}

} // namespace

// ======================================================================================
// C ABI
// ======================================================================================
extern "C" {

int nb_abi_version(void) { return NB_ABI_VERSION; }

void nb_default_config(nb_config *cfg, uint64_t n_max) {
  std::memset(cfg, 0, sizeof(*cfg));
  cfg->n_max = n_max;
  cfg->theta = 50.0f;   // THETA, main.rs:35
  cfg->dt = 0.1f;       // STEP_SIZE, main.rs:34
  cfg->tree_kind = NB_TREE_QUAD;
  cfg->leaf_cap = 0;
  cfg->domain_x0 = 0.f;
  cfg->domain_y0 = 0.f;
  cfg->domain_h = 100000.f; // HEIGHT, main.rs:31
  cfg->force_mode = NB_FORCE_FAST;
}

const char *nb_last_error(const nb_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

void nb_destroy(nb_ctx *c) {
  if (!c) return;
  cudaSetDevice(c->device);
  if (c->stream) cudaStreamSynchronize(c->stream);
  if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
  free_bodies(c->body[0]);
  free_bodies(c->body[1]);
  for (int i = 0; i < 2; ++i) { cudaFree(c->keys[i]); cudaFree(c->vals[i]); }
  cudaFree(c->keys_new);
  cudaFree(c->leafdepth); cudaFree(c->nstart); cudaFree(c->sort.ghist); cudaFree(c->sort.tickets); cudaFree(c->repair.mkey); cudaFree(c->repair.mslot); cudaFree(c->repair.s_slot); cudaFree(c->repair.t_sorted); cudaFree(c->repair.k_sorted); cudaFree(c->repair.slot_sorted); cudaFree(c->sort.status); cudaFree(c->sort.bsum);
  cudaFree(c->T.box); cudaFree(c->T.node); cudaFree(c->T.mass); cudaFree(c->T.depth);
  cudaFree(c->T.bmax); cudaFree(c->T.arrive); cudaFree(c->T.kids);
  for (void *p : c->peer_mapped) cudaIpcCloseMemHandle(p);
  cudaFree(c->peer_flags);
  cudaFree(c->sbody); cudaFree(c->sacc); cudaFree(c->bounds_partial); cudaFree(c->meta); cudaFree(c->aos);
  cudaFree(c->flush_buf);
  if (c->h_meta) cudaFreeHost(c->h_meta);
  if (c->h_bmeta) cudaFreeHost(c->h_bmeta);
  cudaFree(c->bw.nodes); cudaFree(c->bw.meta); cudaFree(c->bw.list_sub);
  for (int i = 0; i < 2; ++i) {
    cudaFree(c->bw.list_big[i]);
    cudaFree(c->tiles.sum[i]); cudaFree(c->tiles.guess[i]);
  }
  for (cudaEvent_t e : c->ev) cudaEventDestroy(e);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  if (c->side) cudaStreamDestroy(c->side);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

nb_ctx *nb_create(const nb_config *cfg) {
  if (!cfg || cfg->n_max == 0 || cfg->n_max > 0x3fff0000ull) {
    fail(nullptr, NB_ERR_INVALID, "nb_create: n_max must be in [1, 2^30)");
    return nullptr;
  }
  if (cfg->tree_kind != NB_TREE_QUAD && cfg->tree_kind != NB_TREE_LBVH && cfg->tree_kind != NB_TREE_BVH_REF) {
    fail(nullptr, NB_ERR_INVALID, "nb_create: unknown tree_kind");
    return nullptr;
  }
  if (cfg->leaf_cap < 0 || cfg->leaf_cap > 64) {
    fail(nullptr, NB_ERR_INVALID, "nb_create: leaf_cap must be in [0, 64]");
    return nullptr;
  }
  if (cfg->faithful_index_bug && cfg->tree_kind != NB_TREE_BVH_REF) {
    fail(nullptr, NB_ERR_INVALID, "nb_create: faithful_index_bug reproduces main.rs:398-423 and needs NB_TREE_BVH_REF");
    return nullptr;
  }
  if (cfg->tree_kind == NB_TREE_BVH_REF && cfg->leaf_cap != 0 && cfg->leaf_cap != 64) {
    fail(nullptr, NB_ERR_INVALID, "nb_create: NB_TREE_BVH_REF uses the reference's TARGET_POINTS = 64 (leaf_cap 0 or 64)");
    return nullptr;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    fail(nullptr, NB_ERR_CUDA,
         std::string("nb_create: no CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count 0") +
             "); this library has no CPU fallback");
    return nullptr;
  }
  if (cfg->device < 0 || cfg->device >= ndev) {
    fail(nullptr, NB_ERR_INVALID, "nb_create: device ordinal out of range");
    return nullptr;
  }
  nb_ctx *c = new nb_ctx();
  c->cfg = *cfg;
  c->device = cfg->device;
  if (const char *e = std::getenv("NB_B200_NO_FUSE")) c->no_fuse = e[0] == '1';
  if (const char *e = std::getenv("NB_B200_NO_REPAIR")) c->no_repair = e[0] == '1';
  if (const char *e = std::getenv("NB_B200_SORT_MIN_LEVELS")) c->sort_min_levels = std::min(24, std::max(8, std::atoi(e)));
  if (const char *e = std::getenv("NB_B200_AOS_PERMUTE_MIN")) c->aos_permute_min = (uint32_t)std::strtoul(e, nullptr, 10);
  c->quad = cfg->tree_kind == NB_TREE_QUAD;
  c->bvh_ref = cfg->tree_kind == NB_TREE_BVH_REF;
  c->leaf_cap = cfg->leaf_cap ? cfg->leaf_cap : (c->quad ? 8 : 64); // MAX_CAPACITY / TARGET_POINTS
  if (c->cfg.domain_h <= 0.f) { c->cfg.domain_x0 = 0.f; c->cfg.domain_y0 = 0.f; c->cfg.domain_h = 100000.f; }
  auto bail = [&](const char *what, cudaError_t err) -> nb_ctx * {
    fail(nullptr, NB_ERR_CUDA, std::string("nb_create: ") + what + ": " + cudaGetErrorString(err));
    nb_destroy(c);
    return nullptr;
  };
#define NB_TRY(call)                                      \
  if ((e = (call)) != cudaSuccess) return bail(#call, e)
  NB_TRY(cudaSetDevice(c->device));
  NB_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  NB_TRY(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
  NB_TRY(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  NB_TRY(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  const size_t na = ((size_t)cfg->n_max + 64 + 32 * kMaxPeers + 1023) / 1024 * 1024; // + the all-gather's rounded-up slices
  c->n_alloc = (uint32_t)na;
  for (int i = 0; i < 2; ++i) {
    if (alloc_bodies(c, c->body[i], na) != NB_OK) { g_create_error = c->err; nb_destroy(c); return nullptr; }
    NB_TRY(dmalloc(&c->keys[i], na));
    if (i == 0) NB_TRY(dmalloc(&c->keys_new, na));
    NB_TRY(dmalloc(&c->vals[i], na));
  }
  NB_TRY(dmalloc(&c->leafdepth, na));
  NB_TRY(dmalloc(&c->nstart, na));
  c->sort.status_words = sort_status_words((uint32_t)na);
  NB_TRY(dmalloc(&c->sort.ghist, (size_t)8192));
  NB_TRY(dmalloc(&c->sort.tickets, (size_t)16));
  NB_TRY(dmalloc(&c->repair.mkey, (size_t)kRepairMax));
  NB_TRY(dmalloc(&c->repair.mslot, (size_t)kRepairMax));
  NB_TRY(dmalloc(&c->repair.s_slot, (size_t)kRepairMax));
  NB_TRY(dmalloc(&c->repair.t_sorted, (size_t)kRepairMax));
  NB_TRY(dmalloc(&c->repair.k_sorted, (size_t)kRepairMax));
  NB_TRY(dmalloc(&c->repair.slot_sorted, (size_t)kRepairMax));
  NB_TRY(dmalloc(&c->sort.status, c->sort.status_words));
  NB_TRY(dmalloc(&c->sort.bsum, na / 256 + na / 262144 + 64)); // block totals of tree_count_kernel (256 bodies) / of the flag scan
  const size_t ncap = 2 * na + 65536;
  c->T.capacity = (uint32_t)std::min<size_t>(ncap, 0xfffffff0u);
  NB_TRY(dmalloc(&c->T.box, (size_t)c->T.capacity));
  NB_TRY(dmalloc(&c->T.node, 2 * (size_t)c->T.capacity));
  NB_TRY(dmalloc(&c->T.mass, (size_t)c->T.capacity));
  NB_TRY(dmalloc(&c->T.depth, (size_t)c->T.capacity));
  NB_TRY(dmalloc(&c->T.bmax, (size_t)c->T.capacity));
  NB_TRY(dmalloc(&c->T.arrive, (size_t)c->T.capacity));
  if (!c->bvh_ref) NB_TRY(dmalloc(&c->T.kids, (size_t)c->T.capacity)); // children table: up-sweep (quad) and traverse_group_kernel
  NB_TRY(dmalloc(&c->sbody, na));
  NB_TRY(dmalloc(&c->sacc, na));
  NB_TRY(dmalloc(&c->bounds_partial, (size_t)kSMs * 4));
  NB_TRY(dmalloc(&c->meta, 1));
  NB_TRY(dmalloc(&c->aos, 6 * na));
  NB_TRY(cudaMallocHost((void **)&c->h_meta, sizeof(Meta)));
  if (c->bvh_ref) {
    // working arrays alias the Morton pipeline's buffers, which this tree kind does not use
    BvhWork &W = c->bw;
    W.wx = reinterpret_cast<float *>(c->keys[0]);
    W.wy = W.wx + na;
    W.F = reinterpret_cast<uint32_t *>(c->keys[1]);
    W.listT = W.F + na;
    W.order = c->vals[0];
    W.seg = c->nstart;
    W.bsum = c->sort.bsum;
    W.node_cap = (uint32_t)std::min<size_t>(na / 4 + 65536, c->T.capacity);
    NB_TRY(dmalloc(&W.nodes, (size_t)W.node_cap));
    const size_t lcap = na / 64 + 1024;
    for (int i = 0; i < 2; ++i) {
      NB_TRY(dmalloc(&W.list_big[i], lcap));
    }
    NB_TRY(dmalloc(&W.list_sub, lcap));
    NB_TRY(dmalloc(&W.meta, 1));
    NB_TRY(cudaMallocHost((void **)&c->h_bmeta, sizeof(BvhMeta)));
    NB_TRY(cudaFuncSetAttribute(bvh_subtree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SubSmem)));
    const size_t slots = (na >> kTileSlotShift) + 8;
    for (int i = 0; i < 2; ++i) {
      NB_TRY(dmalloc(&c->tiles.sum[i], slots));
      NB_TRY(dmalloc(&c->tiles.guess[i], slots));
    }
  }
  Meta m0{};
  m0.fail_seq = 0xffffffffu;
  m0.x0 = c->cfg.domain_x0; m0.y0 = c->cfg.domain_y0; m0.h = c->cfg.domain_h;
  m0.theta2 = theta_squared(c->cfg.theta);
  m0.one = 1u;
  m0.node_base = c->T.node;
  NB_TRY(cudaMemcpy(c->meta, &m0, sizeof(Meta), cudaMemcpyHostToDevice));
  *c->h_meta = m0;
#undef NB_TRY
  return c;
}

int nb_set_params(nb_ctx *c, float theta, float dt, int32_t force_mode) {
  if (!c) return NB_ERR_INVALID;
  if (force_mode != NB_FORCE_EXACT && force_mode != NB_FORCE_FAST) return fail(c, NB_ERR_INVALID, "bad force_mode");
  c->cfg.theta = theta;
  c->cfg.dt = dt;
  c->cfg.force_mode = force_mode;
  NB_CUDA(c, cudaSetDevice(c->device));
  const float t2 = theta_squared(theta); // stream-ordered: kernels already enqueued keep the old value
  NB_CUDA(c, cudaMemcpyAsync(&c->meta->theta2, &t2, sizeof(float), cudaMemcpyHostToDevice, c->stream));
  c->h_meta->theta2 = t2;
  return NB_OK;
}

int nb_set_counting(nb_ctx *c, int enabled) {
  if (!c) return NB_ERR_INVALID;
  c->counting = enabled != 0;
  return NB_OK;
}

void *nb_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes) != cudaSuccess) return nullptr;
  return p;
}
void nb_host_free(void *p) {
  if (p) cudaFreeHost(p);
}

// FP32 pipe microbenchmark (see ubench.cuh).  *rate = lane-operations per second of the kind's unit:
// kinds 0/1 flop/s, kinds 2/3 pair interactions/s.
int nb_measure_fp32(nb_ctx *c, int kind, double *rate) {
  if (!c || !rate || kind < 0 || kind > 3) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  const int blocks = kSMs * 8, threads = 256, iters = 20000;
  float *out = reinterpret_cast<float *>(c->sacc);
  cudaEvent_t e0 = next_event(c), e1 = next_event(c);
  if (!e1) return fail(c, NB_ERR_CUDA, "cudaEventCreate failed");
  double best = 0.0;
  for (int rep = 0; rep < 4; ++rep) {
    NB_CUDA(c, cudaEventRecord(e0, c->stream));
    switch (kind) {
    case 0: fp32_ubench_kernel<0><<<blocks, threads, 0, c->stream>>>(out, iters, 0.5f); break;
    case 1: fp32_ubench_kernel<1><<<blocks, threads, 0, c->stream>>>(out, iters, 0.5f); break;
    case 2: fp32_ubench_kernel<2><<<blocks, threads, 0, c->stream>>>(out, iters, 0.5f); break;
    default: fp32_ubench_kernel<3><<<blocks, threads, 0, c->stream>>>(out, iters, 0.5f); break;
    }
    NB_CUDA(c, cudaEventRecord(e1, c->stream));
    NB_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    NB_CUDA(c, cudaEventElapsedTime(&ms, e0, e1));
    const double lanes = (double)blocks * threads * iters;
    const double per = kind == 0 ? 16.0 * 2.0 : kind == 1 ? 8.0 * 4.0 : 8.0; // flop or interactions per lane-iter
    const double r = lanes * per / (ms * 1e-3);
    if (rep > 0 && r > best) best = r;
  }
  c->ev_used = 0;
  *rate = best;
  return NB_OK;
}

int nb_flush_l2(nb_ctx *c) {
  if (!c) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  if (!c->flush_buf) {
    c->flush_bytes = (size_t)256 << 20;
    NB_CUDA(c, cudaMalloc(&c->flush_buf, c->flush_bytes));
  }
  NB_CUDA(c, cudaMemsetAsync(c->flush_buf, 0x5a, c->flush_bytes, c->stream));
  return NB_OK;
}

int nb_set_bodies_aos(nb_ctx *c, const nb_particle *bodies, size_t n) {
  if (!c || (!bodies && n)) return NB_ERR_INVALID;
  if (n > c->cfg.n_max) return fail(c, NB_ERR_CAPACITY, "nb_set_bodies: n > n_max");
  NB_CUDA(c, cudaSetDevice(c->device));
  c->n = (uint32_t)n;
  c->v_partial = false;
  c->aos_fresh = false;
  c->tree_valid = false;
  c->order_coherent = false;
  c->movers_valid = false;
  if (n == 0) return NB_OK;
  NB_CUDA(c, cudaMemcpyAsync(c->aos, bodies, n * sizeof(nb_particle), cudaMemcpyHostToDevice, c->stream));
  aos_to_soa_kernel<<<cdiv((uint32_t)n, 256), 256, 0, c->stream>>>(c->aos, c->body[c->cur], (uint32_t)n);
  c->aos_fresh = true;
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  return NB_OK;
}

int nb_set_bodies_soa(nb_ctx *c, const float *x, const float *y, const float *vx, const float *vy,
                      const uint32_t *w, size_t n) {
  if (!c || (n && (!x || !y || !vx || !vy || !w))) return NB_ERR_INVALID;
  if (n > c->cfg.n_max) return fail(c, NB_ERR_CAPACITY, "nb_set_bodies: n > n_max");
  NB_CUDA(c, cudaSetDevice(c->device));
  c->n = (uint32_t)n;
  c->v_partial = false;
  c->aos_fresh = false;
  c->tree_valid = false;
  c->order_coherent = false;
  c->movers_valid = false;
  if (n == 0) return NB_OK;
  BodyArrays &b = c->body[c->cur];
  NB_CUDA(c, cudaMemcpyAsync(b.x, x, n * 4, cudaMemcpyHostToDevice, c->stream));
  NB_CUDA(c, cudaMemcpyAsync(b.y, y, n * 4, cudaMemcpyHostToDevice, c->stream));
  NB_CUDA(c, cudaMemcpyAsync(b.vx, vx, n * 4, cudaMemcpyHostToDevice, c->stream));
  NB_CUDA(c, cudaMemcpyAsync(b.vy, vy, n * 4, cudaMemcpyHostToDevice, c->stream));
  NB_CUDA(c, cudaMemcpyAsync(b.m, w, n * 4, cudaMemcpyHostToDevice, c->stream));
  set_identity_kernel<<<cdiv((uint32_t)n, 256), 256, 0, c->stream>>>(b.id, (uint32_t)n);
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  return NB_OK;
}

// compaction of one candidate stream into body slots [slot0, slot0 + want): chunks of at most n_alloc candidates
// (prefix scratch = vals[0]); returns the number of accepted candidates among the first `limit` (0 = until `want`)
static int scene_compact(nb_ctx *c, int kind, uint64_t seed, uint32_t lo, uint32_t ny, uint32_t slot0, uint32_t want,
                         uint64_t limit, uint32_t *accepted) {
  cudaStream_t st = c->stream;
  uint32_t *pre = c->vals[0], *bt = c->sort.bsum, *emitted = c->sort.tickets, *chunk_total = c->sort.tickets + 1;
  NB_CUDA(c, cudaMemsetAsync(emitted, 0, 2 * sizeof(uint32_t), st));
  const uint32_t chunk = std::min<uint32_t>(c->n_alloc, 1u << 24) / 256u * 256u; // bsum holds n_alloc / 256 + 16 blocks
  uint64_t k0 = 0;
  uint32_t got = 0;
  for (int guard = 0; guard < 100000; ++guard) {
    const uint32_t cnt = (uint32_t)(limit ? std::min<uint64_t>(chunk, limit - k0) : chunk);
    if (cnt == 0) break;
    const uint32_t blocks = cdiv(cnt, 256);
    scene_flag_kernel<<<blocks, 256, 0, st>>>(kind, seed, k0, cnt, lo, ny, pre, bt);
    scan_bsum_kernel<<<1, 1024, 0, st>>>(bt, blocks, chunk_total);
    scene_emit_kernel<<<blocks, 256, 0, st>>>(kind, seed, k0, cnt, lo, ny, pre, bt, c->body[c->cur], slot0, want, emitted);
    scene_advance_kernel<<<1, 1, 0, st>>>(emitted, chunk_total);
    NB_CUDA(c, cudaGetLastError());
    k0 += cnt;
    if (limit && k0 >= limit) break;
    if (!limit) { // rejection sampling: stop once enough candidates were accepted (one look per chunk: set-up code)
      NB_CUDA(c, cudaMemcpyAsync(&got, emitted, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
      NB_CUDA(c, cudaStreamSynchronize(st));
      if (got >= want) break;
    }
  }
  NB_CUDA(c, cudaMemcpyAsync(&got, emitted, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  NB_CUDA(c, cudaStreamSynchronize(st));
  if (accepted) *accepted = got;
  return NB_OK;
}

int nb_scene_generate_device(nb_ctx *c, int32_t scene, uint64_t seed, size_t n) {
  if (!c) return NB_ERR_INVALID;
  if (scene != NB_SCENE_DEFAULT && scene != NB_SCENE_UNIFORM_DISC && scene != NB_SCENE_PLUMMER && scene != NB_SCENE_TWO_GALAXY)
    return fail(c, NB_ERR_INVALID, "nb_scene_generate_device: unknown scene");
  NB_CUDA(c, cudaSetDevice(c->device));
  c->tree_valid = false;
  c->order_coherent = false;
  c->movers_valid = false;
  c->v_partial = false;
  c->aos_fresh = false; // the staging buffer no longer describes the state
  cudaStream_t st = c->stream;
  int rc;
  if (scene == NB_SCENE_DEFAULT) { // World::new (main.rs:276-346): 2 cores + grid disc + 100 000 rand_body; n is ignored
    const uint32_t lim = (100000u / 14u) - 1u;
    const uint32_t lo = (uint32_t)((60000.0 - 3900.0) / 14.0), hi = std::min((uint32_t)((60000.0 + 3900.0) / 14.0) + 1u, lim - 1u);
    const uint32_t ny = hi - lo + 1u;
    const uint64_t cands = (uint64_t)ny * ny;
    if (c->cfg.n_max < 100002u) return fail(c, NB_ERR_CAPACITY, "nb_scene_generate_device: the default scene holds ~151k bodies, n_max is too small");
    const uint32_t room = (uint32_t)std::min<uint64_t>(cands, c->cfg.n_max - 100002u);
    uint32_t grid = 0;
    if ((rc = scene_compact(c, 0, seed, lo, ny, 2u, room, cands, &grid)) != NB_OK) return rc;
    if (grid > room) return fail(c, NB_ERR_CAPACITY, "nb_scene_generate_device: the default scene holds " + std::to_string(grid + 100002u) + " bodies > n_max");
    scene_fixed_kernel<<<cdiv(100000u, 256), 256, 0, st>>>(0, seed, c->body[c->cur], 2u + grid, 100000u);
    NB_CUDA(c, cudaGetLastError());
    c->n = 2u + grid + 100000u;
  } else if (scene == NB_SCENE_TWO_GALAXY) {
    if (n < 2) n = 2;
    if (n > c->cfg.n_max) return fail(c, NB_ERR_CAPACITY, "nb_scene_generate_device: n > n_max");
    const uint32_t per0 = (uint32_t)((n - 2) / 2 + ((n - 2) & 1)), per1 = (uint32_t)((n - 2) / 2);
    scene_fixed_kernel<<<1, 32, 0, st>>>(1, seed, c->body[c->cur], 0u, 0u);
    if (per0 && (rc = scene_compact(c, 1, seed, 0u, 1u, 2u, per0, 0, nullptr)) != NB_OK) return rc;
    if (per1 && (rc = scene_compact(c, 2, seed, 0u, 1u, 2u + per0, per1, 0, nullptr)) != NB_OK) return rc;
    c->n = (uint32_t)n;
  } else {
    if (n > c->cfg.n_max) return fail(c, NB_ERR_CAPACITY, "nb_scene_generate_device: n > n_max");
    c->n = (uint32_t)n;
    if (n == 0) return NB_OK;
    scene_kernel<<<cdiv((uint32_t)n, 256), 256, 0, st>>>(scene, seed, c->body[c->cur], (uint32_t)n);
  }
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaStreamSynchronize(st));
  return NB_OK;
}

size_t nb_body_count(const nb_ctx *c) { return c ? c->n : 0; }

int nb_get_bodies_aos(nb_ctx *c, nb_particle *bodies, size_t n) {
  if (!c || (!bodies && n)) return NB_ERR_INVALID;
  if (n < c->n) return fail(c, NB_ERR_CAPACITY, "nb_get_bodies: buffer smaller than body count");
  NB_CUDA(c, cudaSetDevice(c->device));
  if (c->n == 0) return NB_OK;
  { int rv = ensure_velocities(c); if (rv != NB_OK) return rv; }
  soa_to_aos_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(c->body[c->cur], c->n, c->aos);
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaMemcpyAsync(bodies, c->aos, (size_t)c->n * sizeof(nb_particle), cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  return NB_OK;
}

int nb_get_bodies_soa(nb_ctx *c, float *x, float *y, float *vx, float *vy, uint32_t *w, size_t n) {
  if (!c) return NB_ERR_INVALID;
  if (n < c->n) return fail(c, NB_ERR_CAPACITY, "nb_get_bodies: buffer smaller than body count");
  NB_CUDA(c, cudaSetDevice(c->device));
  if (c->n == 0) return NB_OK;
  { int rv = ensure_velocities(c); if (rv != NB_OK) return rv; }
  BodyArrays &b = c->body[c->cur];
  uint32_t *tmp = c->vals[c->sorted_in ^ 1]; // scratch (the tree order lives in vals[sorted_in])
  const uint32_t *srcs[5] = {(const uint32_t *)b.x, (const uint32_t *)b.y, (const uint32_t *)b.vx,
                             (const uint32_t *)b.vy, b.m};
  void *dsts[5] = {x, y, vx, vy, w};
  for (int k = 0; k < 5; ++k) {
    if (!dsts[k]) continue;
    unpermute_word_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(srcs[k], b.id, c->n, tmp);
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaMemcpyAsync(dsts[k], tmp, (size_t)c->n * 4, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return NB_OK;
}

int nb_step(nb_ctx *c, uint32_t n_steps, nb_counting *out) {
  if (!c) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  return do_steps(c, n_steps, c->cfg.dt, out);
}

int nb_profile_step(nb_ctx *c, nb_kernel_time *out, size_t cap, size_t *count) {
  if (!c || !out || !count) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  std::vector<ProfMark> marks;
  g_prof = &marks;
  prof_mark(c->stream, "start");
  const int rc = do_steps(c, 1, c->cfg.dt, nullptr);
  g_prof = nullptr;
  size_t k = 0;
  for (size_t i = 1; i < marks.size(); ++i) {
    float ms = 0.f;
    if (rc == NB_OK && k < cap && cudaEventElapsedTime(&ms, marks[i - 1].ev, marks[i].ev) == cudaSuccess) {
      std::memset(&out[k], 0, sizeof(out[k]));
      std::strncpy(out[k].name, marks[i].name, sizeof(out[k].name) - 1);
      out[k].us = (double)ms * 1e3;
      ++k;
    }
  }
  for (ProfMark &m : marks) cudaEventDestroy(m.ev);
  *count = k;
  return rc;
}

int nb_world_update(nb_ctx *c, nb_particle *bodies, size_t n, float delta, nb_counting *out) {
  if (!c || (!bodies && n)) return NB_ERR_INVALID;
  if (n > c->cfg.n_max) return fail(c, NB_ERR_CAPACITY, "nb_world_update: n > n_max");
  NB_CUDA(c, cudaSetDevice(c->device));
  c->n = (uint32_t)n;
  c->v_partial = false;
  c->aos_fresh = false;
  c->tree_valid = false;
  c->order_coherent = false;
  c->movers_valid = false;
  if (n == 0) return NB_OK;
  NB_CUDA(c, cudaMemcpyAsync(c->aos, bodies, n * sizeof(nb_particle), cudaMemcpyHostToDevice, c->stream));
  aos_to_soa_kernel<<<cdiv((uint32_t)n, 256), 256, 0, c->stream>>>(c->aos, c->body[c->cur], (uint32_t)n);
  c->aos_fresh = true;
  int rc = do_steps(c, 1, delta, out);
  if (rc != NB_OK) return rc;
  // NB_TREE_BVH_REF: the caller's Vec comes back permuted exactly as BVHTree::from leaves it (main.rs:400 hands
  // `self.particles.as_mut()` to the build), because that order is what the next call's tree is built from
  { int rv = ensure_velocities(c); if (rv != NB_OK) return rv; }
  if (c->bvh_ref) soa_to_aos_slot_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(c->body[c->cur], c->n, c->aos);
  else soa_to_aos_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(c->body[c->cur], c->n, c->aos);
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaMemcpyAsync(bodies, c->aos, n * sizeof(nb_particle), cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  return NB_OK;
}

// World::update for a Vec that is sharded over the ranks' hosts: rank r passes bodies [first, first + count) of the
// n_total-body Vec.  Upload 1/G, NVLink to everybody, step, download 1/G (original order).  Collective.
int nb_world_update_shard(nb_ctx *c, nb_particle *slice, size_t first, size_t count, size_t n_total, float delta,
                          nb_counting *out) {
  if (!c || (!slice && count)) return NB_ERR_INVALID;
  if (n_total > c->cfg.n_max || first + count > n_total) return fail(c, NB_ERR_CAPACITY, "nb_world_update_shard: slice outside n_total <= n_max");
  if (c->n_ranks < 2 || !c->peer) return fail(c, NB_ERR_INVALID, "nb_world_update_shard: needs the peer mappings (nb_comm_peer_import)");
  if (c->bvh_ref) return fail(c, NB_ERR_INVALID, "nb_world_update_shard: NB_TREE_BVH_REF keeps the Vec order, use nb_world_update");
  NB_CUDA(c, cudaSetDevice(c->device));
  c->n = (uint32_t)n_total;
  c->v_partial = false;
  c->aos_fresh = false;
  c->tree_valid = false;
  c->order_coherent = false;
  c->movers_valid = false;
  if (n_total == 0) return NB_OK;
  cudaStream_t st = c->stream;
  auto handshake = [&]() {
    ++c->peer_epoch;
    peer_signal_kernel<<<1, 32, 0, st>>>(c->peer_flags_of, c->n_ranks, c->rank, c->peer_epoch);
    peer_wait_kernel<<<1, 32, 0, st>>>(c->peer_flags, c->n_ranks, c->peer_epoch, c->meta);
    g_launches += 2;
  };
  // NB_B200_SHARD_TIMING=1: device time of the call's phases on stderr (debug aid; events only when asked for)
  static const bool timing = std::getenv("NB_B200_SHARD_TIMING") != nullptr;
  cudaEvent_t tev[6] = {};
  auto mark = [&](int k) {
    if (!timing) return;
    cudaEventCreate(&tev[k]);
    cudaEventRecord(tev[k], st);
  };
  mark(0);
  NB_CUDA(c, cudaMemsetAsync(&c->meta->error, 0, sizeof(uint32_t), st));
  if (count) NB_CUDA(c, cudaMemcpyAsync(c->aos + 6 * first, slice, count * sizeof(nb_particle), cudaMemcpyHostToDevice, st));
  mark(1);
  handshake(); // every rank is inside this call: nobody still reads the arrays the pushes below overwrite
  mark(2);
  if (count)
    push_shard_kernel<<<cdiv((uint32_t)count, 256), 256, 0, st>>>(c->aos, c->peer_out[c->cur], c->peer_out[c->cur ^ 1],
                                                                 c->n_ranks, (uint32_t)first, (uint32_t)count);
  handshake(); // every rank's slice has arrived everywhere
  unpack_shard_kernel<<<cdiv(c->n, 256), 256, 0, st>>>(c->body[c->cur], c->body[c->cur ^ 1].vx, c->n);
  g_launches += 2;
  mark(3);
  int rc = do_steps(c, 1, delta, out);
  if (rc != NB_OK) return rc;
  mark(4);
  if (count) {
    soa_to_aos_range_kernel<<<cdiv(c->n, 256), 256, 0, st>>>(c->body[c->cur], c->n, c->aos, (uint32_t)first, (uint32_t)count,
                                                             c->peer_out[c->cur], c->n_ranks, shard_per(c));
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaMemcpyAsync(slice, c->aos + 6 * first, count * sizeof(nb_particle), cudaMemcpyDeviceToHost, st));
  }
  mark(5);
  NB_CUDA(c, cudaStreamSynchronize(st));
  if (timing) {
    float ms[5] = {};
    for (int k = 0; k < 5; ++k) cudaEventElapsedTime(&ms[k], tev[k], tev[k + 1]);
    std::fprintf(stderr, "[shard timing] rank %d: h2d %.3f  wait-all-in %.3f  push+wait %.3f  step %.3f  gather+d2h %.3f ms\n",
                 c->rank, ms[0], ms[1], ms[2], ms[3], ms[4]);
    for (cudaEvent_t e : tev) cudaEventDestroy(e);
  }
  return NB_OK;
}

int nb_build_tree(nb_ctx *c, nb_counting *out) {
  if (!c) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  int rc = NB_OK;
  for (int attempt = 0; attempt < 2; ++attempt) { // a too-shallow depth guess is redone with a full sort
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (out) {
      e0 = next_event(c); e1 = next_event(c);
      if (!e1) return fail(c, NB_ERR_CUDA, "cudaEventCreate failed");
      NB_CUDA(c, cudaEventRecord(e0, c->stream));
    }
    c->batch_steps = 1;
    rc = enqueue_build(c);
    if (rc != NB_OK) return rc;
    if (out) NB_CUDA(c, cudaEventRecord(e1, c->stream));
    rc = sync_and_check(c);
    if (out && rc == NB_OK) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) out->build_s += ms * 1e-3;
    }
    c->ev_used = 0;
    if (rc != NB_OK || !c->resort_hit) break;
  }
  return rc;
}

int nb_accel(nb_ctx *c, float *ax, float *ay, nb_counting *out) {
  if (!c || !ax || !ay) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  if (c->n == 0) return NB_OK;
  int rc;
  for (int attempt = 0; attempt < 2; ++attempt) { // a too-shallow depth guess is redone with a full sort
    cudaEvent_t e0 = next_event(c), e1 = next_event(c), e2 = next_event(c);
    if (!e2) return fail(c, NB_ERR_CUDA, "cudaEventCreate failed");
    NB_CUDA(c, cudaEventRecord(e0, c->stream));
    c->batch_steps = 1;
    if (!c->tree_valid && (rc = enqueue_build(c)) != NB_OK) return rc;
    NB_CUDA(c, cudaEventRecord(e1, c->stream));
    if ((rc = enqueue_traverse(c, 0, c->n)) != NB_OK) return rc;
    NB_CUDA(c, cudaEventRecord(e2, c->stream));
    // scratch: the unsorted ping-pong buffers are free once the tree is built
    float *dax = reinterpret_cast<float *>(c->keys[c->sorted_in ^ 1]);
    float *day = dax + c->n_alloc;
    accel_to_original_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(c->sacc, c->vals[c->sorted_in],
                                                                    c->body[c->cur].id, c->n, dax, day);
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaMemcpyAsync(ax, dax, (size_t)c->n * 4, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(c, cudaMemcpyAsync(ay, day, (size_t)c->n * 4, cudaMemcpyDeviceToHost, c->stream));
    if ((rc = sync_and_check(c)) != NB_OK) return rc;
    if (c->resort_hit) { // the traversal did nothing: start over (full sort)
      c->ev_used = 0;
      continue;
    }
    if (out) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) out->build_s += ms * 1e-3;
      if (cudaEventElapsedTime(&ms, e1, e2) == cudaSuccess) out->traverse_s += ms * 1e-3;
    }
    break;
  }
  c->ev_used = 0;
  absorb_counters(c, out);
  return NB_OK;
}

int nb_accel_direct(nb_ctx *c, size_t first, size_t count, float *ax, float *ay, double *seconds) {
  if (!c || !ax || !ay) return NB_ERR_INVALID;
  if (first + count > c->n) return fail(c, NB_ERR_CAPACITY, "nb_accel_direct: target range exceeds body count");
  NB_CUDA(c, cudaSetDevice(c->device));
  if (count == 0) return NB_OK;
  const uint32_t n = c->n, nt = (uint32_t)count;
  BodyArrays &b = c->body[c->cur];
  uint32_t *slot_of = nullptr;
  double *acc = nullptr;
  float *dax = nullptr;
  struct Scratch { // freed on every return path
    uint32_t *&a; double *&b; float *&d;
    ~Scratch() { cudaFree(a); cudaFree(b); cudaFree(d); }
  } scratch{slot_of, acc, dax};
  NB_CUDA(c, dmalloc(&slot_of, (size_t)n));
  NB_CUDA(c, dmalloc(&acc, 2 * (size_t)nt));
  NB_CUDA(c, dmalloc(&dax, 2 * (size_t)nt));
  slot_of_id_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(b.id, n, slot_of);
  NB_CUDA(c, cudaMemsetAsync(acc, 0, 2 * (size_t)nt * sizeof(double), c->stream));
  // split the sources so that the grid fills the machine even for few targets
  const uint32_t tblocks = cdiv(nt, kDirectThreads);
  uint32_t ychunks = std::max(1u, std::min(cdiv(n, 4096u), cdiv((uint32_t)kSMs * 8u, tblocks)));
  ychunks = std::min(ychunks, 65535u);
  uint32_t chunk = cdiv(cdiv(n, ychunks), kDirectThreads) * kDirectThreads;
  ychunks = cdiv(n, chunk);
  cudaEvent_t e0 = next_event(c), e1 = next_event(c);
  if (!e1) return fail(c, NB_ERR_CUDA, "cudaEventCreate failed");
  NB_CUDA(c, cudaEventRecord(e0, c->stream));
  dim3 grid(tblocks, ychunks);
  if (c->cfg.force_mode == NB_FORCE_EXACT)
    direct_kernel<true><<<grid, kDirectThreads, 0, c->stream>>>(b.x, b.y, b.m, n, slot_of + first, nt, acc, chunk);
  else
    direct_kernel<false><<<grid, kDirectThreads, 0, c->stream>>>(b.x, b.y, b.m, n, slot_of + first, nt, acc, chunk);
  NB_CUDA(c, cudaEventRecord(e1, c->stream));
  direct_finish_kernel<<<cdiv(nt, 256), 256, 0, c->stream>>>(acc, nt, dax, dax + nt);
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaMemcpyAsync(ax, dax, (size_t)nt * 4, cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaMemcpyAsync(ay, dax + nt, (size_t)nt * 4, cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  if (seconds) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) *seconds = ms * 1e-3;
  }
  c->ev_used = 0;
  return NB_OK;
}

int nb_quad_empty(nb_ctx *c, uint32_t *nodes_visited) {
  if (!c || !nodes_visited) return NB_ERR_INVALID;
  return quad_stats(c, nodes_visited, nullptr);
}
int nb_quad_prune(nb_ctx *c, uint32_t *children_pruned) {
  if (!c || !children_pruned) return NB_ERR_INVALID;
  return quad_stats(c, nullptr, children_pruned);
}

int nb_get_tree_info(nb_ctx *c, nb_tree_info *info) {
  if (!c || !info) return NB_ERR_INVALID;
  if (!c->tree_valid) return fail(c, NB_ERR_INVALID, "no valid tree: call nb_build_tree or nb_accel first");
  NB_CUDA(c, cudaSetDevice(c->device));
  int rc = sync_and_check(c);
  if (rc != NB_OK) return rc;
  std::memset(info, 0, sizeof(*info));
  const Meta &m = *c->h_meta;
  info->n_bodies = c->n;
  info->n_in_tree = c->n ? m.n_in : 0;
  info->n_nodes = c->n ? m.n_nodes : 0;
  info->n_leaves = c->n ? m.n_leaves : 0;
  info->max_depth = c->n ? m.max_depth : 0;
  info->overflow_leaves = c->n ? m.overflow_leaves : 0;
  info->root_x0 = m.x0; info->root_y0 = m.y0; info->root_h = m.h;
  info->sort_top_digits = c->n ? (uint32_t)(c->last_redone ? c->last_failed_guess : c->last_top_passes) : 0;
  info->sort_top_levels = c->n ? (uint32_t)(c->last_redone ? c->last_failed_levels : c->last_top_levels) : 0;
  info->sort_repaired = c->n ? (uint32_t)(c->last_redone ? c->last_failed_repaired : c->last_repaired) : 0;
  info->sort_movers = info->sort_repaired && !c->last_redone ? c->h_meta->n_movers : 0;
  info->sort_redone = (c->n && c->last_redone) ? 1u : 0u;
  return NB_OK;
}

int nb_get_tree(nb_ctx *c, const nb_tree_view *v) {
  if (!c || !v) return NB_ERR_INVALID;
  nb_tree_info info;
  int rc = nb_get_tree_info(c, &info);
  if (rc != NB_OK) return rc;
  const uint32_t n = c->n, nn = (uint32_t)info.n_nodes;
  if (n == 0) return NB_OK;
  BodyArrays &b = c->body[c->cur];
  const int w = c->sorted_in;
  if (v->keys && c->bvh_ref) std::memset(v->keys, 0, (size_t)n * 8); // no Morton keys in this tree
  if (v->keys && !c->bvh_ref) {
    // recompute the slot-order keys into the free ping-pong buffer, then to original order
    unsigned long long *tmp = nullptr;
    NB_CUDA(c, dmalloc(&tmp, (size_t)n));
    morton_kernel<kQuadLevels><<<cdiv(cdiv(n, 4), 256), 256, 0, c->stream>>>(b.x, b.y, c->meta, n, c->quad ? 1 : 0, c->keys[w ^ 1],
                                                                  c->vals[w ^ 1]);
    unpermute_key_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(c->keys[w ^ 1], b.id, n, tmp);
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaMemcpyAsync(v->keys, tmp, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(tmp);
  }
  if (v->order) {
    uint32_t *tmp = c->vals[w ^ 1];
    order_to_original_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(c->vals[w], b.id, n, tmp);
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaMemcpyAsync(v->order, tmp, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(c, cudaStreamSynchronize(c->stream));
    // bodies outside the root square: canonical (ascending id) order
    std::sort(v->order + info.n_in_tree, v->order + n);
  }
  if (nn == 0) return NB_OK;
  std::vector<uint4> link(nn);
  std::vector<float4> box(nn), cm(nn), inter(2 * (size_t)nn);
  NB_CUDA(c, cudaMemcpy(inter.data(), c->T.node, 2 * (size_t)nn * sizeof(float4), cudaMemcpyDeviceToHost));
  for (uint32_t i = 0; i < nn; ++i) {
    cm[i] = inter[2 * (size_t)i];
    std::memcpy(&link[i], &inter[2 * (size_t)i + 1], sizeof(uint4));
  }
  NB_CUDA(c, cudaMemcpy(box.data(), c->T.box, (size_t)nn * sizeof(float4), cudaMemcpyDeviceToHost));
  if (v->node_depth) NB_CUDA(c, cudaMemcpy(v->node_depth, c->T.depth, (size_t)nn * 4, cudaMemcpyDeviceToHost));
  if (v->node_mass) NB_CUDA(c, cudaMemcpy(v->node_mass, c->T.mass, (size_t)nn * 4, cudaMemcpyDeviceToHost));
  for (uint32_t i = 0; i < nn; ++i) {
    if (v->node_skip) v->node_skip[i] = link[i].x;
    if (v->node_start) v->node_start[i] = link[i].y;
    if (v->node_count) v->node_count[i] = link[i].z & kCountMask;
    if (v->node_leaf) v->node_leaf[i] = (link[i].z & kLeafBit) ? 1 : 0;
    if (v->node_parent) v->node_parent[i] = link[i].w;
    if (v->node_cogx) v->node_cogx[i] = cm[i].x;
    if (v->node_cogy) v->node_cogy[i] = cm[i].y;
    if (v->node_s2) v->node_s2[i] = cm[i].w;
    if (v->node_x0) v->node_x0[i] = box[i].x;
    if (v->node_y0) v->node_y0[i] = box[i].y;
    if (v->node_x1) v->node_x1[i] = box[i].z;
    if (v->node_y1) v->node_y1[i] = box[i].w;
  }
  return NB_OK;
}

int nb_draw(nb_ctx *c, uint8_t *frame, uint32_t render_h) {
  if (!c || !frame || render_h == 0 || render_h > 16384) return NB_ERR_INVALID;
  NB_CUDA(c, cudaSetDevice(c->device));
  const uint32_t world_h = 100000u; // HEIGHT, main.rs:31
  const uint32_t cell = world_h / render_h; // HEIGHT / RENDER_HEIGHT, integer, main.rs:51
  if (cell == 0) return fail(c, NB_ERR_INVALID, "render_h larger than HEIGHT");
  if ((world_h - 1) / cell >= render_h) return fail(c, NB_ERR_INVALID, "HEIGHT/render_h must cover the world");
  const uint32_t npix = render_h * render_h, n = c->n;
  DrawScratch d{};
  uint32_t *buf = nullptr, *slot_of = nullptr;
  uint8_t *dframe = nullptr;
  NB_CUDA(c, dmalloc(&buf, 3 * (size_t)npix));
  NB_CUDA(c, dmalloc(&slot_of, (size_t)std::max(n, 1u)));
  NB_CUDA(c, dmalloc(&dframe, 4 * (size_t)npix));
  d.hits = buf; d.last = buf + npix; d.heavy = buf + 2 * (size_t)npix;
  NB_CUDA(c, cudaMemsetAsync(buf, 0, 3 * (size_t)npix * 4, c->stream));
  { int rv = ensure_velocities(c); if (rv != NB_OK) return rv; }
  BodyArrays &b = c->body[c->cur];
  if (n) {
    slot_of_id_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(b.id, n, slot_of);
    draw_splat_kernel<<<cdiv(n, 256), 256, 0, c->stream>>>(b, n, render_h, (float)world_h, cell, d, c->bvh_ref);
  }
  draw_compose_kernel<<<cdiv(npix, 256), 256, 0, c->stream>>>(d, c->bvh_ref ? nullptr : slot_of, b, npix, dframe);
  NB_CUDA(c, cudaGetLastError());
  NB_CUDA(c, cudaMemcpyAsync(frame, dframe, 4 * (size_t)npix, cudaMemcpyDeviceToHost, c->stream));
  NB_CUDA(c, cudaStreamSynchronize(c->stream));
  cudaFree(buf); cudaFree(slot_of); cudaFree(dframe);
  return NB_OK;
}

// ---- snapshot ---------------------------------------------------------------------------------
int nb_save_snapshot(nb_ctx *c, const char *path, uint64_t steps_done) {
  if (!c || !path) return NB_ERR_INVALID;
  std::vector<nb_particle> buf(c->n);
  int rc = NB_OK;
  if (c->bvh_ref && c->n) {
    // BVHTree::from starts from the Vec order the previous build left (the partition swaps in place and the f32
    // sums are sequential): the snapshot keeps the SLOT order, with the original index of each body in the pad word
    NB_CUDA(c, cudaSetDevice(c->device));
    { int rv = ensure_velocities(c); if (rv != NB_OK) return rv; }
    soa_to_aos_slot_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(c->body[c->cur], c->n, c->aos, true);
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaMemcpyAsync(buf.data(), c->aos, (size_t)c->n * sizeof(nb_particle), cudaMemcpyDeviceToHost, c->stream));
    NB_CUDA(c, cudaStreamSynchronize(c->stream));
  } else
    rc = nb_get_bodies_aos(c, buf.data(), buf.size());
  if (rc != NB_OK) return rc;
  FILE *f = fopen(path, "wb");
  if (!f) return fail(c, NB_ERR_INVALID, std::string("cannot open ") + path);
  const char magic[8] = {'N', 'B', 'B', '2', 'S', 'N', 'A', 'P'};
  const uint32_t ver[2] = {2u, c->bvh_ref ? 1u : 0u}; // flags bit 0: slot order, pad = original index
  const uint64_t hdr[2] = {(uint64_t)c->n, steps_done};
  bool ok = fwrite(magic, 1, 8, f) == 8 && fwrite(ver, 4, 2, f) == 2 && fwrite(hdr, 8, 2, f) == 2 &&
            fwrite(buf.data(), sizeof(nb_particle), buf.size(), f) == buf.size();
  ok = (fclose(f) == 0) && ok;
  return ok ? NB_OK : fail(c, NB_ERR_INVALID, std::string("short write to ") + path);
}

int nb_load_snapshot(nb_ctx *c, const char *path, uint64_t *steps_done) {
  if (!c || !path) return NB_ERR_INVALID;
  FILE *f = fopen(path, "rb");
  if (!f) return fail(c, NB_ERR_INVALID, std::string("cannot open ") + path);
  char magic[8];
  uint32_t ver[2];
  uint64_t hdr[2];
  bool ok = fread(magic, 1, 8, f) == 8 && fread(ver, 4, 2, f) == 2 && fread(hdr, 8, 2, f) == 2 &&
            std::memcmp(magic, "NBB2SNAP", 8) == 0 && (ver[0] == 1u || ver[0] == 2u);
  if (!ok) { fclose(f); return fail(c, NB_ERR_INVALID, std::string("not a snapshot: ") + path); }
  if (hdr[0] > c->cfg.n_max) { fclose(f); return fail(c, NB_ERR_CAPACITY, "snapshot holds more bodies than n_max"); }
  std::vector<nb_particle> buf(hdr[0]);
  ok = fread(buf.data(), sizeof(nb_particle), buf.size(), f) == buf.size();
  fclose(f);
  if (!ok) return fail(c, NB_ERR_INVALID, std::string("truncated snapshot: ") + path);
  if (steps_done) *steps_done = hdr[1];
  int rc = nb_set_bodies_aos(c, buf.data(), buf.size());
  if (rc == NB_OK && ver[0] == 2u && (ver[1] & 1u) && !buf.empty()) { // slot-order snapshot: ids from the pad words
    aos_to_soa_kernel<<<cdiv(c->n, 256), 256, 0, c->stream>>>(c->aos, c->body[c->cur], c->n, true);
    c->aos_fresh = false; // the ids come from the pad words: permute_from_aos_kernel would number the bodies by slot
    NB_CUDA(c, cudaGetLastError());
    NB_CUDA(c, cudaStreamSynchronize(c->stream));
  }
  return rc;
}

// ---- multi-GPU -------------------------------------------------------------------------------
int nb_comm_unique_id(uint8_t id[NB_UNIQUE_ID_BYTES]) {
  std::string err;
  if (!g_nccl.load(err)) return fail(nullptr, NB_ERR_NCCL, err);
  ncclUniqueId u;
  ncclResult_t r = g_nccl.GetUniqueId(&u);
  if (r != ncclSuccess) return fail(nullptr, NB_ERR_NCCL, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
  static_assert(sizeof(ncclUniqueId) <= NB_UNIQUE_ID_BYTES, "unique id size");
  std::memset(id, 0, NB_UNIQUE_ID_BYTES);
  std::memcpy(id, &u, sizeof(u));
  return NB_OK;
}

int nb_comm_init(nb_ctx *c, int rank, int n_ranks, const uint8_t id[NB_UNIQUE_ID_BYTES]) {
  if (!c || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return NB_ERR_INVALID;
  if (n_ranks == 1) { c->rank = 0; c->n_ranks = 1; return NB_OK; }
  if (c->cfg.faithful_index_bug) return fail(c, NB_ERR_INVALID, "faithful_index_bug is single-GPU only");
  std::string err;
  if (!g_nccl.load(err)) return fail(c, NB_ERR_NCCL, err);
  NB_CUDA(c, cudaSetDevice(c->device));
  ncclUniqueId u;
  std::memcpy(&u, id, sizeof(u));
  ncclResult_t r = g_nccl.CommInitRank(&c->comm, n_ranks, u, rank);
  if (r != ncclSuccess) return nccl_fail(c, r, "ncclCommInitRank");
  c->rank = rank;
  c->n_ranks = n_ranks;
  return NB_OK;
}

// Block layout: 8 x cudaIpcMemHandle_t for body[parity].{x,y,vx,vy}, 1 for the flag array, u64 array capacity.
int nb_comm_peer_export(nb_ctx *c, uint8_t block[NB_PEER_BLOCK_BYTES]) {
  if (!c || !block) return NB_ERR_INVALID;
  static_assert(9 * sizeof(cudaIpcMemHandle_t) + 8 <= NB_PEER_BLOCK_BYTES, "peer block size");
  if (c->n_ranks < 2 || c->n_ranks > kMaxPeers) return fail(c, NB_ERR_INVALID, "peer exchange needs 2..8 ranks (nb_comm_init first)");
  NB_CUDA(c, cudaSetDevice(c->device));
  if (!c->peer_flags) {
    NB_CUDA(c, cudaMalloc(&c->peer_flags, kMaxPeers * sizeof(uint32_t)));
    NB_CUDA(c, cudaMemset(c->peer_flags, 0, kMaxPeers * sizeof(uint32_t)));
  }
  std::memset(block, 0, NB_PEER_BLOCK_BYTES);
  cudaIpcMemHandle_t *h = reinterpret_cast<cudaIpcMemHandle_t *>(block);
  for (int p = 0; p < 2; ++p) {
    float *arrs[4] = {c->body[p].x, c->body[p].y, c->body[p].vx, c->body[p].vy};
    for (int a = 0; a < 4; ++a) NB_CUDA(c, cudaIpcGetMemHandle(&h[4 * p + a], arrs[a]));
  }
  NB_CUDA(c, cudaIpcGetMemHandle(&h[8], c->peer_flags));
  const uint64_t cap = c->n_alloc;
  std::memcpy(block + 9 * sizeof(cudaIpcMemHandle_t), &cap, 8);
  return NB_OK;
}

int nb_comm_peer_import(nb_ctx *c, const uint8_t *blocks) {
  if (!c || !blocks) return NB_ERR_INVALID;
  if (c->n_ranks < 2 || c->n_ranks > kMaxPeers || !c->peer_flags)
    return fail(c, NB_ERR_INVALID, "nb_comm_peer_import: call nb_comm_init and nb_comm_peer_export first");
  if (c->peer) return fail(c, NB_ERR_INVALID, "nb_comm_peer_import: already imported");
  NB_CUDA(c, cudaSetDevice(c->device));
  for (int r = 0; r < c->n_ranks; ++r) {
    const uint8_t *blk = blocks + (size_t)r * NB_PEER_BLOCK_BYTES;
    const cudaIpcMemHandle_t *h = reinterpret_cast<const cudaIpcMemHandle_t *>(blk);
    uint64_t cap = 0;
    std::memcpy(&cap, blk + 9 * sizeof(cudaIpcMemHandle_t), 8);
    if (cap != c->n_alloc) return fail(c, NB_ERR_INVALID, "nb_comm_peer_import: ranks were created with different n_max");
    void *ptr[9];
    for (int k = 0; k < 9; ++k) {
      if (r == c->rank) {
        float *own[8] = {c->body[0].x, c->body[0].y, c->body[0].vx, c->body[0].vy,
                         c->body[1].x, c->body[1].y, c->body[1].vx, c->body[1].vy};
        ptr[k] = k < 8 ? (void *)own[k] : (void *)c->peer_flags;
      } else {
        cudaIpcMemHandle_t hh;
        std::memcpy(&hh, &h[k], sizeof(hh));
        NB_CUDA(c, cudaIpcOpenMemHandle(&ptr[k], hh, cudaIpcMemLazyEnablePeerAccess));
        c->peer_mapped.push_back(ptr[k]);
      }
    }
    for (int p = 0; p < 2; ++p) {
      c->peer_out[p].x[r] = (float *)ptr[4 * p + 0];
      c->peer_out[p].y[r] = (float *)ptr[4 * p + 1];
      c->peer_out[p].vx[r] = (float *)ptr[4 * p + 2];
      c->peer_out[p].vy[r] = (float *)ptr[4 * p + 3];
    }
    c->peer_flags_of.flags[r] = (uint32_t *)ptr[8];
  }
  c->peer = true;
  return NB_OK;
}

int nb_comm_peer_enable(nb_ctx *c, int on) {
  if (!c) return NB_ERR_INVALID;
  if (on && c->peer_flags_of.flags[c->n_ranks > 0 ? c->n_ranks - 1 : 0] == nullptr)
    return fail(c, NB_ERR_INVALID, "nb_comm_peer_enable: no imported peer mappings");
  c->peer = on != 0;
  return NB_OK;
}

void nb_shard_range(uint64_t n, int rank, int n_ranks, uint64_t *first, uint64_t *count) {
  if (n_ranks <= 1) {
    *first = 0;
    *count = n;
    return;
  }
  // whole 32-body groups: traverse_group_kernel's sums depend on which 32 bodies share a warp, so shard starts are
  // multiples of 32 and every rank forms the groups a single GPU forms (state bit-identical to one GPU)
  const uint64_t per = ((n + (uint64_t)n_ranks - 1) / (uint64_t)n_ranks + 31u) / 32u * 32u;
  const uint64_t f = std::min<uint64_t>(n, per * (uint64_t)rank);
  *first = f;
  *count = std::min<uint64_t>(per, n - f);
}

int nb_comm_owned_range(nb_ctx *c, uint64_t *first, uint64_t *count) {
  if (!c || !first || !count) return NB_ERR_INVALID;
  uint32_t f, n;
  owned_range(c, f, n);
  *first = f;
  *count = n;
  return NB_OK;
}

} // extern "C"

// ======================================================================================
// Allocation-free direct sum (no counterpart in the reference).
// nb_accel_direct cudaMallocs its fp64 accumulators on every call; callers that must not allocate (CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc) get the same tiled O(N^2) sum with the accumulators in a
// file-scope device array instead.  The array is held by every GPU that loads the library until the process exits.
// ======================================================================================
namespace {

constexpr uint32_t kNoAllocMaxTargets = NB_NOALLOC_MAX_TARGETS; // 2 doubles each: 1 MiB
static __device__ __align__(256) double g_noalloc_acc[2 * (size_t)kNoAllocMaxTargets];

// In host code `g_noalloc_acc` is a host-side stand-in, NOT a device address: the kernels get the address
// cudaGetSymbolAddress returns, looked up once per device with that device current.
constexpr int kNoAllocMaxDevices = 64;
double *g_noalloc_ptr[kNoAllocMaxDevices] = {};
std::mutex g_noalloc_mutex;

// direct_kernel (direct.cuh) with the targets named by position instead of through a slot table: target t is body
// first + t of the arrays passed in.  Same tiles, same fp32 pair terms and per-tile partial sums, same fp64
// accumulation across tiles and source chunks.
template <bool EXACT>
__global__ void __launch_bounds__(kDirectThreads) direct_noalloc_kernel(const float *__restrict__ x,
                                                                        const float *__restrict__ y,
                                                                        const uint32_t *__restrict__ m, uint32_t n,
                                                                        uint32_t first, uint32_t nt,
                                                                        double *__restrict__ acc, uint32_t chunk) {
  __shared__ float4 tile[kDirectThreads];
  const uint32_t t = blockIdx.x * kDirectThreads + threadIdx.x;
  const bool valid = t < nt;
  const uint32_t slot = first + (valid ? t : 0u);
  const float px = x[slot], py = y[slot];
  const uint32_t s0 = blockIdx.y * chunk;
  const uint32_t s1 = (s0 + chunk < n) ? s0 + chunk : n;
  double dax = 0.0, day = 0.0;
  for (uint32_t base = s0; base < s1; base += kDirectThreads) {
    const uint32_t j = base + threadIdx.x;
    tile[threadIdx.x] = (j < s1) ? make_float4(x[j], y[j], (float)m[j], 0.f) : make_float4(px, py, 0.f, 0.f);
    __syncthreads();
    const uint32_t cnt = (s1 - base < (uint32_t)kDirectThreads) ? s1 - base : (uint32_t)kDirectThreads;
    float fax = 0.f, fay = 0.f;
    if (cnt == (uint32_t)kDirectThreads) {
#pragma unroll 8
      for (int q = 0; q < kDirectThreads; ++q) {
        const float4 b = tile[q];
        pair_force<EXACT>(px, py, b.x, b.y, b.z, fax, fay);
      }
    } else {
      for (uint32_t q = 0; q < cnt; ++q) {
        const float4 b = tile[q];
        pair_force<EXACT>(px, py, b.x, b.y, b.z, fax, fay);
      }
    }
    dax += (double)fax;
    day += (double)fay;
    __syncthreads();
  }
  if (valid) {
    atomicAdd(&acc[2 * (size_t)t], dax);
    atomicAdd(&acc[2 * (size_t)t + 1], day);
  }
}

} // namespace

extern "C" int nb_accel_direct_noalloc(const float *d_x, const float *d_y, const uint32_t *d_weight, size_t n,
                                       size_t first, size_t count, int32_t force_mode, float *d_ax, float *d_ay,
                                       void *stream) {
  if (count == 0) return NB_OK;
  if (!d_x || !d_y || !d_weight || !d_ax || !d_ay) return fail(nullptr, NB_ERR_INVALID, "nb_accel_direct_noalloc: null pointer");
  if (force_mode != NB_FORCE_EXACT && force_mode != NB_FORCE_FAST) return fail(nullptr, NB_ERR_INVALID, "nb_accel_direct_noalloc: force_mode");
  if (count > kNoAllocMaxTargets)
    return fail(nullptr, NB_ERR_CAPACITY,
                "nb_accel_direct_noalloc: " + std::to_string(count) + " targets exceed NB_NOALLOC_MAX_TARGETS = " +
                    std::to_string(kNoAllocMaxTargets) + " (the static scratch array); split the call or use nb_accel_direct");
  if (n > 0xffffffffull || first + count > n) return fail(nullptr, NB_ERR_CAPACITY, "nb_accel_direct_noalloc: targets outside [0, n) or n >= 2^32");
  int dev = 0;
  NB_CUDA(nullptr, cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(nullptr, NB_ERR_INVALID, "nb_accel_direct_noalloc: device ordinal out of range");
  double *acc = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) { // set-up, once per device, with that device current
      double *p = nullptr;
      NB_CUDA(nullptr, cudaGetSymbolAddress((void **)&p, g_noalloc_acc));
      g_noalloc_ptr[dev] = p;
    }
    acc = g_noalloc_ptr[dev];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const uint32_t un = (uint32_t)n, nt = (uint32_t)count;
  // the array keeps the last call's sums: the accumulators of this call's targets start at zero
  NB_CUDA(nullptr, cudaMemsetAsync(acc, 0, 2 * (size_t)nt * sizeof(double), st));
  const uint32_t tblocks = cdiv(nt, kDirectThreads);
  uint32_t ychunks = std::max(1u, std::min(cdiv(un, 4096u), cdiv((uint32_t)kSMs * 8u, tblocks)));
  ychunks = std::min(ychunks, 65535u);
  const uint32_t chunk = cdiv(cdiv(un, ychunks), kDirectThreads) * kDirectThreads;
  ychunks = cdiv(un, chunk);
  const dim3 grid(tblocks, ychunks);
  if (force_mode == NB_FORCE_EXACT)
    direct_noalloc_kernel<true><<<grid, kDirectThreads, 0, st>>>(d_x, d_y, d_weight, un, (uint32_t)first, nt, acc, chunk);
  else
    direct_noalloc_kernel<false><<<grid, kDirectThreads, 0, st>>>(d_x, d_y, d_weight, un, (uint32_t)first, nt, acc, chunk);
  direct_finish_kernel<<<cdiv(nt, 256), 256, 0, st>>>(acc, nt, d_ax, d_ay);
  g_launches += 2;
  NB_CUDA(nullptr, cudaGetLastError());
  return NB_OK;
}
