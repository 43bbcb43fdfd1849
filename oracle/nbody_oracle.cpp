// nbody_oracle.cpp — CPU ORACLE.  TEST INFRASTRUCTURE ONLY.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (libnbody_b200.so) never links, loads or calls it.
//
// What it is: a C++ restatement of the per-step hot path of KristinnVikarJ/nbody-simulation,
// literal in operation order, compiled with -ffp-contract=off (Rust never contracts a*b+c to
// FMA) and IEEE divides.  Every function cites the reference lines it follows
// (paths relative to /root/reference).
//
// PARITY STATUS: **parity unpinned**.  The reference ships no tests, fixtures or golden vectors
// (SURVEY.md §4), cannot be built here (no cargo/rustc, 232 un-vendored crates, needs a
// window), and is Rust, so neither oracle/_ref nor imported fixtures exist.  The restatement is
// pinned only by hand-derived known-answer tests (tests/test_oracle_kat.py) and by the
// cross-check "literal quad_tree.rs insert-built tree == Morton-pipeline tree" below.
// Third-party arithmetic relied on from memory (crates absent from /root/reference):
//   pathfinder_geometry 0.5.1 / pathfinder_simd 0.5.2 (Cargo.lock:1136-1147): lane-wise IEEE
//     f32 (+,-,*,/ as addps/subps/mulps/divps; min/max as minps/maxps; square_length =
//     x*x + y*y); Vector2F / f32 == lane-wise divide by splat.
//   partition 0.1.2 (Cargo.lock:1130-1131): in-place two-pointer (Hoare style) partition,
//     returns (predicate-true side, predicate-false side).
//   rayon 1.7.0: ordered collect => per-body results independent of thread count.
//
// Two families live here:
//   (A) LITERAL restatements: bvh_tree.rs (live path), quad_tree.rs (orphan; traversal carried
//       over from main.rs:349-386), World::update, pair force.
//   (B) PIPELINE restatement: the Morton-key -> stable sort -> tree -> up-sweep -> stackless
//       traversal formulation the GPU implements, written independently (recursive, serial)
//       so GPU output can be compared bit for bit.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <limits>
#include <memory>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/nbody_b200.h" // struct layouts only (nb_tree_view, nb_counting)

namespace {

// ---------------------------------------------------------------------------------------
// Vector2F (pathfinder_geometry) — lane-wise f32, SSE min/max semantics.
// ---------------------------------------------------------------------------------------
struct V2 {
  float x, y;
};
inline V2 operator+(V2 a, V2 b) { return {a.x + b.x, a.y + b.y}; }
inline V2 operator-(V2 a, V2 b) { return {a.x - b.x, a.y - b.y}; }
inline V2 operator*(V2 a, float s) { return {a.x * s, a.y * s}; }
inline V2 operator/(V2 a, float s) { return {a.x / s, a.y / s}; }
inline V2 vmin(V2 a, V2 b) { return {a.x < b.x ? a.x : b.x, a.y < b.y ? a.y : b.y}; } // minps
inline V2 vmax(V2 a, V2 b) { return {a.x > b.x ? a.x : b.x, a.y > b.y ? a.y : b.y}; } // maxps
inline float square_length(V2 a) { return a.x * a.x + a.y * a.y; }

// src/main.rs:228-232
inline float dist2(V2 pos1, V2 pos2) { return square_length(pos1 - pos2); }

// src/main.rs:234-253
inline void calculate_gravity(V2 particle1, V2 particle2, V2 &accel, float force) {
  V2 diff = particle2 - particle1;
  float sum = std::fabs(diff.x) + std::fabs(diff.y);
  if (!std::isnormal(sum)) return; // :241
  float distance = square_length(diff);
  if (distance < 0.001f) distance = 0.001f; // :247-249
  accel = accel + (diff * force) / (sum * distance); // :252
}

// src/main.rs:193-198 (+ id: bookkeeping only, never enters arithmetic)
struct Particle {
  V2 position, velocity;
  uint32_t weight;
  uint32_t id;
};

// ---------------------------------------------------------------------------------------
// (A1) LITERAL src/bvh_tree.rs
// ---------------------------------------------------------------------------------------
struct RectB { // bvh_tree.rs:8-21
  V2 offset, size;
  bool contains(V2 o) const {
    return o.y > offset.y && o.x > offset.x && o.x < (offset.x + size.x) &&
           o.y < (offset.y + size.y);
  }
};

struct BVH { // bvh_tree.rs:23-35
  bool is_leaf;
  RectB boundary;
  uint32_t total_mass;
  V2 cog;
  BVH *child[2];
  const Particle *pts;
  size_t len;
};

constexpr size_t TARGET_POINTS = 64; // bvh_tree.rs:37

struct BVHArena {
  std::deque<BVH> pool;
  BVH *alloc() {
    pool.emplace_back();
    return &pool.back();
  }
  bool overflow = false;
};

// partition 0.1.2 (from memory): left = predicate true.
template <class P> size_t partition_crate(Particle *d, size_t len, P pred) {
  if (len == 0) return 0;
  size_t l = 0, r = len - 1;
  for (;;) {
    while (l < len && pred(d[l])) ++l;
    while (r > 0 && !pred(d[r])) --r;
    if (l >= r) return l;
    std::swap(d[l], d[r]);
  }
}

// bvh_tree.rs:40-54
BVH *bvh_make_leaf(BVHArena &A, const Particle *pts, size_t len) {
  V2 mn{std::numeric_limits<float>::max(), std::numeric_limits<float>::max()}, mx{0.f, 0.f};
  for (size_t i = 0; i < len; ++i) {
    mn = vmin(mn, pts[i].position);
    mx = vmax(mx, pts[i].position);
  }
  BVH *n = A.alloc();
  n->is_leaf = true;
  n->boundary = {mn, mx - mn};
  n->total_mass = 0;
  n->cog = {0.f, 0.f};
  n->child[0] = n->child[1] = nullptr;
  n->pts = pts;
  n->len = len;
  return n;
}

// bvh_tree.rs:57-96
BVH *bvh_from(BVHArena &A, Particle *pts, size_t len, int depth) {
  if (depth > 4000) { // the reference would overflow its stack (all-coincident input)
    A.overflow = true;
    return bvh_make_leaf(A, pts, len);
  }
  V2 mn{std::numeric_limits<float>::max(), std::numeric_limits<float>::max()}, mx{0.f, 0.f},
      sum{0.f, 0.f};
  for (size_t i = 0; i < len; ++i) { // :58-61
    mn = vmin(mn, pts[i].position);
    mx = vmax(mx, pts[i].position);
    sum = sum + pts[i].position;
  }
  RectB bounds{mn, mx - mn};
  V2 halved = sum / (float)len; // :67
  size_t half_len = len / 2;
  size_t cx = 0, cy = 0;
  for (size_t i = 0; i < len; ++i) {
    cx += pts[i].position.x > halved.x;
    cy += pts[i].position.y > halved.y;
  }
  size_t hori = half_len > cx ? half_len - cx : cx - half_len; // :71
  size_t vert = half_len > cy ? half_len - cy : cy - half_len; // :72
  size_t split;
  if (vert > hori) // :73-77
    split = partition_crate(pts, len, [&](const Particle &p) { return p.position.x > halved.x; });
  else
    split = partition_crate(pts, len, [&](const Particle &p) { return p.position.y > halved.y; });
  Particle *left = pts;
  size_t llen = split;
  Particle *right = pts + split;
  size_t rlen = len - split;
  BVH *l = llen > TARGET_POINTS ? bvh_from(A, left, llen, depth + 1) : bvh_make_leaf(A, left, llen);
  BVH *r = rlen > TARGET_POINTS ? bvh_from(A, right, rlen, depth + 1) : bvh_make_leaf(A, right, rlen);
  BVH *n = A.alloc();
  n->is_leaf = false;
  n->boundary = bounds;
  n->total_mass = 0;
  n->cog = {0.f, 0.f};
  n->child[0] = l;
  n->child[1] = r;
  n->pts = pts;
  n->len = len;
  return n;
}

// bvh_tree.rs:98-116
V2 bvh_get_cog(const BVH *n) {
  if (n->is_leaf) {
    V2 acc{0.f, 0.f};
    for (size_t i = 0; i < n->len; ++i) acc = acc + n->pts[i].position;
    return acc / (float)n->len;
  }
  return n->cog;
}
// bvh_tree.rs:118-131
uint32_t bvh_get_mass(const BVH *n) {
  if (n->is_leaf) {
    uint32_t a = 0;
    for (size_t i = 0; i < n->len; ++i) a += n->pts[i].weight;
    return a;
  }
  return n->total_mass;
}
// bvh_tree.rs:133-158
void bvh_calculate_gravity(BVH *n) {
  if (n->is_leaf) return;
  bvh_calculate_gravity(n->child[0]);
  bvh_calculate_gravity(n->child[1]);
  uint32_t m0 = bvh_get_mass(n->child[0]), m1 = bvh_get_mass(n->child[1]);
  uint32_t mass = m0 + m1;
  V2 big = (bvh_get_cog(n->child[0]) * (float)m0) + (bvh_get_cog(n->child[1]) * (float)m1);
  n->cog = big / (float)mass;
  n->total_mass = mass;
}

struct Counters {
  uint64_t interactions = 0, tests = 0;
};

// src/main.rs:349-386
void bvh_sum_gravity(V2 p, const BVH *t, V2 &accel, float theta, Counters &c) {
  if (t->is_leaf) {
    for (size_t i = 0; i < t->len; ++i)
      calculate_gravity(p, t->pts[i].position, accel, (float)t->pts[i].weight);
    c.interactions += t->len;
  } else {
    c.tests++;
    V2 sz = t->boundary.size;
    V2 tmp = vmax(sz, V2{sz.y, sz.x}); // :371
    if (!t->boundary.contains(p) && tmp.x * tmp.y < dist2(p, t->cog) * theta * theta) {
      calculate_gravity(p, t->cog, accel, (float)t->total_mass);
      c.interactions++;
    } else {
      bvh_sum_gravity(p, t->child[0], accel, theta, c);
      bvh_sum_gravity(p, t->child[1], accel, theta, c);
    }
  }
}

// ---------------------------------------------------------------------------------------
// (A2) LITERAL src/quad_tree.rs (SmallParticle carries its weight: the field is commented out
// at main.rs:201-204 but read at quad_tree.rs:144 — the evident intent is restored).
// ---------------------------------------------------------------------------------------
struct SmallParticle {
  V2 position;
  uint32_t weight;
  uint32_t id;
};
struct RectQ { // quad_tree.rs:7-33
  V2 offset;
  float height, height2;
  static RectQ make(V2 off, float h) { return {off, h, h * h}; }
  bool contains(V2 o) const {
    return o.y > offset.y && o.x > offset.x && o.x < (offset.x + height) &&
           o.y < (offset.y + height);
  }
  V2 off(float x, float y) const { return {offset.x + x, offset.y + y}; }
};
struct Quad { // quad_tree.rs:35-51
  RectQ boundary;
  V2 cog;
  bool is_leaf;
  // Leaf
  uint8_t count;
  SmallParticle slots[8];
  bool present[8];
  // Root
  uint8_t flags;
  uint32_t total_mass;
  Quad *children[4];
};
struct QuadArena {
  std::deque<Quad> pool;
  bool overflow = false;
  Quad *make(RectQ b) { // quad_tree.rs:55-64
    pool.emplace_back();
    Quad *q = &pool.back();
    q->boundary = b;
    q->cog = {0.f, 0.f};
    q->is_leaf = true;
    q->count = 0;
    for (int i = 0; i < 8; ++i) q->present[i] = false;
    q->flags = 0;
    q->total_mass = 0;
    for (int i = 0; i < 4; ++i) q->children[i] = nullptr;
    return q;
  }
};
constexpr uint8_t MAX_CAPACITY = 8; // quad_tree.rs:54

void quad_insert(QuadArena &A, Quad *q, const SmallParticle &pt, int depth);

// quad_tree.rs:209-227
void quad_subdivide(QuadArena &A, Quad *q, int depth) {
  SmallParticle cur[8];
  int n = q->count;
  for (int i = 0; i < n; ++i) cur[i] = q->slots[i];
  q->is_leaf = false;
  q->flags = 0;
  for (int i = 0; i < 4; ++i) q->children[i] = nullptr;
  q->total_mass = (uint32_t)n;
  q->cog = {0.f, 0.f};
  q->count = 0;
  for (int i = 0; i < n; ++i) quad_insert(A, q, cur[i], depth);
}

// quad_tree.rs:153-207
void quad_insert(QuadArena &A, Quad *q, const SmallParticle &pt, int depth) {
  if (depth > 200) { // the reference recurses forever on > 8 coincident bodies
    A.overflow = true;
    return;
  }
  if (q->is_leaf) {
    if (q->count == MAX_CAPACITY) {
      quad_subdivide(A, q, depth);
      quad_insert(A, q, pt, depth);
    } else {
      q->slots[q->count] = pt;
      q->present[q->count] = true;
      q->count++;
    }
    return;
  }
  float half_height = q->boundary.height / 2.0f; // :172
  float hori_half = q->boundary.offset.x + half_height;
  float vert_half = q->boundary.offset.y + half_height;
  bool north = pt.position.y > vert_half;
  bool west = pt.position.x > hori_half;
  int child = ((int)north << 1) + (int)west; // :179
  if ((q->flags & (1 << child)) == 0) {
    V2 off;
    switch (child) { // :182-188
    case 0: off = q->boundary.offset; break;
    case 1: off = q->boundary.off(half_height, 0.0f); break;
    case 2: off = q->boundary.off(0.0f, half_height); break;
    default: off = q->boundary.off(half_height, half_height); break;
    }
    q->children[child] = A.make(RectQ::make(off, half_height));
    q->flags |= (uint8_t)(1 << child);
  }
  quad_insert(A, q->children[child], pt, depth + 1);
}

// quad_tree.rs:139-151
uint32_t quad_get_total_mass(const Quad *q) {
  if (q->is_leaf) {
    uint32_t a = 0;
    for (int i = 0; i < 8; ++i)
      if (q->present[i]) a += q->slots[i].weight;
    return a;
  }
  return q->total_mass;
}

// quad_tree.rs:229-270
void quad_calculate_gravity(Quad *q) {
  if (q->is_leaf) {
    if (q->count > 0) {
      V2 acc{0.f, 0.f};
      for (int i = 0; i < 8; ++i)
        if (q->present[i]) acc = acc + q->slots[i].position;
      q->cog = acc / (float)q->count;
    }
    return;
  }
  for (int i = 0; i < 4; ++i)
    if (q->children[i]) quad_calculate_gravity(q->children[i]);
  uint32_t mass = 0;
  for (int i = 0; i < 4; ++i)
    if (q->children[i]) mass += quad_get_total_mass(q->children[i]);
  V2 big{0.f, 0.f};
  for (int i = 0; i < 4; ++i)
    if (q->children[i])
      big = big + (q->children[i]->cog * (float)quad_get_total_mass(q->children[i]));
  q->cog = big / (float)mass;
  q->total_mass = mass;
}

// quad_tree.rs:66-89
uint32_t quad_empty(Quad *q) {
  if (q->is_leaf) {
    q->count = 0;
    for (int i = 0; i < 8; ++i) q->present[i] = false;
    return 1;
  }
  if (q->total_mass == 0) return 1;
  q->total_mass = 0;
  uint32_t sum = 0;
  for (int i = 0; i < 4; ++i)
    if (q->children[i]) sum += quad_empty(q->children[i]);
  return sum + 1;
}

// quad_tree.rs:94-137
uint32_t quad_prune(Quad *q) {
  uint32_t sum = 0;
  if (q->is_leaf) return 0;
  for (int i = 0; i < 4; ++i) {
    bool child_empty = false;
    Quad *c = q->children[i];
    if (c) {
      if (c->is_leaf) {
        if (c->count == 0) child_empty = true;
      } else {
        if (c->total_mass == 0) child_empty = true;
        else sum += quad_prune(c);
      }
    }
    if (child_empty) {
      q->children[i] = nullptr;
      q->flags ^= (uint8_t)(1 << i);
      sum += 1;
    }
  }
  return sum;
}

// Traversal carried over from main.rs:349-386 (quad_tree.rs has none): leaves summed directly,
// Root accepted iff !contains(p) && max(w,h)^2 < dist2*theta*theta with size = (height,height),
// children visited in index order 0..3.
void quad_sum_gravity(V2 p, const Quad *t, V2 &accel, float theta, Counters &c) {
  if (t->is_leaf) {
    for (int i = 0; i < 8; ++i)
      if (t->present[i])
        calculate_gravity(p, t->slots[i].position, accel, (float)t->slots[i].weight);
    c.interactions += t->count;
  } else {
    c.tests++;
    float h = t->boundary.height;
    if (!t->boundary.contains(p) && h * h < dist2(p, t->cog) * theta * theta) {
      calculate_gravity(p, t->cog, accel, (float)t->total_mass);
      c.interactions++;
    } else {
      for (int i = 0; i < 4; ++i)
        if (t->children[i]) quad_sum_gravity(p, t->children[i], accel, theta, c);
    }
  }
}

// ---------------------------------------------------------------------------------------
// Flat pre-order tree (what nb_get_tree exports)
// ---------------------------------------------------------------------------------------
struct FlatTree {
  size_t n = 0, n_in = 0;
  std::vector<uint64_t> keys;   // original order
  std::vector<uint32_t> order;  // tree order -> original id
  std::vector<int32_t> depth;
  std::vector<uint32_t> start, count, skip, parent, mass;
  std::vector<uint8_t> leaf;
  std::vector<float> cogx, cogy, x0, y0, x1, y1, s2;
  std::vector<float> sx, sy, sm; // bodies in tree order (mass as f32)
  uint32_t overflow_leaves = 0;
  float rx0 = 0, ry0 = 0, rh = 0;
  uint32_t push(int32_t d, uint32_t st, uint32_t cnt, uint32_t par) {
    uint32_t i = (uint32_t)depth.size();
    depth.push_back(d);
    start.push_back(st);
    count.push_back(cnt);
    skip.push_back(0);
    parent.push_back(par);
    mass.push_back(0);
    leaf.push_back(0);
    cogx.push_back(0.f);
    cogy.push_back(0.f);
    x0.push_back(0.f);
    y0.push_back(0.f);
    x1.push_back(0.f);
    y1.push_back(0.f);
    s2.push_back(0.f);
    return i;
  }
};

// Stackless pre-order walk; per body identical in visiting order and arithmetic to the
// recursive main.rs:349-386 (child 0 first; leaf bodies in leaf order).
inline void flat_sum_gravity(const FlatTree &t, V2 p, float theta, V2 &accel, Counters &c) {
  const uint32_t nn = (uint32_t)t.depth.size();
  uint32_t cur = 0;
  while (cur < nn) {
    if (t.leaf[cur]) {
      uint32_t s = t.start[cur], e = s + t.count[cur];
      for (uint32_t j = s; j < e; ++j) calculate_gravity(p, V2{t.sx[j], t.sy[j]}, accel, t.sm[j]);
      c.interactions += t.count[cur];
      cur = t.skip[cur];
    } else {
      c.tests++;
      bool contains = p.y > t.y0[cur] && p.x > t.x0[cur] && p.x < t.x1[cur] && p.y < t.y1[cur];
      V2 cog{t.cogx[cur], t.cogy[cur]};
      if (!contains && t.s2[cur] < dist2(p, cog) * theta * theta) {
        calculate_gravity(p, cog, accel, (float)t.mass[cur]);
        c.interactions++;
        cur = t.skip[cur];
      } else
        cur = cur + 1;
    }
  }
}

// ---------------------------------------------------------------------------------------
// (B) PIPELINE restatement
// ---------------------------------------------------------------------------------------
constexpr int KEY_LEVELS = 24;                       // quadtree levels encoded (2 bits each)
constexpr uint64_t KEY_OUT = (uint64_t)1 << 48;      // body outside the root square (quad)

// In-domain test: the caller-side filter the reference implies (within_bounds, main.rs:223-226).
inline bool in_domain(float px, float py, float x0, float y0, float h) {
  return py < (y0 + h) && px < (x0 + h) && py >= y0 && px >= x0;
}

// Morton key = the child indices quad_tree.rs:172-194 would take, level by level, using the
// same f32 midline chain (half = height/2; mid = offset + half; bit = coord > mid;
// child offset = mid or offset).  y bit above x bit (quad_tree.rs:179).
inline uint64_t morton_key(float px, float py, float x0, float y0, float h) {
  float ox = x0, oy = y0, hh = h;
  uint64_t key = 0;
  for (int l = 0; l < KEY_LEVELS; ++l) {
    hh = hh / 2.0f;
    float mx = ox + hh, my = oy + hh;
    bool bx = px > mx, by = py > my;
    key = (key << 2) | ((uint64_t)by << 1) | (uint64_t)bx;
    if (bx) ox = mx;
    if (by) oy = my;
  }
  return key;
}

inline int digit_at(uint64_t key, int level /*1-based*/) {
  return (int)((key >> (2 * (KEY_LEVELS - level))) & 3);
}

void fill_sorted_bodies(FlatTree &t, const float *x, const float *y, const uint32_t *w) {
  t.sx.resize(t.n);
  t.sy.resize(t.n);
  t.sm.resize(t.n);
  for (size_t i = 0; i < t.n; ++i) {
    uint32_t id = t.order[i];
    t.sx[i] = x[id];
    t.sy[i] = y[id];
    t.sm[i] = (float)w[id];
  }
}

void pipe_quad_rec(FlatTree &t, size_t lo, size_t hi, int depth, float ox, float oy, float h,
                   uint32_t parent, int cap) {
  uint32_t me = t.push(depth, (uint32_t)lo, (uint32_t)(hi - lo), parent);
  t.x0[me] = ox;
  t.y0[me] = oy;
  t.x1[me] = ox + h;
  t.y1[me] = oy + h;
  t.s2[me] = h * h;
  if (hi - lo <= (size_t)cap || depth == KEY_LEVELS) {
    t.leaf[me] = 1;
    if (hi - lo > (size_t)cap) t.overflow_leaves++;
    std::sort(t.order.begin() + lo, t.order.begin() + hi); // insertion order == ascending id
  } else {
    float half = h / 2.0f;
    float mx = ox + half, my = oy + half;
    size_t p = lo;
    for (int q = 0; q < 4; ++q) {
      size_t e = p;
      while (e < hi && digit_at(t.keys[t.order[e]], depth + 1) == q) ++e;
      if (e > p) pipe_quad_rec(t, p, e, depth + 1, (q & 1) ? mx : ox, (q & 2) ? my : oy, half, me, cap);
      p = e;
    }
  }
  t.skip[me] = (uint32_t)t.depth.size();
}

// Up-sweep over the flat tree in reverse pre-order (children before parents).
// kind 0 (quad): quad_tree.rs:229-270.  kind 1 (bvh): bvh_tree.rs:98-158 + tight bounds :40-66.
void flat_upsweep(FlatTree &t, const uint32_t *w, int kind) {
  const uint32_t nn = (uint32_t)t.depth.size();
  for (uint32_t ii = nn; ii-- > 0;) {
    uint32_t s = t.start[ii], e = s + t.count[ii];
    if (kind == 1) { // tight AABB: min init f32::MAX, max init 0 (bvh_tree.rs:42,59)
      V2 mn{std::numeric_limits<float>::max(), std::numeric_limits<float>::max()}, mx{0.f, 0.f};
      for (uint32_t j = s; j < e; ++j) {
        V2 p{t.sx[j], t.sy[j]};
        mn = vmin(mn, p);
        mx = vmax(mx, p);
      }
      V2 size = mx - mn;
      t.x0[ii] = mn.x;
      t.y0[ii] = mn.y;
      t.x1[ii] = mn.x + size.x;
      t.y1[ii] = mn.y + size.y;
      V2 tmp = vmax(size, V2{size.y, size.x});
      t.s2[ii] = tmp.x * tmp.y;
    }
    if (t.leaf[ii]) {
      V2 acc{0.f, 0.f};
      uint32_t m = 0;
      for (uint32_t j = s; j < e; ++j) {
        acc = acc + V2{t.sx[j], t.sy[j]};
        m += w[t.order[j]];
      }
      V2 c = acc / (float)t.count[ii];
      t.cogx[ii] = c.x;
      t.cogy[ii] = c.y;
      t.mass[ii] = m;
    } else {
      uint32_t m = 0;
      V2 big{0.f, 0.f};
      bool first = true;
      for (uint32_t c = ii + 1; c < t.skip[ii]; c = t.skip[c]) {
        m += t.mass[c];
        V2 term = V2{t.cogx[c], t.cogy[c]} * (float)t.mass[c];
        if (kind == 1 && first) big = term; // bvh_tree.rs:150-152: a*m0 + b*m1, no zero seed
        else big = big + term;              // quad_tree.rs:259-264: fold from (0,0)
        first = false;
      }
      V2 c = big / (float)m;
      t.cogx[ii] = c.x;
      t.cogy[ii] = c.y;
      t.mass[ii] = m;
    }
  }
}

// LBVH: binary radix tree over the 48-bit keys (bit 47 first), bucket leaves of <= cap bodies.
void pipe_lbvh_rec(FlatTree &t, size_t lo, size_t hi, uint32_t parent, int cap) {
  uint64_t kf = t.keys[t.order[lo]], kl = t.keys[t.order[hi - 1]];
  uint64_t diff = kf ^ kl;
  int pre = diff ? (__builtin_clzll(diff) - 16) : 48; // common prefix length in bits
  bool leaf = (hi - lo <= (size_t)cap) || diff == 0;
  uint32_t me = t.push(pre, (uint32_t)lo, (uint32_t)(hi - lo), parent);
  if (leaf) {
    t.leaf[me] = 1;
    if (hi - lo > (size_t)cap) t.overflow_leaves++;
    std::sort(t.order.begin() + lo, t.order.begin() + hi);
  } else {
    int bit = 47 - pre; // first differing bit
    size_t a = lo, b = hi; // first index whose `bit` is set
    while (a < b) {
      size_t m = (a + b) / 2;
      if ((t.keys[t.order[m]] >> bit) & 1) b = m;
      else a = m + 1;
    }
    pipe_lbvh_rec(t, lo, a, me, cap);
    pipe_lbvh_rec(t, a, hi, me, cap);
  }
  t.skip[me] = (uint32_t)t.depth.size();
}

struct OrcTree {
  int kind; // 0 pipe-quad, 1 pipe-lbvh, 2 literal quad_tree.rs, 3 literal bvh_tree.rs
  FlatTree flat;
  std::vector<float> x, y;
  std::vector<uint32_t> w;
  // literal structures kept alive for the recursive traversals
  std::unique_ptr<QuadArena> qa;
  Quad *qroot = nullptr;
  std::unique_ptr<BVHArena> ba;
  BVH *broot = nullptr;
  std::vector<Particle> particles;
  bool overflow = false;
};

void flatten_quad(const Quad *q, int depth, uint32_t parent, FlatTree &t) {
  uint32_t me = t.push(depth, (uint32_t)t.order.size(), 0, parent);
  t.x0[me] = q->boundary.offset.x;
  t.y0[me] = q->boundary.offset.y;
  t.x1[me] = q->boundary.offset.x + q->boundary.height;
  t.y1[me] = q->boundary.offset.y + q->boundary.height;
  t.s2[me] = q->boundary.height * q->boundary.height;
  t.cogx[me] = q->cog.x;
  t.cogy[me] = q->cog.y;
  t.mass[me] = quad_get_total_mass(q);
  if (q->is_leaf) {
    t.leaf[me] = 1;
    for (int i = 0; i < 8; ++i)
      if (q->present[i]) t.order.push_back(q->slots[i].id);
  } else {
    for (int i = 0; i < 4; ++i)
      if (q->children[i]) flatten_quad(q->children[i], depth + 1, me, t);
  }
  t.count[me] = (uint32_t)t.order.size() - t.start[me];
  t.skip[me] = (uint32_t)t.depth.size();
}

void flatten_bvh(const BVH *b, int depth, uint32_t parent, FlatTree &t, const Particle *base) {
  uint32_t me = t.push(depth, (uint32_t)(b->pts - base), (uint32_t)b->len, parent);
  t.x0[me] = b->boundary.offset.x;
  t.y0[me] = b->boundary.offset.y;
  t.x1[me] = b->boundary.offset.x + b->boundary.size.x;
  t.y1[me] = b->boundary.offset.y + b->boundary.size.y;
  V2 sz = b->boundary.size;
  V2 tmp = vmax(sz, V2{sz.y, sz.x});
  t.s2[me] = tmp.x * tmp.y;
  V2 c = bvh_get_cog(b);
  t.cogx[me] = c.x;
  t.cogy[me] = c.y;
  t.mass[me] = bvh_get_mass(b);
  if (b->is_leaf) t.leaf[me] = 1;
  else {
    flatten_bvh(b->child[0], depth + 1, me, t, base);
    flatten_bvh(b->child[1], depth + 1, me, t, base);
  }
  t.skip[me] = (uint32_t)t.depth.size();
}

double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

} // namespace

extern "C" {

// ---- known-answer-test hooks --------------------------------------------------------------
void orc_pair_force(float p1x, float p1y, float p2x, float p2y, float m, float *ax, float *ay) {
  V2 a{*ax, *ay};
  calculate_gravity(V2{p1x, p1y}, V2{p2x, p2y}, a, m);
  *ax = a.x;
  *ay = a.y;
}
// main.rs:419-423
void orc_integrate(float *x, float *y, float *vx, float *vy, const float *ax, const float *ay,
                   size_t n, float delta) {
  for (size_t i = 0; i < n; ++i) {
    V2 v = V2{vx[i], vy[i]} + V2{ax[i], ay[i]} * delta;
    V2 velo = v * delta;
    V2 p = V2{x[i], y[i]} + velo;
    vx[i] = v.x;
    vy[i] = v.y;
    x[i] = p.x;
    y[i] = p.y;
  }
}
// main.rs:370-372 with a bvh_tree.rs Rectangle(offset,size)
int orc_opening_accept(float px, float py, float ox, float oy, float w, float h, float cx, float cy,
                       float theta) {
  RectB r{{ox, oy}, {w, h}};
  V2 tmp = vmax(r.size, V2{r.size.y, r.size.x});
  return (!r.contains(V2{px, py}) && tmp.x * tmp.y < dist2(V2{px, py}, V2{cx, cy}) * theta * theta) ? 1 : 0;
}
uint64_t orc_morton_key(float px, float py, float x0, float y0, float h) {
  return in_domain(px, py, x0, y0, h) ? morton_key(px, py, x0, y0, h) : KEY_OUT;
}

// ---- trees ----------------------------------------------------------------------------------
// kind: 0 pipeline quadtree, 1 pipeline LBVH, 2 literal quad_tree.rs, 3 literal bvh_tree.rs
void *orc_tree_build(int kind, const float *x, const float *y, const uint32_t *w, size_t n, float x0,
                     float y0, float h, int cap, int domain_auto) {
  OrcTree *T = new OrcTree();
  T->kind = kind;
  T->x.assign(x, x + n);
  T->y.assign(y, y + n);
  T->w.assign(w, w + n);
  FlatTree &t = T->flat;
  t.n = n;
  if (h <= 0.f) { x0 = 0.f; y0 = 0.f; h = 100000.f; }
  if (kind == 1 && domain_auto && n > 0) { // bounds reduce (exact min/max)
    float mnx = x[0], mny = y[0], mxx = x[0], mxy = y[0];
    for (size_t i = 1; i < n; ++i) {
      mnx = std::fmin(mnx, x[i]); mny = std::fmin(mny, y[i]);
      mxx = std::fmax(mxx, x[i]); mxy = std::fmax(mxy, y[i]);
    }
    x0 = mnx; y0 = mny;
    float ww = mxx - mnx, hh = mxy - mny;
    h = ww > hh ? ww : hh;
  }
  t.rx0 = x0; t.ry0 = y0; t.rh = h;
  if (cap <= 0) cap = (kind == 0 || kind == 2) ? 8 : 64;
  t.keys.resize(n);
  if (kind == 0 || kind == 2)
    for (size_t i = 0; i < n; ++i)
      t.keys[i] = in_domain(x[i], y[i], x0, y0, h) ? morton_key(x[i], y[i], x0, y0, h) : KEY_OUT;
  else
    for (size_t i = 0; i < n; ++i) t.keys[i] = morton_key(x[i], y[i], x0, y0, h); // clamps

  if (kind == 0 || kind == 1) {
    t.order.resize(n);
    for (size_t i = 0; i < n; ++i) t.order[i] = (uint32_t)i;
    std::stable_sort(t.order.begin(), t.order.end(),
                     [&](uint32_t a, uint32_t b) { return t.keys[a] < t.keys[b]; });
    size_t n_in = n;
    if (kind == 0) {
      n_in = 0;
      while (n_in < n && t.keys[t.order[n_in]] != KEY_OUT) ++n_in;
    }
    t.n_in = n_in;
    if (n_in > 0) {
      if (kind == 0) pipe_quad_rec(t, 0, n_in, 0, x0, y0, h, 0xffffffffu, cap);
      else pipe_lbvh_rec(t, 0, n_in, 0xffffffffu, cap);
    }
    fill_sorted_bodies(t, x, y, w);
    flat_upsweep(t, w, kind);
  } else if (kind == 2) {
    T->qa.reset(new QuadArena());
    T->qroot = T->qa->make(RectQ::make(V2{x0, y0}, h));
    std::vector<uint32_t> outside;
    for (size_t i = 0; i < n; ++i) {
      if (t.keys[i] == KEY_OUT) { outside.push_back((uint32_t)i); continue; }
      quad_insert(*T->qa, T->qroot, SmallParticle{{x[i], y[i]}, w[i], (uint32_t)i}, 0);
    }
    T->overflow = T->qa->overflow;
    quad_calculate_gravity(T->qroot);
    if (n - outside.size() > 0) flatten_quad(T->qroot, 0, 0xffffffffu, t);
    t.n_in = t.order.size();
    for (uint32_t id : outside) t.order.push_back(id);
    fill_sorted_bodies(t, x, y, w);
  } else {
    T->particles.resize(n);
    for (size_t i = 0; i < n; ++i) T->particles[i] = Particle{{x[i], y[i]}, {0.f, 0.f}, w[i], (uint32_t)i};
    T->ba.reset(new BVHArena());
    T->broot = bvh_from(*T->ba, T->particles.data(), n, 0); // always a Root at the top (bvh_tree.rs:90)
    T->overflow = T->ba->overflow;
    bvh_calculate_gravity(T->broot);
    flatten_bvh(T->broot, 0, 0xffffffffu, t, T->particles.data());
    t.order.resize(n);
    for (size_t i = 0; i < n; ++i) t.order[i] = T->particles[i].id;
    t.n_in = n;
    fill_sorted_bodies(t, x, y, w);
  }
  return T;
}

void orc_tree_free(void *h) { delete (OrcTree *)h; }

int orc_tree_info(void *h, nb_tree_info *info) {
  OrcTree *T = (OrcTree *)h;
  const FlatTree &t = T->flat;
  std::memset(info, 0, sizeof(*info));
  info->n_bodies = t.n;
  info->n_in_tree = t.n_in;
  info->n_nodes = t.depth.size();
  uint64_t nl = 0;
  int32_t md = 0;
  for (size_t i = 0; i < t.depth.size(); ++i) {
    nl += t.leaf[i];
    md = std::max(md, t.depth[i]);
  }
  info->n_leaves = nl;
  info->max_depth = (uint32_t)md;
  info->overflow_leaves = t.overflow_leaves + (T->overflow ? 1u : 0u);
  info->root_x0 = t.rx0;
  info->root_y0 = t.ry0;
  info->root_h = t.rh;
  return 0;
}

int orc_tree_export(void *h, const nb_tree_view *v) {
  OrcTree *T = (OrcTree *)h;
  const FlatTree &t = T->flat;
  size_t nn = t.depth.size();
#define CP(dst, src, cnt) if (v->dst) std::memcpy(v->dst, (src).data(), (cnt) * sizeof((src)[0]))
  CP(keys, t.keys, t.n);
  CP(order, t.order, t.n);
  CP(node_depth, t.depth, nn);
  CP(node_start, t.start, nn);
  CP(node_count, t.count, nn);
  CP(node_skip, t.skip, nn);
  CP(node_parent, t.parent, nn);
  CP(node_leaf, t.leaf, nn);
  CP(node_mass, t.mass, nn);
  CP(node_cogx, t.cogx, nn);
  CP(node_cogy, t.cogy, nn);
  CP(node_x0, t.x0, nn);
  CP(node_y0, t.y0, nn);
  CP(node_x1, t.x1, nn);
  CP(node_y1, t.y1, nn);
  CP(node_s2, t.s2, nn);
#undef CP
  return 0;
}

// Accelerations of `nt` targets (original ids; NULL = bodies 0..nt-1).  OpenMP
// schedule(dynamic,5000) mirrors rayon with_min_len(5000) (main.rs:406-416); threads<=0 = all.
// literal != 0 on kinds 2/3 walks the pointer tree recursively exactly like the reference.
int orc_tree_accel(void *h, float theta, const uint32_t *targets, size_t nt, int literal, float *ax,
                   float *ay, uint64_t *interactions, uint64_t *tests, int threads, double *seconds) {
  OrcTree *T = (OrcTree *)h;
  uint64_t ti = 0, tt = 0;
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#else
  threads = 1;
#endif
  double t0 = now_s();
#pragma omp parallel for schedule(dynamic, 5000) num_threads(threads) reduction(+ : ti, tt)
  for (long long k = 0; k < (long long)nt; ++k) {
    uint32_t id = targets ? targets[k] : (uint32_t)k;
    V2 p{T->x[id], T->y[id]};
    V2 acc{0.f, 0.f};
    Counters c;
    if (literal && T->kind == 2) quad_sum_gravity(p, T->qroot, acc, theta, c);
    else if (literal && T->kind == 3) bvh_sum_gravity(p, T->broot, acc, theta, c);
    else flat_sum_gravity(T->flat, p, theta, acc, c);
    ax[k] = acc.x;
    ay[k] = acc.y;
    ti += c.interactions;
    tt += c.tests;
  }
  if (seconds) *seconds = now_s() - t0;
  if (interactions) *interactions = ti;
  if (tests) *tests = tt;
  return 0;
}

// quad_tree.rs:66-137 on a literal quadtree: returns empty() then prune() counts.
int orc_quad_empty_prune(void *h, uint32_t *emptied, uint32_t *pruned) {
  OrcTree *T = (OrcTree *)h;
  if (T->kind != 2) return 1;
  *emptied = quad_empty(T->qroot);
  *pruned = quad_prune(T->qroot);
  return 0;
}

// The same two methods on a FRESH literal tree after calculate_gravity, each on its own: `emptied` = what empty()
// would return (counted without mutating), then `pruned` = prune() (mutates: do not reuse the handle afterwards).
static uint32_t quad_empty_visits(const Quad *q) { // quad_tree.rs:66-89, read-only
  if (q->is_leaf) return 1;
  if (q->total_mass == 0) return 1;
  uint32_t sum = 0;
  for (int i = 0; i < 4; ++i)
    if (q->children[i]) sum += quad_empty_visits(q->children[i]);
  return sum + 1;
}
int orc_quad_fresh_stats(void *h, uint32_t *emptied, uint32_t *pruned) {
  OrcTree *T = (OrcTree *)h;
  if (T->kind != 2) return 1;
  *emptied = quad_empty_visits(T->qroot);
  *pruned = quad_prune(T->qroot);
  return 0;
}

// ---- direct sums ----------------------------------------------------------------------------
// fp64 truth of the same force law (main.rs:235-253) on the f32 inputs.
int orc_direct_f64(const float *x, const float *y, const uint32_t *w, size_t n, const uint32_t *targets,
                   size_t nt, double *ax, double *ay, int threads) {
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
  for (long long k = 0; k < (long long)nt; ++k) {
    uint32_t id = targets ? targets[k] : (uint32_t)k;
    double px = x[id], py = y[id], sx = 0, sy = 0;
    for (size_t j = 0; j < n; ++j) {
      double dx = (double)x[j] - px, dy = (double)y[j] - py;
      double sum = std::fabs(dx) + std::fabs(dy);
      if (!std::isnormal((float)sum)) continue;
      double d = dx * dx + dy * dy;
      if (d < (double)0.001f) d = (double)0.001f;
      double inv = (double)w[j] / (sum * d);
      sx += dx * inv;
      sy += dy * inv;
    }
    ax[k] = sx;
    ay[k] = sy;
  }
  return 0;
}
// fp32 sequential direct sum, literal pair force, sources in index order.
int orc_direct_f32(const float *x, const float *y, const uint32_t *w, size_t n, const uint32_t *targets,
                   size_t nt, float *ax, float *ay, int threads) {
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 64) num_threads(threads)
  for (long long k = 0; k < (long long)nt; ++k) {
    uint32_t id = targets ? targets[k] : (uint32_t)k;
    V2 p{x[id], y[id]}, acc{0.f, 0.f};
    for (size_t j = 0; j < n; ++j) calculate_gravity(p, V2{x[j], y[j]}, acc, (float)w[j]);
    ax[k] = acc.x;
    ay[k] = acc.y;
  }
  return 0;
}

// ---- World (main.rs:37-39, 388-425) ---------------------------------------------------------
struct OrcWorld {
  std::vector<Particle> particles; // for kind 3 this Vec is permuted in place every step
  size_t n;
};

void *orc_world_new(const float *x, const float *y, const float *vx, const float *vy, const uint32_t *w,
                    size_t n) {
  OrcWorld *W = new OrcWorld();
  W->n = n;
  W->particles.resize(n);
  for (size_t i = 0; i < n; ++i) W->particles[i] = Particle{{x[i], y[i]}, {vx[i], vy[i]}, w[i], (uint32_t)i};
  return W;
}
void orc_world_free(void *h) { delete (OrcWorld *)h; }

// Bodies by ORIGINAL id.
void orc_world_get(void *h, float *x, float *y, float *vx, float *vy, uint32_t *w) {
  OrcWorld *W = (OrcWorld *)h;
  for (const Particle &p : W->particles) {
    if (x) x[p.id] = p.position.x;
    if (y) y[p.id] = p.position.y;
    if (vx) vx[p.id] = p.velocity.x;
    if (vy) vy[p.id] = p.velocity.y;
    if (w) w[p.id] = p.weight;
  }
}

// One World::update(delta).  kind as in orc_tree_build (0/1 pipeline trees, 2 literal quadtree
// rebuilt fresh each step, 3 literal bvh_tree.rs = the reference's live path).
// faithful_bug != 0 (kind 3 only) applies acceleration[idx] to the *permuted* particle idx exactly
// as main.rs:419-423 does; otherwise acceleration follows its body (identity).
// sample_stride > 1 computes the traversal only for every stride-th body (timing of infeasible
// sizes; the state is then NOT advanced) and scales nothing: counting/interactions cover the sample.
int orc_world_update(void *h, int kind, float delta, float theta, float x0, float y0, float dh, int cap,
                     int faithful_bug, int threads, size_t sample_stride, nb_counting *out) {
  OrcWorld *W = (OrcWorld *)h;
  size_t n = W->n;
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#else
  threads = 1;
#endif
  if (sample_stride < 1) sample_stride = 1;
  double t0 = now_s();
  std::vector<Particle> cloned = W->particles; // main.rs:398
  std::vector<V2> acceleration(n, V2{0.f, 0.f});
  uint64_t ti = 0, tt = 0;
  double t1, t2;
  if (kind == 3) {
    BVHArena A;
    BVH *root = bvh_from(A, W->particles.data(), n, 0); // permutes W->particles (main.rs:400)
    bvh_calculate_gravity(root);
    t1 = now_s();
#pragma omp parallel for schedule(dynamic, 5000) num_threads(threads) reduction(+ : ti, tt)
    for (long long i = 0; i < (long long)n; i += (long long)sample_stride) {
      V2 acc{0.f, 0.f};
      Counters c;
      bvh_sum_gravity(cloned[i].position, root, acc, theta, c);
      acceleration[i] = acc;
      ti += c.interactions;
      tt += c.tests;
    }
    t2 = now_s();
    if (sample_stride == 1) {
      // acceleration[i] belongs to cloned[i]; cloned order == pre-permutation Vec order.
      if (faithful_bug) {
        for (size_t idx = 0; idx < n; ++idx) { // main.rs:419-423 verbatim
          Particle &p = W->particles[idx];
          p.velocity = p.velocity + acceleration[idx] * delta;
          V2 velo = p.velocity * delta;
          p.position = p.position + velo;
        }
      } else {
        std::vector<uint32_t> pos_of_id(n);
        for (size_t k = 0; k < n; ++k) pos_of_id[cloned[k].id] = (uint32_t)k;
        for (size_t idx = 0; idx < n; ++idx) {
          Particle &p = W->particles[idx];
          p.velocity = p.velocity + acceleration[pos_of_id[p.id]] * delta;
          V2 velo = p.velocity * delta;
          p.position = p.position + velo;
        }
      }
    }
  } else if (kind == 2) {
    // literal quad_tree.rs, timed as the reference would run it: QuadTree::new, insert every in-bounds body
    // (within_bounds, main.rs:223-226), calculate_gravity.  No Morton keys, no flattening: the reference's
    // insert computes neither, and the recursive walk below needs neither.
    if (dh <= 0.f) { x0 = 0.f; y0 = 0.f; dh = 100000.f; }
    QuadArena A;
    Quad *root = A.make(RectQ::make(V2{x0, y0}, dh));
    for (size_t i = 0; i < n; ++i) {
      const Particle &q = cloned[i];
      if (in_domain(q.position.x, q.position.y, x0, y0, dh))
        quad_insert(A, root, SmallParticle{q.position, q.weight, (uint32_t)i}, 0);
    }
    quad_calculate_gravity(root);
    t1 = now_s();
#pragma omp parallel for schedule(dynamic, 5000) num_threads(threads) reduction(+ : ti, tt)
    for (long long i = 0; i < (long long)n; i += (long long)sample_stride) {
      V2 acc{0.f, 0.f};
      Counters c;
      quad_sum_gravity(cloned[i].position, root, acc, theta, c);
      acceleration[i] = acc;
      ti += c.interactions;
      tt += c.tests;
    }
    t2 = now_s();
    if (sample_stride == 1) {
      for (size_t idx = 0; idx < n; ++idx) {
        Particle &p = W->particles[idx];
        p.velocity = p.velocity + acceleration[idx] * delta;
        V2 velo = p.velocity * delta;
        p.position = p.position + velo;
      }
    }
  } else {
    std::vector<float> x(n), y(n);
    std::vector<uint32_t> w(n);
    for (size_t i = 0; i < n; ++i) {
      x[i] = cloned[i].position.x;
      y[i] = cloned[i].position.y;
      w[i] = cloned[i].weight;
    }
    OrcTree *T = (OrcTree *)orc_tree_build(kind, x.data(), y.data(), w.data(), n, x0, y0, dh, cap, 0);
    t1 = now_s();
#pragma omp parallel for schedule(dynamic, 5000) num_threads(threads) reduction(+ : ti, tt)
    for (long long i = 0; i < (long long)n; i += (long long)sample_stride) {
      V2 acc{0.f, 0.f};
      Counters c;
      if (kind == 2) quad_sum_gravity(cloned[i].position, T->qroot, acc, theta, c);
      else flat_sum_gravity(T->flat, cloned[i].position, theta, acc, c);
      acceleration[i] = acc;
      ti += c.interactions;
      tt += c.tests;
    }
    t2 = now_s();
    orc_tree_free(T);
    if (sample_stride == 1) {
      for (size_t idx = 0; idx < n; ++idx) {
        Particle &p = W->particles[idx];
        p.velocity = p.velocity + acceleration[idx] * delta;
        V2 velo = p.velocity * delta;
        p.position = p.position + velo;
      }
    }
  }
  double t3 = now_s();
  if (out) {
    out->build_s += t1 - t0;
    out->traverse_s += t2 - t1;
    out->integrate_s += t3 - t2;
    out->steps += 1;
    out->interactions += ti;
    out->node_tests += tt;
  }
  return 0;
}


// ---- design tool: statistics of a group-cooperative walk (G tree-order-adjacent bodies walk the
// union of their private walks, every body keeping its own decisions).  Not used by any test of
// the product; it sizes the GPU traversal (SIMD efficiency of warp-cooperative walks).
// out[0]=groups, [1]=internal node visits, [2]=leaf visits, [3]=sum of unmasked bodies at internal
// visits (= opening tests), [4]=accepted (node interactions), [5]=leaf-body pair interactions,
// [6]=sum over leaf visits of leaf body count (uniform loop trips), [7]=sum of unmasked bodies at leaf visits,
// [8]/[9]/[10]=internal visits where every unmasked body accepts / none accepts / they disagree,
// [11]=internal visits with all G bodies unmasked
int orc_group_walk_stats(void *h, float theta, int G, double *out) {
  OrcTree *T = (OrcTree *)h;
  const FlatTree &t = T->flat;
  const uint32_t nn = (uint32_t)t.depth.size();
  const size_t n = t.n;
  double acc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
#pragma omp parallel
  {
    double a[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    std::vector<uint32_t> resume(G);
#pragma omp for schedule(dynamic, 64)
    for (long long g0 = 0; g0 < (long long)n; g0 += G) {
      int gs = (int)std::min<size_t>(G, n - g0);
      for (int k = 0; k < gs; ++k) resume[k] = 0;
      a[0] += 1;
      uint32_t cur = 0;
      while (cur < nn) {
        if (t.leaf[cur]) {
          int act = 0;
          for (int k = 0; k < gs; ++k) act += (cur >= resume[k]);
          a[2] += 1; a[6] += t.count[cur]; a[7] += act; a[5] += (double)act * t.count[cur];
          cur = cur + 1;
        } else {
          a[1] += 1;
          bool any_open = false;
          int n_act = 0, n_acc = 0;
          V2 cog{t.cogx[cur], t.cogy[cur]};
          for (int k = 0; k < gs; ++k) {
            if (cur < resume[k]) continue;
            a[3] += 1;
            n_act++;
            V2 p{t.sx[g0 + k], t.sy[g0 + k]};
            bool contains = p.y > t.y0[cur] && p.x > t.x0[cur] && p.x < t.x1[cur] && p.y < t.y1[cur];
            if (!contains && t.s2[cur] < dist2(p, cog) * theta * theta) { a[4] += 1; n_acc++; resume[k] = t.skip[cur]; }
            else any_open = true;
          }
          a[n_acc == n_act ? 8 : (n_acc == 0 ? 9 : 10)] += 1;
          if (n_act == gs) a[11] += 1;
          cur = any_open ? cur + 1 : t.skip[cur];
        }
      }
    }
#pragma omp critical
    for (int k = 0; k < 12; ++k) acc[k] += a[k];
  }
  for (int k = 0; k < 12; ++k) out[k] = acc[k];
  return 0;
}

// Histogram of leaf-loop trips by the number of unmasked bodies in the group at the leaf visit:
// hist[a] = sum over leaf visits with a unmasked bodies of the leaf's body count (a = 0..G).  Kernel-design aid.
int orc_group_walk_leaf_hist(void *h, float theta, int G, double *hist) {
  OrcTree *T = (OrcTree *)h;
  const FlatTree &t = T->flat;
  const uint32_t nn = (uint32_t)t.depth.size();
  const size_t n = t.n;
  std::vector<double> acc(G + 1, 0.0);
#pragma omp parallel
  {
    std::vector<double> a(G + 1, 0.0);
    std::vector<uint32_t> resume(G);
#pragma omp for schedule(dynamic, 64)
    for (long long g0 = 0; g0 < (long long)n; g0 += G) {
      int gs = (int)std::min<size_t>(G, n - g0);
      for (int k = 0; k < gs; ++k) resume[k] = 0;
      uint32_t cur = 0;
      while (cur < nn) {
        if (t.leaf[cur]) {
          int act = 0;
          for (int k = 0; k < gs; ++k) act += (cur >= resume[k]);
          a[act] += t.count[cur];
          cur = cur + 1;
        } else {
          bool any_open = false;
          V2 cog{t.cogx[cur], t.cogy[cur]};
          for (int k = 0; k < gs; ++k) {
            if (cur < resume[k]) continue;
            V2 p{t.sx[g0 + k], t.sy[g0 + k]};
            bool contains = p.y > t.y0[cur] && p.x > t.x0[cur] && p.x < t.x1[cur] && p.y < t.y1[cur];
            if (!contains && t.s2[cur] < dist2(p, cog) * theta * theta) resume[k] = t.skip[cur];
            else any_open = true;
          }
          cur = any_open ? cur + 1 : t.skip[cur];
        }
      }
    }
#pragma omp critical
    for (int k = 0; k <= G; ++k) acc[k] += a[k];
  }
  for (int k = 0; k <= G; ++k) hist[k] = acc[k];
  return 0;
}

// ---- design tool (round 2): statistics of a group BREADTH-FIRST walk with a conservative group-box classification
// (sized csrc/traverse_group.cuh; the numbers in profiles/r02a_traverse_group_kernel_ncu.md and DESIGN.md K5 come
// from here: scripts/group_walk_stats.py).  G tree-order-adjacent bodies share a frontier of (node, mask) entries; an
// entry is "all accept" when the node passes the opening test for the nearest point of the group's bounding box,
// "all open" when it fails for the farthest corner, "mixed" otherwise (every body of the mask then decides for
// itself, exactly as main.rs:370-372).  out[0]=groups, [1]=entries classified, [2]=all-accept with a full mask,
// [3]=all-accept with a partial mask, [4]=all-open, [5]=mixed, [6]=full-mask items (nodes + leaf bodies),
// [7]=partial-mask items, [8]=sum of popcounts over partial-mask items, [9]=sum over groups of the largest per-body
// partial-item count, [10]=frontier batches of 32 entries, [11]=pair interactions, [12]=leaf entries, [13]=sum over
// groups of the largest per-body item count, [14]=levels.  hist[0][p] / hist[1][p]: node / body items by popcount p.
int orc_group_bfs_stats(void *h, float theta, int G, double *out, double *hist /* [2][33] */) {
  OrcTree *T = (OrcTree *)h;
  const FlatTree &t = T->flat;
  const size_t n = t.n_in;
  const int NS = 15;
  double acc[NS] = {0};
  const float th2 = theta * theta;
#pragma omp parallel
  {
    double a[NS] = {0};
    double hl[66] = {0};
    std::vector<std::pair<uint32_t, uint32_t>> fr, nx;
    std::vector<uint32_t> pcnt(G), tcnt(G);
#pragma omp for schedule(dynamic, 64)
    for (long long g0 = 0; g0 < (long long)n; g0 += G) {
      int gs = (int)std::min<size_t>(G, n - g0);
      const uint32_t full = gs == 32 ? 0xffffffffu : ((1u << gs) - 1u);
      float gx0 = 1e30f, gy0 = 1e30f, gx1 = -1e30f, gy1 = -1e30f;
      for (int k = 0; k < gs; ++k) {
        gx0 = std::min(gx0, t.sx[g0 + k]); gx1 = std::max(gx1, t.sx[g0 + k]);
        gy0 = std::min(gy0, t.sy[g0 + k]); gy1 = std::max(gy1, t.sy[g0 + k]);
      }
      for (int k = 0; k < gs; ++k) pcnt[k] = tcnt[k] = 0;
      a[0] += 1;
      fr.clear();
      fr.push_back({0u, full});
      auto item = [&](uint32_t mask, uint32_t cnt, int isbody) {
        hl[isbody * 33 + __builtin_popcount(mask)] += cnt;
        a[11] += (double)__builtin_popcount(mask) * cnt;
        for (int k = 0; k < gs; ++k) if (mask >> k & 1) tcnt[k] += cnt;
        if (mask == full) a[6] += cnt;
        else {
          a[7] += cnt; a[8] += (double)__builtin_popcount(mask) * cnt;
          for (int k = 0; k < gs; ++k) if (mask >> k & 1) pcnt[k] += cnt;
        }
      };
      while (!fr.empty()) {
        a[14] += 1;
        a[10] += (double)((fr.size() + 31) / 32);
        nx.clear();
        for (auto [cur, mask] : fr) {
          a[1] += 1;
          if (t.leaf[cur]) { a[12] += 1; item(mask, t.count[cur], 1); continue; }
          float cx = t.cogx[cur], cy = t.cogy[cur];
          float ddx = std::max(std::max(gx0 - cx, cx - gx1), 0.f), ddy = std::max(std::max(gy0 - cy, cy - gy1), 0.f);
          float dmin2 = ddx * ddx + ddy * ddy;
          float ex = std::max(cx - gx0, gx1 - cx), ey = std::max(cy - gy0, gy1 - cy);
          float dmax2 = ex * ex + ey * ey;
          bool disjoint = t.x1[cur] <= gx0 || t.x0[cur] >= gx1 || t.y1[cur] <= gy0 || t.y0[cur] >= gy1;
          uint32_t open = 0;
          if (disjoint && t.s2[cur] < dmin2 * th2 * 0.9999f) {
            a[mask == full ? 2 : 3] += 1;
            item(mask, 1, 0);
          } else if (!(t.s2[cur] < dmax2 * th2 * 1.0001f)) {
            a[4] += 1;
            open = mask;
          } else {
            a[5] += 1;
            uint32_t am = 0;
            for (int k = 0; k < gs; ++k) {
              if (!(mask >> k & 1)) continue;
              V2 p{t.sx[g0 + k], t.sy[g0 + k]};
              bool contains = p.y > t.y0[cur] && p.x > t.x0[cur] && p.x < t.x1[cur] && p.y < t.y1[cur];
              if (!contains && t.s2[cur] < dist2(p, V2{cx, cy}) * theta * theta) am |= 1u << k;
            }
            if (am) item(am, 1, 0);
            open = mask & ~am;
          }
          if (open)
            for (uint32_t c = cur + 1; c < t.skip[cur]; c = t.skip[c]) nx.push_back({c, open});
        }
        fr.swap(nx);
      }
      uint32_t mp = 0, mt = 0;
      for (int k = 0; k < gs; ++k) { mp = std::max(mp, pcnt[k]); mt = std::max(mt, tcnt[k]); }
      a[9] += mp; a[13] += mt;
    }
#pragma omp critical
    {
      for (int k = 0; k < NS; ++k) acc[k] += a[k];
      if (hist) for (int k = 0; k < 66; ++k) hist[k] += hl[k];
    }
  }
  for (int k = 0; k < NS; ++k) out[k] = acc[k];
  return 0;
}

int orc_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

} // extern "C"
