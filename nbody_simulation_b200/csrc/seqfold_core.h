// seqfold_core.h — bit-exact PARALLEL evaluation of a SEQUENTIAL f32 fold  acc = fl(acc + x[i]).
//
// Why: BVHTree::from (reference src/bvh_tree.rs:58-67) takes the split point of every node from
// `sum.add(p.position)` folded left to right in f32.  At 1M bodies the accumulator's ulp reaches 4096
// while the addends are ~5e4, so the mean — and with it which bodies go left — depends on the exact
// rounding sequence.  A tree-shaped parallel sum gives a different tree.  This header reproduces the
// sequential result exactly, in parallel:
//
//   While the accumulator stays in one binade [2^k, 2^(k+1)) its ulp u = 2^(k-23) is fixed and
//   A = |acc|/u is an integer in [2^23, 2^24).  Adding x is then integer arithmetic:
//        A' = A + n + r,   n = floor(t), t = sign(acc)*x/u,  r = round-to-nearest-even of frac(t)
//   and r depends on the running state only through the PARITY of A (ties to even).  A chunk of
//   consecutive addends is therefore a function  parity -> (total, min, max of the running floor),
//   and such functions compose associatively: a block scan gives every chunk its exact incoming A.
//   min/max tell whether the accumulator left the binade inside the chunk; the first addend where it
//   does is applied with a real f32 add and the scan restarts in the new binade.
//
// Host+device code (NB_HD) so the CPU test tier can check it against a plain loop.
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define NB_HD __host__ __device__ __forceinline__
#else
#define NB_HD inline
#endif
#ifdef __CUDA_ARCH__
#define NB_UNROLL _Pragma("unroll")
#else
#define NB_UNROLL
#endif

namespace nb {
namespace sf {

constexpr int32_t kBig = 1 << 28;      // saturation bound of every summary field
constexpr int32_t kLoA = 1 << 23;      // A in [kLoA, kHiA)
constexpr int32_t kHiA = 1 << 24;

NB_HD uint32_t f2u(float f) {
#ifdef __CUDA_ARCH__
  return __float_as_uint(f);
#else
  uint32_t u;
  memcpy(&u, &f, 4);
  return u;
#endif
}
NB_HD float u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}
NB_HD float fadd(float a, float b) {
#ifdef __CUDA_ARCH__
  return __fadd_rn(a, b);
#else
  volatile float r = a + b; // the oracle and this file are compiled without fast-math
  return r;
#endif
}
NB_HD int32_t clampi(int64_t v) { return v > kBig ? kBig : (v < -kBig ? -kBig : (int32_t)v); }

// The accumulator as sign * A * 2^(E-150), E = biased exponent (>= 1), A in [2^23, 2^24).
struct Binade {
  uint32_t E;    // biased exponent of acc
  uint32_t sign; // 0 / 0x80000000
  int32_t A;     // 24-bit significand
};
// false: acc is zero, subnormal or not finite -> take plain f32 steps until it is normal.
NB_HD bool binade_of(float acc, Binade &b) {
  const uint32_t u = f2u(acc);
  b.E = (u >> 23) & 0xffu;
  b.sign = u & 0x80000000u;
  b.A = (int32_t)((u & 0x007fffffu) | 0x00800000u);
  return b.E >= 1u && b.E <= 254u;
}
NB_HD float binade_value(const Binade &b, int32_t A) { // A in [2^23, 2^24)
  return u2f(b.sign | (b.E << 23) | ((uint32_t)A & 0x007fffffu));
}

// One addend relative to a binade: t = sign(acc)*x/u = n + f, f in [0,1);
// c = 0 (f < 1/2, round down), 1 (f == 1/2, tie), 2 (f > 1/2, round up);  bad: cannot be handled here.
struct QElem {
  int32_t n;
  int32_t c; // 0,1,2 or 3 = bad
};
NB_HD QElem quantize(float x, const Binade &b) {
  const uint32_t u = f2u(x);
  const uint32_t ex = (u >> 23) & 0xffu;
  QElem q;
  q.n = 0;
  q.c = 0;
  if (ex == 255u) { q.c = 3; return q; } // inf / nan
  uint32_t m = u & 0x007fffffu;
  if (ex != 0u) m |= 0x00800000u;
  if (m == 0u) return q; // +-0
  const int sh = (int)b.E - (int)(ex ? ex : 1u); // x = m * 2^(ex'-150); u = 2^(E-150); t = m >> sh
  if (sh < 0) { q.c = 3; return q; } // addend in a higher binade than the accumulator
  uint32_t ip, cls;
  bool exact;
  if (sh == 0) { ip = m; cls = 0; exact = true; }
  else if (sh >= 26) { ip = 0; cls = 0; exact = false; } // below a quarter ulp
  else {
    ip = m >> sh;
    const uint32_t r = m & ((1u << sh) - 1u), half = 1u << (sh - 1);
    exact = (r == 0u);
    cls = r < half ? 0u : (r == half ? 1u : 2u);
  }
  if (((u ^ b.sign) & 0x80000000u) == 0u) { // same sign as the accumulator
    q.n = (int32_t)ip;
    q.c = (int32_t)cls;
  } else if (exact) {
    q.n = -(int32_t)ip;
    q.c = 0;
  } else {
    q.n = -(int32_t)ip - 1;
    q.c = 2 - (int32_t)cls;
  }
  return q;
}

// Chunk summary for both incoming parities p of A.
struct Summary {
  int32_t tot[2];  // A_out - A_in
  int32_t lo[2];   // min over addends of (running - A_in) + n   (the floor before rounding)
  int32_t hi[2];   // max of the same
};
NB_HD Summary identity() {
  Summary s;
  s.tot[0] = s.tot[1] = 0;
  s.lo[0] = s.lo[1] = kBig;
  s.hi[0] = s.hi[1] = -kBig;
  return s;
}
// append one addend to a summary
NB_HD void push(Summary &s, const QElem &q) {
  NB_UNROLL
  for (int p = 0; p < 2; ++p) {
    if (q.c == 3) { s.lo[p] = -kBig; s.hi[p] = kBig; continue; }
    const int64_t f = (int64_t)s.tot[p] + q.n;
    const int32_t fl = clampi(f);
    s.lo[p] = fl < s.lo[p] ? fl : s.lo[p];
    s.hi[p] = fl > s.hi[p] ? fl : s.hi[p];
    const int32_t odd = (int32_t)((uint32_t)(p + fl) & 1u);
    s.tot[p] = clampi(f + (q.c == 2 ? 1 : (q.c == 1 ? odd : 0)));
  }
}
// push() for a quantised addend known not to be bad and a chunk short enough (<= 64 addends of |n| <= 2^24)
// that the int32 fields cannot overflow: (c >> 1) is "round up", (c & (f ^ p) & 1) is "tie and floor odd".
NB_HD void push_fast(Summary &s, int32_t n, int32_t c) {
  NB_UNROLL
  for (int p = 0; p < 2; ++p) {
    const int32_t f = s.tot[p] + n;
    s.lo[p] = f < s.lo[p] ? f : s.lo[p];
    s.hi[p] = f > s.hi[p] ? f : s.hi[p];
    s.tot[p] = f + ((c >> 1) | (c & (f ^ p) & 1));
  }
}
NB_HD void poison(Summary &s) { // a chunk holding an addend the binade cannot take
  s.lo[0] = s.lo[1] = -kBig;
  s.hi[0] = s.hi[1] = kBig;
}
// h = g after f
NB_HD Summary compose(const Summary &f, const Summary &g) {
  Summary h;
  NB_UNROLL
  for (int p = 0; p < 2; ++p) {
    const int q = (int)((uint32_t)(p + f.tot[p]) & 1u);
    h.tot[p] = clampi((int64_t)f.tot[p] + g.tot[q]);
    const int32_t l2 = clampi((int64_t)f.tot[p] + g.lo[q]), h2 = clampi((int64_t)f.tot[p] + g.hi[q]);
    h.lo[p] = f.lo[p] < l2 ? f.lo[p] : l2;
    h.hi[p] = f.hi[p] > h2 ? f.hi[p] : h2;
    if (g.hi[q] < g.lo[q]) { h.lo[p] = f.lo[p]; h.hi[p] = f.hi[p]; } // g empty
  }
  return h;
}
// Every add of the summarised run, started from significand A, stays inside the binade (conservative:
// floor in [2^23, 2^24 - 2], so the rounded value is < 2^24 too).
NB_HD bool valid_from(const Summary &s, int32_t A) {
  const int p = A & 1;
  if (s.hi[p] < s.lo[p]) return true; // empty
  return (int64_t)A + s.lo[p] >= kLoA && (int64_t)A + s.hi[p] <= kHiA - 2;
}
NB_HD bool step_valid(const QElem &q, int32_t A) {
  return q.c != 3 && (int64_t)A + q.n >= kLoA && (int64_t)A + q.n <= kHiA - 2;
}
NB_HD int32_t step_apply(const QElem &q, int32_t A) {
  const int32_t fl = A + q.n;
  return fl + (q.c == 2 ? 1 : (q.c == 1 ? (fl & 1) : 0));
}

// Reference implementation of the windowed algorithm, serial (host tests; also the shape of the device
// kernel): `chunk` addends per "thread", `threads` chunks per window.
inline float fold_windowed_serial(const float *x, size_t n, float acc, int chunk, int threads,
                                  unsigned long long *exact_steps = nullptr) {
  size_t pos = 0;
  unsigned long long slow = 0;
  while (pos < n) {
    Binade b;
    if (!binade_of(acc, b)) { acc = fadd(acc, x[pos++]); ++slow; continue; }
    const size_t wlen = (n - pos < (size_t)chunk * threads) ? n - pos : (size_t)chunk * threads;
    // per-chunk summaries, scanned
    Summary pre = identity();
    int32_t A = b.A;
    bool broke = false;
    for (size_t c0 = 0; c0 < wlen && !broke; c0 += chunk) {
      const size_t ce = c0 + chunk < wlen ? c0 + chunk : wlen;
      Summary s = identity();
      for (size_t i = c0; i < ce; ++i) push(s, quantize(x[pos + i], b));
      const int32_t Ain = b.A + pre.tot[b.A & 1];
      if (valid_from(s, Ain)) { pre = compose(pre, s); A = b.A + pre.tot[b.A & 1]; continue; }
      // first invalid chunk: walk it
      int32_t Ar = Ain;
      for (size_t i = c0; i < ce; ++i) {
        const QElem q = quantize(x[pos + i], b);
        if (!step_valid(q, Ar)) {
          acc = fadd(binade_value(b, Ar), x[pos + i]);
          ++slow;
          pos += i + 1;
          broke = true;
          break;
        }
        Ar = step_apply(q, Ar);
      }
    }
    if (!broke) { acc = binade_value(b, A); pos += wlen; }
  }
  if (exact_steps) *exact_steps = slow;
  return acc;
}

} // namespace sf
} // namespace nb
