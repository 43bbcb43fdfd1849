// Host build of nbody_simulation_b200/csrc/seqfold_core.h for the CPU test tier (g++, no CUDA):
// the windowed parity-automaton fold against a plain sequential f32 loop.
#include "../../nbody_simulation_b200/csrc/seqfold_core.h"
#include <stddef.h>

#include <algorithm>
#include <vector>
// The device kernel's shape (csrc/bvh_ref.cuh block_seqfold): chunk summaries, a Hillis-Steele inclusive
// scan with compose() (exercises associativity), the first invalid chunk walks.
static float fold_scan(const float *x, size_t n, float acc, int chunk, int threads) {
  using namespace nb::sf;
  size_t pos = 0;
  std::vector<Summary> own(threads), inc(threads), tmp(threads);
  while (pos < n) {
    Binade b;
    if (!binade_of(acc, b)) { acc = fadd(acc, x[pos++]); continue; }
    const size_t W = (size_t)chunk * threads, wlen = n - pos < W ? n - pos : W;
    for (int t = 0; t < threads; ++t) {
      own[t] = identity();
      bool bad = false;
      for (int j = 0; j < chunk; ++j) { // as window_scan() does on the device (chunk <= 32)
        size_t i = (size_t)t * chunk + j;
        if (i < wlen) {
          QElem q = quantize(x[pos + i], b);
          bad = bad || q.c == 3;
          push_fast(own[t], q.n, q.c);
        }
      }
      if (bad) poison(own[t]);
      inc[t] = own[t];
    }
    for (int o = 1; o < threads; o <<= 1) {
      for (int t = 0; t < threads; ++t) tmp[t] = t >= o ? compose(inc[t - o], inc[t]) : inc[t];
      inc.swap(tmp);
    }
    int bad = -1;
    for (int t = 0; t < threads && bad < 0; ++t) {
      Summary pre = t ? inc[t - 1] : identity();
      if (valid_from(pre, b.A) && !valid_from(own[t], b.A + pre.tot[b.A & 1])) bad = t;
    }
    if (bad < 0) { acc = binade_value(b, b.A + inc[threads - 1].tot[b.A & 1]); pos += wlen; continue; }
    Summary pre = bad ? inc[bad - 1] : identity();
    int32_t Ar = b.A + pre.tot[b.A & 1];
    for (int j = 0; j < chunk; ++j) {
      size_t i = (size_t)bad * chunk + j;
      QElem q = quantize(x[pos + i], b);
      if (!step_valid(q, Ar)) { acc = fadd(binade_value(b, Ar), x[pos + i]); pos += i + 1; break; }
      Ar = step_apply(q, Ar);
    }
  }
  return acc;
}

// The device pipeline for big nodes (csrc/bvh_ref.cuh): fp64 tile prefix -> binade guess -> tile summary ->
// sequential resolve with plain folds where a guess does not hold.  Returns the sum; *plain = tiles folded plainly.
static float fold_tiles(const float *x, size_t n, int tile, unsigned long long *plain) {
  using namespace nb::sf;
  size_t nt = (n + tile - 1) / tile;
  std::vector<double> pre(nt, 0.0);
  double run = 0.0;
  for (size_t j = 0; j < nt; ++j) {
    pre[j] = run;
    size_t e = std::min(n, (j + 1) * (size_t)tile);
    double t = 0.0;
    for (size_t i = j * tile; i < e; ++i) t += (double)x[i];
    run += t;
  }
  float acc = 0.f;
  unsigned long long np = 0;
  for (size_t j = 0; j < nt; ++j) {
    size_t b0 = j * (size_t)tile, e = std::min(n, b0 + tile);
    bool fast = false;
    Binade g, b;
    if (j > 0 && binade_of((float)pre[j], g) && binade_of(acc, b) && g.E == b.E && g.sign == b.sign) {
      Summary all = identity();
      const size_t per = (size_t)tile / 32; // addends per lane, composed in lane order
      for (size_t c0 = b0; c0 < e; c0 += per) {
        Summary own = identity();
        bool bad = false;
        for (size_t i = c0; i < std::min(e, c0 + per); ++i) {
          QElem q = quantize(x[i], g);
          bad = bad || q.c == 3;
          push_fast(own, q.n, q.c);
        }
        if (bad) poison(own);
        all = compose(all, own);
      }
      if (valid_from(all, b.A)) { acc = binade_value(b, b.A + all.tot[b.A & 1]); fast = true; }
    }
    if (!fast) {
      ++np;
      for (size_t i = b0; i < e; ++i) acc = fadd(acc, x[i]);
    }
  }
  if (plain) *plain = np;
  return acc;
}

extern "C" {
float nbt_fold_tiles(const float *x, size_t n, int tile, unsigned long long *plain) { return fold_tiles(x, n, tile, plain); }
float nbt_fold_scan(const float *x, size_t n, float acc, int chunk, int threads) {
  return fold_scan(x, n, acc, chunk, threads);
}
float nbt_fold_plain(const float *x, size_t n, float acc) {
  for (size_t i = 0; i < n; ++i) acc = nb::sf::fadd(acc, x[i]);
  return acc;
}
float nbt_fold_windowed(const float *x, size_t n, float acc, int chunk, int threads, unsigned long long *slow) {
  return nb::sf::fold_windowed_serial(x, n, acc, chunk, threads, slow);
}
}
