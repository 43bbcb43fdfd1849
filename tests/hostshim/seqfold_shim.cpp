// Host build of nbody_simulation_b200/csrc/seqfold_core.h for the CPU test tier (g++, no CUDA):
// the windowed parity-automaton fold against a plain sequential f32 loop.
#include "../../nbody_simulation_b200/csrc/seqfold_core.h"
#include <stddef.h>

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

extern "C" {
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
