// scene.cpp — host-side synthetic scenes (no GPU needed).
//
// NB_SCENE_DEFAULT is a seeded restatement of `World::new()` (reference src/main.rs:276-346) with
// its helpers rand_disc / rand_body / rotate_right (main.rs:255-273).  The reference draws from
// unseeded thread-local RNGs (rand::thread_rng main.rs:278, fastrand::f32 main.rs:256-257), so its
// scene is only statistically reproducible; here every draw is a counter-based hash of
// (seed, stream, index) so any host regenerates identical bits.
// The other scenes are the BASELINE.json configs made concrete in SURVEY.md §8(d).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/nbody_b200.h"

namespace {

inline uint64_t mix64(uint64_t z) { // splitmix64 finaliser
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// uniform f32 in [0,1): 24 random mantissa bits (what fastrand::f32 / rand's Standard give)
inline float u01(uint64_t seed, uint32_t stream, uint64_t idx) {
  uint64_t r = mix64(mix64(seed ^ (0xD1B54A32D192ED03ull * (stream + 1))) + idx * 0x9E3779B97F4A7C15ull);
  return (float)(r >> 40) * (1.0f / 16777216.0f);
}

struct V2 { float x, y; };
inline float dist2(V2 a, V2 b) { float dx = a.x - b.x, dy = a.y - b.y; return dx * dx + dy * dy; }

constexpr float TAU = 6.28318530717958647692f; // std::f32::consts::TAU

// main.rs:255-258
inline V2 rand_disc(uint64_t seed, uint32_t stream, uint64_t idx) {
  float theta = u01(seed, stream, 2 * idx) * TAU;
  float r = u01(seed, stream, 2 * idx + 1);
  return {std::cos(theta) * r, std::sin(theta) * r};
}

inline nb_particle mk(float x, float y, float vx, float vy, uint32_t w) {
  nb_particle p;
  p.x = x; p.y = y; p.vx = vx; p.vy = vy; p.weight = w; p._pad = 0;
  return p;
}

struct Sink {
  nb_particle *out;
  size_t cap, n = 0;
  void push(const nb_particle &p) {
    if (n < cap && out) out[n] = p;
    ++n;
  }
};

// main.rs:276-346
void scene_default(uint64_t seed, Sink &s) {
  const uint32_t HEIGHT = 100000;
  const V2 circle1{35000.0f, 35000.0f}, circle2{60000.0f, 60000.0f};
  const float c1lenr2 = 15000000.0f;
  s.push(mk(circle1.x, circle1.y, 200.0f, 250.0f, 75000000u)); // :282-286
  s.push(mk(circle2.x, circle2.y, 0.0f, 0.0f, 750000u));       // :287-291
  const uint32_t lim = (HEIGHT / 14) - 1;                       // :316
  // Only grid points within sqrt(c1lenr2) of circle2 can pass :319; restricting the loop bounds
  // keeps the reference's x-outer / y-inner push order.
  const uint32_t lo = (uint32_t)((60000.0 - 3900.0) / 14.0), hi = (uint32_t)((60000.0 + 3900.0) / 14.0) + 1;
  for (uint32_t x = lo; x < lim && x <= hi; ++x) {
    for (uint32_t y = lo; y < lim && y <= hi; ++y) {
      V2 pos{(float)x * 14.0f, (float)y * 14.0f};
      float d2 = dist2(pos, circle2);
      if (d2 < c1lenr2 && d2 > 500000.0f) {
        float hi_r = (c1lenr2 - d2) + 1.0f;                 // gen_range(0f32..hi_r), :321
        float g = u01(seed, 0, (uint64_t)x * 8192 + y) * hi_r;
        if (g > 6000000.0f) {
          float k = std::sqrt(std::sqrt(750000.0f) / d2);    // :323-324
          V2 d{pos.x - circle2.x, pos.y - circle2.y};
          V2 rot{d.y, -d.x};                                 // rotate_right :271-273
          s.push(mk(pos.x, pos.y, rot.x * k, rot.y * k, 1u));
        }
      }
    }
  }
  const V2 offset{50000.0f, 50000.0f};
  for (uint64_t i = 0; i < 100000; ++i) { // :334-335, rand_body :260-269
    V2 pos = rand_disc(seed, 1, i), vel = rand_disc(seed, 2, i);
    s.push(mk(pos.x * 25000.0f + offset.x, pos.y * 25000.0f + offset.y, vel.x, vel.y, 1u));
  }
}

void scene_uniform_disc(uint64_t seed, size_t n, Sink &s) {
  for (uint64_t i = 0; i < n; ++i) {
    double a = 6.283185307179586 * (double)u01(seed, 3, 2 * i);
    double r = 45000.0 * std::sqrt((double)u01(seed, 3, 2 * i + 1));
    s.push(mk((float)(50000.0 + r * std::cos(a)), (float)(50000.0 + r * std::sin(a)), 0.f, 0.f, 1u));
  }
}

void scene_plummer(uint64_t seed, size_t n, Sink &s) {
  for (uint64_t i = 0; i < n; ++i) {
    double a = 6.283185307179586 * (double)u01(seed, 4, 2 * i);
    double u = 0.999 * (double)u01(seed, 4, 2 * i + 1);
    double r = 2000.0 * std::sqrt(u / (1.0 - u)); // projected Plummer, R < 63.2k
    float x = (float)(50000.0 + r * std::cos(a)), y = (float)(50000.0 + r * std::sin(a));
    // clip to the reference domain [0, HEIGHT)
    if (!(x >= 0.f && x < 100000.f)) x = x < 0.f ? 0.f : 99999.99f;
    if (!(y >= 0.f && y < 100000.f)) y = y < 0.f ? 0.f : 99999.99f;
    s.push(mk(x, y, 0.f, 0.f, 1u));
  }
}

// Two copies of the World::new disc recipe (main.rs:316-332), continuous instead of on a grid,
// scaled x4 in radius (r in (2828, 12000)): heavy core + rotating disc with the reference's
// taper P(keep | d2) = max(0, 1 - 6e6/(1.5e7 - d2/16 + 1)) and speed law rot90cw(d)*sqrt(sqrt(M)/d2).
void scene_two_galaxy(uint64_t seed, size_t n, Sink &s) {
  const V2 c[2] = {{35000.0f, 35000.0f}, {60000.0f, 60000.0f}};
  const V2 cv[2] = {{200.0f, 250.0f}, {0.0f, 0.0f}};
  const uint32_t cm[2] = {75000000u, 750000u};
  if (n < 2) n = 2;
  size_t per[2] = {(n - 2) / 2 + ((n - 2) & 1), (n - 2) / 2};
  s.push(mk(c[0].x, c[0].y, cv[0].x, cv[0].y, cm[0]));
  s.push(mk(c[1].x, c[1].y, cv[1].x, cv[1].y, cm[1]));
  for (int g = 0; g < 2; ++g) {
    uint64_t ctr = 0;
    for (size_t i = 0; i < per[g]; ++i) {
      for (;;) {
        double d2s = 500000.0 + (9000000.0 - 500000.0) * (double)u01(seed, 5 + g, 3 * ctr); // unscaled d^2
        double keep = 1.0 - 6000000.0 / (15000000.0 - d2s + 1.0);
        double t = (double)u01(seed, 5 + g, 3 * ctr + 1);
        double a = 6.283185307179586 * (double)u01(seed, 5 + g, 3 * ctr + 2);
        ++ctr;
        if (t < keep) {
          double r = 4.0 * std::sqrt(d2s);
          V2 pos{(float)((double)c[g].x + r * std::cos(a)), (float)((double)c[g].y + r * std::sin(a))};
          float d2 = dist2(pos, c[g]);
          float k = std::sqrt(std::sqrt((float)cm[g]) / d2);
          V2 d{pos.x - c[g].x, pos.y - c[g].y};
          s.push(mk(pos.x, pos.y, d.y * k + cv[g].x, -d.x * k + cv[g].y, 1u));
          break;
        }
      }
    }
  }
}

} // namespace

extern "C" size_t nb_scene_generate(int32_t scene, uint64_t seed, size_t n_request, nb_particle *out,
                                    size_t cap) {
  Sink s{out, cap};
  switch (scene) {
  case NB_SCENE_DEFAULT: scene_default(seed, s); break;
  case NB_SCENE_UNIFORM_DISC: scene_uniform_disc(seed, n_request, s); break;
  case NB_SCENE_PLUMMER: scene_plummer(seed, n_request, s); break;
  case NB_SCENE_TWO_GALAXY: scene_two_galaxy(seed, n_request, s); break;
  default: return 0;
  }
  return s.n;
}
