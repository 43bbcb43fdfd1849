// ubench.cuh — FP32 pipe microbenchmarks: the measured denominators of the traversal / direct-sum
// roofline (MEASURED_PEAKS.json carries HBM and bf16 numbers only).
//   kind 0: FFMA   — 16 independent scalar FMA chains per lane         (2 flop per lane-instruction)
//   kind 1: FFMA2  — 8 independent packed f32x2 FMA chains per lane    (4 flop per lane-instruction)
//   kind 2: the FAST pair force (main.rs:235-253) on register-resident sources: interactions/s ceiling
//   kind 3: the same with packed f32x2 math (two sources per iteration)
#pragma once
#include "common.cuh"

namespace nb {

template <int KIND> __global__ void __launch_bounds__(256) fp32_ubench_kernel(float *out, int iters, float seed) {
  const float t = (float)threadIdx.x * 1e-6f + seed;
  if (KIND == 0) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = t + i;
    const float m = 1.0000001f, c = 1e-7f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += a[i];
    if (s == 12345.678f) out[0] = s;
  } else if (KIND == 1) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = make_float2(t + i, t - i);
    const float2 m = make_float2(1.0000001f, 0.9999999f), c = make_float2(1e-7f, 2e-7f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = __ffma2_rn(a[i], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    if (s == 12345.678f) out[0] = s;
  } else if (KIND == 2) {
    float ax = 0.f, ay = 0.f;
    const float px = t, py = 2.f * t;
    float sx = 3.f + t, sy = 5.f - t;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        pair_force<false>(px, py, sx + (float)i, sy, 1.0f, ax, ay);
      }
      sx += 1e-3f;
    }
    if (ax + ay == 12345.678f) out[0] = ax;
  } else {
    float2 ax = make_float2(0.f, 0.f), ay = ax;
    const float2 npx = make_float2(-t, -t), npy = make_float2(-2.f * t, -2.f * t);
    float2 sx = make_float2(3.f + t, 4.f + t), sy = make_float2(5.f - t, 6.f - t);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 dx = __fadd2_rn(make_float2(sx.x + (float)i, sx.y + (float)i), npx);
        const float2 dy = __fadd2_rn(sy, npy);
        float2 d2 = __ffma2_rn(dy, dy, __fmul2_rn(dx, dx));
        const float2 sum = make_float2(fabsf(dx.x) + fabsf(dy.x), fabsf(dx.y) + fabsf(dy.y));
        d2 = make_float2(fmaxf(d2.x, 0.001f), fmaxf(d2.y, 0.001f));
        const float2 den = __ffma2_rn(sum, d2, make_float2(1e-20f, 1e-20f));
        const float2 w = make_float2(rcp_approx(den.x), rcp_approx(den.y));
        ax = __ffma2_rn(dx, w, ax);
        ay = __ffma2_rn(dy, w, ay);
      }
      sx.x += 1e-3f;
    }
    if (ax.x + ay.y == 12345.678f) out[0] = ax.x;
  }
}

} // namespace nb
