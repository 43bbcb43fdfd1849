// morton.cuh — K1: bounds reduce + Morton key generation (HBM-bound, 128-bit vector loads of the
// SoA position arrays).
#pragma once
#include "common.cuh"

namespace nb {

// Morton key = the child indices QuadTree::insert would take level by level
// (reference src/quad_tree.rs:172-194): half = height/2; mid = offset + half; bit = coord > mid;
// child offset = mid (bit set) or offset.  Same f32 midline chain as the reference, so the key
// is the literal insertion path to any depth; child = (north<<1) + west (quad_tree.rs:179) puts
// the y bit above the x bit.
// LEVELS < kQuadLevels: only the top LEVELS levels of the path, the bits below are zero.  The build asks for as many
// levels as its top-digit sort covers (radix_sort.cuh): a leaf that would need more is caught by tree_count_kernel
// (equal truncated keys share all 24 levels) and the step is redone with the full chain.
template <int LEVELS = kQuadLevels>
__device__ __forceinline__ unsigned long long morton_chain(float px, float py, float x0, float y0, float h) {
  float ox = x0, oy = y0, hh = h;
  unsigned long long key = 0;
#pragma unroll
  for (int l = 0; l < LEVELS; ++l) {
    hh = __fmul_rn(hh, 0.5f); // == hh / 2.0 exactly
    float mx = __fadd_rn(ox, hh), my = __fadd_rn(oy, hh);
    bool bx = px > mx, by = py > my;
    key = (key << 2) | ((unsigned long long)by << 1) | (unsigned long long)bx;
    ox = bx ? mx : ox;
    oy = by ? my : oy;
  }
  return key << (2 * (kQuadLevels - LEVELS));
}

// The caller-side filter the reference implies for its quadtree (within_bounds, main.rs:223-226).
__device__ __forceinline__ bool in_domain(float px, float py, float x0, float y0, float h) {
  return py < __fadd_rn(y0, h) && px < __fadd_rn(x0, h) && py >= y0 && px >= x0;
}

template <int LEVELS = kQuadLevels>
__global__ void __launch_bounds__(256) morton_kernel(const float *__restrict__ x, const float *__restrict__ y,
                                                     const Meta *__restrict__ meta, uint32_t n, int quad_mode,
                                                     unsigned long long *__restrict__ keys,
                                                     uint32_t *__restrict__ vals) {
  const float x0 = meta->x0, y0 = meta->y0, h = meta->h;
  uint32_t i4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u;
  if (i4 >= n) return;
  float px[4], py[4];
  if (i4 + 3 < n) {
    float4 vx = *reinterpret_cast<const float4 *>(x + i4);
    float4 vy = *reinterpret_cast<const float4 *>(y + i4);
    px[0] = vx.x; px[1] = vx.y; px[2] = vx.z; px[3] = vx.w;
    py[0] = vy.x; py[1] = vy.y; py[2] = vy.z; py[3] = vy.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      px[k] = (i4 + k < n) ? x[i4 + k] : 0.f;
      py[k] = (i4 + k < n) ? y[i4 + k] : 0.f;
    }
  }
  unsigned long long kk[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    kk[k] = morton_chain<LEVELS>(px[k], py[k], x0, y0, h);
    if (quad_mode && !in_domain(px[k], py[k], x0, y0, h)) kk[k] = kKeyOut;
  }
  if (i4 + 3 < n) {
    reinterpret_cast<ulonglong2 *>(keys + i4)[0] = make_ulonglong2(kk[0], kk[1]);
    reinterpret_cast<ulonglong2 *>(keys + i4)[1] = make_ulonglong2(kk[2], kk[3]);
    *reinterpret_cast<uint4 *>(vals + i4) = make_uint4(i4, i4 + 1, i4 + 2, i4 + 3);
  } else {
    for (int k = 0; k < 4 && i4 + k < n; ++k) {
      keys[i4 + k] = kk[k];
      vals[i4 + k] = i4 + k;
    }
  }
}

// ---- bounds reduce (LBVH domain_auto): exact min/max, order independent -------------------------
__global__ void __launch_bounds__(256) bounds_partial_kernel(const float *__restrict__ x, const float *__restrict__ y,
                                                             uint32_t n, float4 *__restrict__ partial) {
  float mnx = INFINITY, mny = INFINITY, mxx = -INFINITY, mxy = -INFINITY;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float a = x[i], b = y[i];
    mnx = fminf(mnx, a); mxx = fmaxf(mxx, a);
    mny = fminf(mny, b); mxy = fmaxf(mxy, b);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o));
    mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
    mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
  }
  __shared__ float4 s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = make_float4(mnx, mny, mxx, mxy);
  __syncthreads();
  if (threadIdx.x == 0) {
    float4 r = s[0];
    for (int w = 1; w < 8; ++w) {
      r.x = fminf(r.x, s[w].x); r.y = fminf(r.y, s[w].y);
      r.z = fmaxf(r.z, s[w].z); r.w = fmaxf(r.w, s[w].w);
    }
    partial[blockIdx.x] = r;
  }
}

__global__ void bounds_final_kernel(const float4 *__restrict__ partial, int np, Meta *meta) {
  float4 r = make_float4(INFINITY, INFINITY, -INFINITY, -INFINITY);
  for (int i = threadIdx.x; i < np; i += 32) {
    float4 p = partial[i];
    r.x = fminf(r.x, p.x); r.y = fminf(r.y, p.y);
    r.z = fmaxf(r.z, p.z); r.w = fmaxf(r.w, p.w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    r.x = fminf(r.x, __shfl_xor_sync(0xffffffffu, r.x, o));
    r.y = fminf(r.y, __shfl_xor_sync(0xffffffffu, r.y, o));
    r.z = fmaxf(r.z, __shfl_xor_sync(0xffffffffu, r.z, o));
    r.w = fmaxf(r.w, __shfl_xor_sync(0xffffffffu, r.w, o));
  }
  if (threadIdx.x == 0) {
    float w = __fsub_rn(r.z, r.x), hgt = __fsub_rn(r.w, r.y);
    meta->x0 = r.x;
    meta->y0 = r.y;
    meta->h = w > hgt ? w : hgt;
  }
}

} // namespace nb
