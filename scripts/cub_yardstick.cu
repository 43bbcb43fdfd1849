// Library yardstick: cub::DeviceRadixSort::SortPairs on the same (u64 key, u32 value) pairs and
// key width as csrc/radix_sort.cuh.  Not part of the product; built and run only by scripts/cub_yardstick.sh.
#include <cub/device/device_radix_sort.cuh>
#include <cstdio>
#include <cstdlib>
#include <vector>
int main(int argc, char **argv) {
  const size_t n = argc > 1 ? atol(argv[1]) : 1000000;
  const int bits = argc > 2 ? atoi(argv[2]) : 49;
  std::vector<unsigned long long> hk(n);
  std::vector<unsigned> hv(n);
  unsigned long long s = 88172645463325252ull;
  for (size_t i = 0; i < n; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; hk[i] = s & ((1ull << 48) - 1); hv[i] = (unsigned)i; }
  unsigned long long *k0, *k1; unsigned *v0, *v1;
  cudaMalloc(&k0, n * 8); cudaMalloc(&k1, n * 8); cudaMalloc(&v0, n * 4); cudaMalloc(&v1, n * 4);
  void *tmp = nullptr; size_t tb = 0;
  cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, v0, v1, (int)n, 0, bits);
  cudaMalloc(&tmp, tb);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e9f;
  for (int r = 0; r < 12; ++r) {
    cudaMemcpy(k0, hk.data(), n * 8, cudaMemcpyHostToDevice); cudaMemcpy(v0, hv.data(), n * 4, cudaMemcpyHostToDevice);
    cudaEventRecord(e0);
    cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, v0, v1, (int)n, 0, bits);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r > 1 && ms < best) best = ms;
  }
  printf("cub SortPairs n=%zu bits=%d: %.3f ms (best of 10)\n", n, bits, best);
  return 0;
}
