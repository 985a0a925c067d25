// Yardsticks for the lab5 sort, NOT part of the library (tools/ only, never linked into it):
//   * cub::DeviceRadixSort::SortKeys - the vendor's keys-only radix sort on this GPU;
//   * std::sort on the host (one thread) - the CPU baseline SURVEY section 8 names.
// Usage: sort_yardstick <log2 n> [cpu_log2_n]   -> one JSON object on stdout.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cub/device/device_radix_sort.cuh>
#include <random>
#include <vector>

#define CK(x)                                                                     \
  do {                                                                            \
    cudaError_t e_ = (x);                                                         \
    if (e_ != cudaSuccess) {                                                      \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(1);                                                                    \
    }                                                                             \
  } while (0)

__global__ void fill_random(int* k, size_t n, unsigned seed) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    unsigned long long x = (i + 1) * 0x9E3779B97F4A7C15ull + seed;
    x ^= x >> 30, x *= 0xBF58476D1CE4E5B9ull, x ^= x >> 27, x *= 0x94D049BB133111EBull, x ^= x >> 31;
    k[i] = (int)x;
  }
}

int main(int argc, char** argv) {
  const int lg = argc > 1 ? atoi(argv[1]) : 26;
  const int cpu_lg = argc > 2 ? atoi(argv[2]) : 24;
  const size_t n = size_t(1) << lg;
  int *in, *out;
  CK(cudaMalloc(&in, n * 4));
  CK(cudaMalloc(&out, n * 4));
  fill_random<<<1184, 256>>>(in, n, 12345u);
  size_t tmp_bytes = 0;
  CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, in, out, n));
  void* tmp;
  CK(cudaMalloc(&tmp, tmp_bytes));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  std::vector<float> t;
  for (int it = 0; it < 13; ++it) {
    CK(cudaEventRecord(e0));
    CK(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, in, out, n));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (it >= 3) t.push_back(ms);
  }
  std::sort(t.begin(), t.end());
  const float cub_ms = t[t.size() / 2];
  // CPU baseline: std::sort of 2^cpu_lg keys, one thread
  const size_t m = size_t(1) << cpu_lg;
  std::vector<int> h(m);
  CK(cudaMemcpy(h.data(), in, m * 4, cudaMemcpyDeviceToHost));
  const auto c0 = std::chrono::steady_clock::now();
  std::sort(h.begin(), h.end());
  const double cpu_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - c0).count();
  printf("{\"n\": \"2^%d\", \"cub_sortkeys_ms\": %.4f, \"cub_Gkeys_s\": %.1f, \"cub_temp_bytes\": %zu, "
         "\"cpu_std_sort\": {\"n\": \"2^%d\", \"ms\": %.1f, \"Mkeys_s\": %.1f}}\n",
         lg, cub_ms, n / cub_ms / 1e6, tmp_bytes, cpu_lg, cpu_ms, m / cpu_ms / 1e3);
  return 0;
}
