// Host driver for the reference's lab1 kernel at bandwidth-relevant sizes.
//
// The reference reads its vectors from stdin as text, which makes n = 2^25 .. 2^28
// impractical through its own main().  This driver generates the inputs on the device
// and times the reference's kernel - its body is NOT in this repository: the build
// script (baseline/build_lab1_driver.sh) cuts `__global__ void kernel(...)` out of the
// untouched /root/reference/lab1/src/to_plot.cu into baseline/_ref/lab1_kernel.inc, which
// is included below.  Launch shapes: the reference's own <<<512,512>>> and a sweep.
// Output: one JSON object on stdout.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "_ref/lab1_kernel.inc"  // defines: __global__ void kernel(double*, double*, double*, int)

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e_ = (x);                                                             \
    if (e_ != cudaSuccess) {                                                          \
      fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));     \
      exit(1);                                                                        \
    }                                                                                 \
  } while (0)

__global__ void fill(double* a, double* b, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    a[i] = 1.0 + (double)(i % 1000) * 0.5;
    b[i] = 0.25 * (double)(i % 77);
  }
}
__global__ void check(const double* a, const double* b, const double* c, long long n, int* bad) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    if (c[i] != a[i] - b[i]) atomicAdd(bad, 1);
}

static float median_ms(int grid, int block, double* a, double* b, double* c, int n, int iters) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  std::vector<float> t;
  for (int it = 0; it < iters + 3; ++it) {
    CK(cudaEventRecord(e0));
    kernel<<<grid, block>>>(a, b, c, n);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (it >= 3) t.push_back(ms);  // three warm-up launches
  }
  std::sort(t.begin(), t.end());
  return t[t.size() / 2];
}

int main(int argc, char** argv) {
  const int lg = argc > 1 ? atoi(argv[1]) : 25;
  const long long n = 1ll << lg;  // the kernel indexes with int: n <= 2^30 keeps idx + offset in range
  double *a, *b, *c;
  int* bad;
  CK(cudaMalloc(&a, n * 8));
  CK(cudaMalloc(&b, n * 8));
  CK(cudaMalloc(&c, n * 8));
  CK(cudaMalloc(&bad, 4));
  CK(cudaMemset(bad, 0, 4));
  fill<<<1184, 256>>>(a, b, n);
  CK(cudaDeviceSynchronize());
  const int grids[] = {512, 148, 296, 592, 1184, 2368, 4736, 16384, 65536};
  const int blocks[] = {512, 128, 256, 1024};
  printf("{\"n\": \"2^%d\", \"bytes\": %lld, \"shapes\": [", lg, 24 * n);
  float best = 1e30f;
  int bg = 0, bb = 0;
  bool first = true;
  for (int g : grids)
    for (int bl : blocks) {
      const float ms = median_ms(g, bl, a, b, c, (int)n, 10);
      printf("%s{\"grid\": %d, \"block\": %d, \"ms\": %.5f, \"GBps\": %.1f}", first ? "" : ", ", g, bl, ms,
             24.0 * n / ms / 1e6);
      first = false;
      if (ms < best) best = ms, bg = g, bb = bl;
    }
  // the same loop, same events, around this repository's kernel (library path in argv[2])
  float own_ms = -1.f;
  if (argc > 2) {
    typedef int (*vecsub_fn)(const double*, const double*, double*, size_t, void*);
    void* lib = dlopen(argv[2], RTLD_NOW);
    vecsub_fn fn = lib ? (vecsub_fn)dlsym(lib, "b200_vecsub_f64") : nullptr;
    if (fn) {
      cudaEvent_t e0, e1;
      CK(cudaEventCreate(&e0));
      CK(cudaEventCreate(&e1));
      std::vector<float> t;
      for (int it = 0; it < 13; ++it) {
        CK(cudaEventRecord(e0));
        if (fn(a, b, c, (size_t)n, nullptr) != 0) break;
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 3) t.push_back(ms);
      }
      if (t.size() == 10) {
        std::sort(t.begin(), t.end());
        own_ms = t[5];
      }
    }
  }
  check<<<1184, 256>>>(a, b, c, n, bad);
  int h_bad = 0;
  CK(cudaMemcpy(&h_bad, bad, 4, cudaMemcpyDeviceToHost));
  printf("], \"best\": {\"grid\": %d, \"block\": %d, \"ms\": %.5f, \"GBps\": %.1f}, \"this_build_same_loop\": "
         "{\"ms\": %.5f, \"GBps\": %.1f}, \"wrong_elements\": %d}\n",
         bg, bb, best, 24.0 * n / best / 1e6, own_ms, own_ms > 0 ? 24.0 * n / own_ms / 1e6 : 0.0, h_bad);
  return h_bad != 0;
}
