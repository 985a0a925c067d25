// lab1: c = a - b over doubles.  ref: lab1/src/main.cu:22-29 (scalar grid-stride
// kernel, LDG.E.64 x2 + DADD + STG.E.64 per element).
//
// B200 design: pure HBM stream, 24 B/element.  Each thread moves 256-bit
// vectors (ld.global.nc.v4.f64, new on sm_100) with four independent vector
// pairs in flight before the first use, no L1 allocation on the way in and
// evict-first on the way out.  One CTA per 32 KB slab, balanced by the hardware.
#include <cstdlib>

#include "common.cuh"

namespace b200 {

// default shape: measured on B200 over {128,256,512} threads x {1,2,4} vectors in flight x cache
// hints on/off - all within 1.5% (6.14-6.27 TB/s at n = 2^25, 6.87-6.99 at 2^28); 512 x 2 with
// hints is the best at both sizes
constexpr int kVsThreads = 512;
constexpr int kVsUnroll = 2;  // 2 x 32 B per array per thread in flight: 1024 vectors = 32 KB per array and CTA

__device__ __forceinline__ void ld_plain_f64x4(const double* p, double (&v)[4]) {
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ void st_plain_f64x4(double* p, const double (&v)[4]) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
}

// One CTA per slab of THREADS x UNROLL 256-bit vectors of each array, no loop: every thread
// issues its 2 x UNROLL loads back to back, then subtracts and stores.  The hardware CTA
// scheduler balances the slabs over the SMs (no serialised grid-stride tail).
// HINTS: no L1 allocation on the way in, evict-first on the way out.
template <int THREADS, int UNROLL, bool HINTS>
__global__ void __launch_bounds__(THREADS)
    vecsub_v4_kernel(const double* __restrict__ a, const double* __restrict__ b,
                     double* __restrict__ c, size_t n_vec4) {
  const size_t base = (size_t)blockIdx.x * (THREADS * UNROLL) + threadIdx.x;
  double va[UNROLL][4], vb[UNROLL][4];
  auto ld = [](const double* p, double (&v)[4]) {
    if (HINTS) ld_stream_f64x4(p, v);
    else ld_plain_f64x4(p, v);
  };
  auto st = [](double* p, const double (&v)[4]) {
    if (HINTS) st_stream_f64x4(p, v);
    else st_plain_f64x4(p, v);
  };
  if (base + (UNROLL - 1) * THREADS < n_vec4) {  // full slab (all but the last CTA)
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      ld(a + 4 * (base + u * THREADS), va[u]);
      ld(b + 4 * (base + u * THREADS), vb[u]);
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
      for (int k = 0; k < 4; ++k) va[u][k] = va[u][k] - vb[u][k];
      st(c + 4 * (base + u * THREADS), va[u]);
    }
  } else {
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const size_t i = base + u * THREADS;
      if (i < n_vec4) {
        ld(a + 4 * i, va[u]);
        ld(b + 4 * i, vb[u]);
#pragma unroll
        for (int k = 0; k < 4; ++k) va[u][k] = va[u][k] - vb[u][k];
        st(c + 4 * i, va[u]);
      }
    }
  }
}

// Scalar grid-stride kernel: the tail of the vector kernel, unaligned buffers,
// and the launch-geometry sweep (any <<<grid, block>>> the user asks for).
__global__ void vecsub_scalar_kernel(const double* __restrict__ a, const double* __restrict__ b,
                                     double* __restrict__ c, size_t begin, size_t n) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    c[i] = a[i] - b[i];
}

}  // namespace b200

using namespace b200;

extern "C" int b200_vecsub_f64(const double* a, const double* b, double* c, size_t n,
                               b200_stream_t s) {
  if (n == 0) return 0;
  cudaStream_t st = as_stream(s);
  const bool aligned = ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)c)) & 31u) == 0;
  size_t n_vec4 = aligned ? n / 4 : 0;
  const int sms = device_info().sm_count;
  if (n_vec4) {
    // B200LABS_VECSUB="threads,unroll,hints" picks another shape of the kernel (tuning sweeps)
    static const int shape = [] {
      const char* e = getenv("B200LABS_VECSUB");
      int t = kVsThreads, u = kVsUnroll, h = 1;
      if (e) sscanf(e, "%d,%d,%d", &t, &u, &h);
      return t * 100 + u * 10 + h;
    }();
#define VS_CASE(T, U, H)                                                                    \
  case T * 100 + U * 10 + H:                                                                \
    vecsub_v4_kernel<T, U, H != 0><<<(unsigned)((n_vec4 + T * U - 1) / (T * U)), T, 0, st>>>(a, b, c, n_vec4); \
    break;
    switch (shape) {
      VS_CASE(256, 4, 1) VS_CASE(256, 2, 1) VS_CASE(256, 1, 1) VS_CASE(128, 4, 1) VS_CASE(128, 2, 1)
      VS_CASE(128, 1, 1) VS_CASE(256, 4, 0) VS_CASE(256, 2, 0) VS_CASE(256, 1, 0) VS_CASE(128, 4, 0)
      VS_CASE(128, 2, 0) VS_CASE(128, 1, 0) VS_CASE(512, 1, 1) VS_CASE(512, 2, 1) VS_CASE(512, 1, 0)
      default: B200_REQUIRE(false, "vecsub: unknown kernel shape %d", shape);
    }
#undef VS_CASE
    B200_LAUNCH_OK();
  }
  size_t done = n_vec4 * 4;
  if (done < n) {
    size_t rest = n - done;
    size_t want = (rest + 255) / 256;
    int grid = (int)(want < (size_t)sms * 8 ? want : (size_t)sms * 8);
    vecsub_scalar_kernel<<<grid, 256, 0, st>>>(a, b, c, done, n);
    B200_LAUNCH_OK();
  }
  return 0;
}

extern "C" int b200_vecsub_f64_geom(const double* a, const double* b, double* c, size_t n,
                                    int grid, int block, b200_stream_t s) {
  B200_REQUIRE(grid > 0 && block > 0, "vecsub_geom: grid and block must be positive (got %d, %d)",
               grid, block);
  vecsub_scalar_kernel<<<grid, block, 0, as_stream(s)>>>(a, b, c, 0, n);
  B200_LAUNCH_OK();
  return 0;
}

extern "C" int b200_preload_lab1(void) {
  if (int rc = preload_kernel(vecsub_v4_kernel<kVsThreads, kVsUnroll, true>)) return rc;
  return preload_kernel(vecsub_scalar_kernel);
}
