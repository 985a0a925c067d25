// lab1: c = a - b over doubles.  ref: lab1/src/main.cu:22-29 (scalar grid-stride
// kernel, LDG.E.64 x2 + DADD + STG.E.64 per element).
//
// B200 design: pure HBM stream, 24 B/element.  Each thread moves 256-bit
// vectors (ld.global.nc.v4.f64, new on sm_100) with four independent vector
// pairs in flight before the first use, no L1 allocation on the way in and
// evict-first on the way out.  One CTA per 32 KB slab, balanced by the hardware.
#include "common.cuh"

namespace b200 {

constexpr int kVsThreads = 256;
constexpr int kVsUnroll = 4;  // 4 x 32 B per array per thread in flight
constexpr int kVsPerBlock = kVsThreads * kVsUnroll;  // 1024 vectors = 32 KB per array

// One CTA per 32 KB slab of each array, no loop: every thread issues its eight
// 256-bit loads back to back, then subtracts and stores.  The hardware CTA
// scheduler balances the slabs over the SMs (no serialised grid-stride tail).
__global__ void __launch_bounds__(kVsThreads)
    vecsub_v4_kernel(const double* __restrict__ a, const double* __restrict__ b,
                     double* __restrict__ c, size_t n_vec4) {
  const size_t base = (size_t)blockIdx.x * kVsPerBlock + threadIdx.x;
  double va[kVsUnroll][4], vb[kVsUnroll][4];
  if (base + (kVsUnroll - 1) * kVsThreads < n_vec4) {  // full slab (all but the last CTA)
#pragma unroll
    for (int u = 0; u < kVsUnroll; ++u) {
      ld_stream_f64x4(a + 4 * (base + u * kVsThreads), va[u]);
      ld_stream_f64x4(b + 4 * (base + u * kVsThreads), vb[u]);
    }
#pragma unroll
    for (int u = 0; u < kVsUnroll; ++u) {
#pragma unroll
      for (int k = 0; k < 4; ++k) va[u][k] = va[u][k] - vb[u][k];
      st_stream_f64x4(c + 4 * (base + u * kVsThreads), va[u]);
    }
  } else {
#pragma unroll
    for (int u = 0; u < kVsUnroll; ++u) {
      const size_t i = base + u * kVsThreads;
      if (i < n_vec4) {
        ld_stream_f64x4(a + 4 * i, va[u]);
        ld_stream_f64x4(b + 4 * i, vb[u]);
#pragma unroll
        for (int k = 0; k < 4; ++k) va[u][k] = va[u][k] - vb[u][k];
        st_stream_f64x4(c + 4 * i, va[u]);
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
    const size_t grid = (n_vec4 + kVsPerBlock - 1) / kVsPerBlock;
    vecsub_v4_kernel<<<(unsigned)grid, kVsThreads, 0, st>>>(a, b, c, n_vec4);
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
  if (int rc = preload_kernel(vecsub_v4_kernel)) return rc;
  return preload_kernel(vecsub_scalar_kernel);
}
