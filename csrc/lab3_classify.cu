// lab3: per-pixel Mahalanobis classifier.
// ref: lab3/src/main.cu:40-76 (classify_kernel: fp64 quadratic form per pixel per
// class, strict-< argmin, class index written into the alpha byte) and
// :102-152 (class mean / covariance / inverse computed ON THE HOST).
//
// B200 design:
//  * statistics on the device as EXACT 64-bit integer moments (n, S, SS) — integer
//    sums are associative, so any sharding of the sample points over CTAs or GPUs
//    gives bit-identical parameters (the multi-GPU path all-reduces these ten
//    integers per class inside the kernel, see mgpu_lab3.cu);
//  * DIRECT classify: 4 pixels per thread (one 128-bit load/store), parameters in
//    shared memory, the reference's DFMA chain verbatim (fp64-pipe bound);
//  * LUT classify: an RGBA8 image has at most 2^24 distinct colours, so for big
//    images classify every colour once with the same DFMA chain into a 16 MiB
//    table that lives in the 126 MB L2, then the image pass is a byte gather at
//    HBM speed.  Decisions are bit-identical to DIRECT by construction.
#include <cfloat>
#include <cstdlib>

#include "common.cuh"
#include "lab3_math.cuh"

namespace b200 {


// ------------------------------------------------------------------ statistics
constexpr int kMomThreads = 256;

__global__ void __launch_bounds__(kMomThreads)
    lab3_moments_kernel(const uint32_t* __restrict__ img, int w, int64_t y0, int64_t h_local,
                        const int2* __restrict__ pts, const int* __restrict__ class_off,
                        unsigned long long* __restrict__ moments) {
  const int c = blockIdx.y;
  const int begin = class_off[c], end = class_off[c + 1];
  unsigned long long m[kNM];
#pragma unroll
  for (int k = 0; k < kNM; ++k) m[k] = 0;
  for (int i = begin + blockIdx.x * kMomThreads + threadIdx.x; i < end;
       i += gridDim.x * kMomThreads) {
    const int2 p = __ldg(pts + i);
    const int64_t yl = (int64_t)p.y - y0;
    if (yl < 0 || yl >= h_local || p.x < 0 || p.x >= w) continue;  // another band owns it
    const uint32_t px = __ldg(img + yl * w + p.x);
    const unsigned r = px & 0xff, g = (px >> 8) & 0xff, b = (px >> 16) & 0xff;
    m[0] += 1;
    m[1] += r, m[2] += g, m[3] += b;
    m[4] += r * r, m[5] += r * g, m[6] += r * b;
    m[7] += g * g, m[8] += g * b, m[9] += b * b;
  }
  __shared__ unsigned long long red[kMomThreads / 32][kNM];
#pragma unroll
  for (int k = 0; k < kNM; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m[k] += __shfl_down_sync(0xffffffffu, m[k], o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < kNM; ++k) red[warp][k] = m[k];
  }
  __syncthreads();
  if (threadIdx.x < kNM) {
    unsigned long long s = 0;
    for (int wv = 0; wv < kMomThreads / 32; ++wv) s += red[wv][threadIdx.x];
    if (s) atomicAdd(moments + (size_t)c * kNM + threadIdx.x, s);
  }
}

__global__ void lab3_finalize_kernel(const long long* __restrict__ moments, int nc,
                                     double* __restrict__ params) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < nc) lab3_finalize_class(moments + (size_t)c * kNM, params + (size_t)c * kNP);
}

// ------------------------------------------------------------------- classify
// (maha / classify_group / set_alpha / load_params_smem live in lab3_math.cuh)

constexpr int kClsThreads = 256;

// DIRECT: 4 px per thread, persistent grid-stride over 16-byte groups.
__global__ void __launch_bounds__(kClsThreads)
    lab3_classify_direct_kernel(uint32_t* __restrict__ img, int64_t n_vec, int nc,
                                const double* __restrict__ params) {
  __shared__ double sp[kMaxC * kNP];
  load_params_smem(sp, params, nc);
  const int64_t stride = (int64_t)gridDim.x * kClsThreads;
  for (int64_t i = (int64_t)blockIdx.x * kClsThreads + threadIdx.x; i < n_vec; i += stride) {
    uint4 v = *reinterpret_cast<const uint4*>(img + 4 * i);
    uint32_t px[4] = {v.x, v.y, v.z, v.w};
    int best[4];
    classify_group<4>(px, nc, sp, best);
    v.x = set_alpha(v.x, best[0]);
    v.y = set_alpha(v.y, best[1]);
    v.z = set_alpha(v.z, best[2]);
    v.w = set_alpha(v.w, best[3]);
    *reinterpret_cast<uint4*>(img + 4 * i) = v;
  }
}

// scalar: tail pixels, unaligned images and the launch-geometry sweep
__global__ void lab3_classify_scalar_kernel(uint32_t* __restrict__ img, int64_t begin, int64_t n,
                                            int nc, const double* __restrict__ params) {
  __shared__ double sp[kMaxC * kNP];
  load_params_smem(sp, params, nc);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t px[1] = {img[i]};
    int best[1];
    classify_group<1>(px, nc, sp, best);
    img[i] = set_alpha(px[0], best[0]);
  }
}

// LUT build: thread t classifies colours 4t..4t+3 (R fastest) -> one 32-bit store
__global__ void __launch_bounds__(kClsThreads)
    lab3_lut_build_kernel(uint32_t* __restrict__ lut4, int nc, const double* __restrict__ params) {
  __shared__ double sp[kMaxC * kNP];
  load_params_smem(sp, params, nc);
  const uint32_t t = blockIdx.x * kClsThreads + threadIdx.x;  // 2^22 threads
  uint32_t px[4] = {4 * t, 4 * t + 1, 4 * t + 2, 4 * t + 3};  // colour id == low 24 bits
  int best[4];
  classify_group<4>(px, nc, sp, best);
  lut4[t] = (uint32_t)(best[0] & 0xff) | ((uint32_t)(best[1] & 0xff) << 8) |
            ((uint32_t)(best[2] & 0xff) << 16) | ((uint32_t)(best[3] & 0xff) << 24);
}

// LUT build, hoisted form (see lut_build_block): one CTA per (G, 64 R values), all B
__global__ void __launch_bounds__(kLutThreads)
    lab3_lut_build_hoisted_kernel(uint8_t* __restrict__ lut, int nc, const double* __restrict__ params) {
  extern __shared__ __align__(16) uint8_t lut_smem_raw[];
  LutSmem& S = *reinterpret_cast<LutSmem*>(lut_smem_raw);
  const int g = blockIdx.x / (256 / kLutR), r0 = (blockIdx.x % (256 / kLutR)) * kLutR;
  lut_build_block(S, nc, params, g, r0, 0, 256, [&](uint32_t id, uint8_t cls) { lut[id] = cls; });
}

// LUT + cell table in one pass (see lut_build_cells_cta): one CTA per (B chunk, G cell, 64 R)
__global__ void __launch_bounds__(kLutThreads)
    lab3_lut_build2_kernel(uint8_t* __restrict__ lut, uint8_t* __restrict__ cells, int nc,
                           const double* __restrict__ params) {
  extern __shared__ __align__(16) uint8_t lut_smem_raw[];
  Lut2Smem& S = *reinterpret_cast<Lut2Smem*>(lut_smem_raw);
  const int rblock = blockIdx.x & 3, gcell = (blockIdx.x >> 2) & 63, chunk = blockIdx.x >> 8;
  lut_build_cells_cta(
      S, nc, params, gcell, rblock * kLutR, chunk * kLut2B,
      [&](uint32_t id, const uint4& v) { *reinterpret_cast<uint4*>(lut + id) = v; },
      [&](uint32_t idx, const uint4& v) {
        if (cells) *reinterpret_cast<uint4*>(cells + idx) = v;
      });
}

__device__ __forceinline__ uint32_t lut_gather(const uint8_t* p) {
#ifdef B200_LUT_GATHER_LDG
  return __ldg(p);
#else
  // the 16 MiB table lives in L2; a random colour practically never hits the 256 KB L1,
  // so do not allocate the line there
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
#endif
}

// LUT apply: 4 px per thread, byte gathers through the read-only path
__global__ void __launch_bounds__(kClsThreads)
    lab3_lut_apply_kernel(uint32_t* __restrict__ img, int64_t n_vec,
                          const uint8_t* __restrict__ lut) {
  const int64_t stride = (int64_t)gridDim.x * kClsThreads;
  for (int64_t i = (int64_t)blockIdx.x * kClsThreads + threadIdx.x; i < n_vec; i += stride) {
    uint4 v = ld_stream_u4(img + 4 * i);
    const uint32_t c0 = lut_gather(lut + (v.x & 0xffffffu)), c1 = lut_gather(lut + (v.y & 0xffffffu)),
                   c2 = lut_gather(lut + (v.z & 0xffffffu)), c3 = lut_gather(lut + (v.w & 0xffffffu));
    v.x = (v.x & 0xffffffu) | (c0 << 24);
    v.y = (v.y & 0xffffffu) | (c1 << 24);
    v.z = (v.z & 0xffffffu) | (c2 << 24);
    v.w = (v.w & 0xffffffu) | (c3 << 24);
    st_stream_u4(img + 4 * i, v);
  }
}

// ---- two-level apply --------------------------------------------------------------
// A random gather into the 16 MiB table costs one L2 sector per pixel, and that sector
// rate — not HBM — bounds the plain apply kernel (profiles/ncu_prof_lab3_lut.md).  Class
// regions are solid bodies in colour space, so most 4x4x8 cells of colours (R>>2, G>>2,
// B>>3) map to ONE class: a 2^17-entry table of those cells fits in shared memory
// (128 KB), and only pixels whose cell straddles a class boundary (entry kMixed) still
// go to the full table.  Decisions stay bit-identical: a cell is marked uniform only
// when all of its 128 full-table entries agree.
constexpr int kCoarseBits = 17;
constexpr size_t kCoarseBytes = size_t(1) << kCoarseBits;
constexpr uint32_t kMixed = 0xffu;  // never a class index (classes <= 32)
constexpr int kApplyThreads = 1024;

__device__ __forceinline__ uint32_t coarse_index(uint32_t v) {
  return ((v >> 2) & 0x3fu) | ((v >> 4) & 0xfc0u) | ((v >> 7) & 0x1f000u);
}

// thread t owns cell (rc, gc, bc) = (t & 63, (t >> 6) & 63, t >> 12): 8 x 4 words of 4 R values
__global__ void __launch_bounds__(256)
    lab3_lut_coarsen_kernel(const uint32_t* __restrict__ lut32, uint8_t* __restrict__ coarse) {
  const uint32_t t = blockIdx.x * 256 + threadIdx.x;
  const uint32_t rc = t & 63u, g0 = ((t >> 6) & 63u) * 4u, b0 = (t >> 12) * 8u;
  const uint32_t first = lut32[(b0 << 14) | (g0 << 6) | rc] & 0xffu;
  const uint32_t want = first * 0x01010101u;
  bool same = true;
#pragma unroll
  for (uint32_t b = 0; b < 8; ++b)
#pragma unroll
    for (uint32_t g = 0; g < 4; ++g) same &= lut32[((b0 + b) << 14) | ((g0 + g) << 6) | rc] == want;
  coarse[t] = (uint8_t)(same ? first : kMixed);
}

template <int U>
__global__ void __launch_bounds__(kApplyThreads, 1)
    lab3_lut_apply_coarse_kernel(uint32_t* __restrict__ img, int64_t n_vec,
                                 const uint8_t* __restrict__ lut,
                                 const uint8_t* __restrict__ coarse) {
  extern __shared__ __align__(16) uint8_t sc[];
  for (int i = threadIdx.x; i < (int)(kCoarseBytes / 16); i += kApplyThreads)
    reinterpret_cast<uint4*>(sc)[i] = __ldg(reinterpret_cast<const uint4*>(coarse) + i);
  __syncthreads();
  auto with_class = [&](uint32_t v) {
    uint32_t c = sc[coarse_index(v)];
    if (c == kMixed) c = lut_gather(lut + (v & 0xffffffu));
    return (v & 0xffffffu) | (c << 24);
  };
  auto apply4 = [&](uint4 v) {
    v.x = with_class(v.x), v.y = with_class(v.y), v.z = with_class(v.z), v.w = with_class(v.w);
    return v;
  };
  const int64_t stride = (int64_t)gridDim.x * kApplyThreads;
  int64_t i = (int64_t)blockIdx.x * kApplyThreads + threadIdx.x;
  for (; i + (U - 1) * stride < n_vec; i += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = ld_stream_u4(img + 4 * (i + u * stride));
#pragma unroll
    for (int u = 0; u < U; ++u) st_stream_u4(img + 4 * (i + u * stride), apply4(v[u]));
  }
  for (; i < n_vec; i += stride) st_stream_u4(img + 4 * i, apply4(ld_stream_u4(img + 4 * i)));
}

__global__ void lab3_lut_apply_scalar_kernel(uint32_t* __restrict__ img, int64_t begin, int64_t n,
                                             const uint8_t* __restrict__ lut) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t v = img[i];
    img[i] = (v & 0xffffffu) | ((uint32_t)__ldg(lut + (v & 0xffffffu)) << 24);
  }
}

constexpr size_t kLutBytes = size_t(1) << 24;
constexpr int64_t kLutMinPixels = int64_t(1) << 25;

// Image pass against a complete table.  `coarse` (kCoarseBytes, may be null) is filled
// here and enables the two-level kernel; B200LABS_LAB3_COARSE=0 forces the plain one.
static int run_lut_apply(uint32_t* img, int64_t npix, const uint8_t* lut, uint8_t* coarse,
                         cudaStream_t st, bool coarse_ready = false) {
  const int sms = device_info().sm_count;
  const bool aligned = (((uintptr_t)img) & 15u) == 0;
  const int64_t n_vec = aligned ? npix / 4 : 0;
  const int64_t tail = n_vec * 4;
  auto grid_for = [&](int64_t items, int threads, int per_sm) {
    int64_t want = (items + threads - 1) / threads;
    int64_t cap = (int64_t)sms * per_sm;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
  };
  static const bool coarse_off = [] {
    const char* e = getenv("B200LABS_LAB3_COARSE");
    return e && !strcmp(e, "0");
  }();
  // below ~2^22 pixels loading 128 KB per SM costs more than the gathers it saves
  const bool two_level = coarse && !coarse_off && (((uintptr_t)coarse) & 15u) == 0 &&
                         n_vec >= (int64_t(1) << 20);
  if (n_vec && two_level) {
    static PerDeviceOnce attr_once;
    if (int rc = attr_once.run("lab3", [&]() -> int {
      B200_CUDA_OK(cudaFuncSetAttribute(lab3_lut_apply_coarse_kernel<4>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)kCoarseBytes));
          return 0;
        }))
      return rc;
    if (!coarse_ready) {  // the table came without its cell table: derive it
      lab3_lut_coarsen_kernel<<<(unsigned)(kCoarseBytes / 256), 256, 0, st>>>((const uint32_t*)lut,
                                                                             coarse);
      B200_LAUNCH_OK();
    }
    lab3_lut_apply_coarse_kernel<4><<<grid_for(n_vec, kApplyThreads, 1), kApplyThreads,
                                      kCoarseBytes, st>>>(img, n_vec, lut, coarse);
    B200_LAUNCH_OK();
  } else if (n_vec) {
    lab3_lut_apply_kernel<<<grid_for(n_vec, kClsThreads, 8), kClsThreads, 0, st>>>(img, n_vec, lut);
    B200_LAUNCH_OK();
  }
  if (tail < npix) {
    lab3_lut_apply_scalar_kernel<<<grid_for(npix - tail, kClsThreads, 8), kClsThreads, 0, st>>>(
        img, tail, npix, lut);
    B200_LAUNCH_OK();
  }
  return 0;
}

static int pick_variant(int64_t npix, int variant, size_t scratch_bytes) {
  if (variant != B200_CLASSIFY_AUTO) return variant;
  return (npix >= kLutMinPixels && scratch_bytes >= kLutBytes) ? B200_CLASSIFY_LUT
                                                               : B200_CLASSIFY_DIRECT;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_lab3_moments(const void* img, int w, int64_t y0, int64_t h_local,
                                 const int* pts, const int* class_off, int nc, long long* moments,
                                 b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kMaxC, "lab3: class count %d outside 1..%d", nc, kMaxC);
  // class sizes live on the device; size the grid for the documented maximum
  // (2^19 points per class) — blocks past a class's end exit immediately.
  dim3 grid(64, nc);
  lab3_moments_kernel<<<grid, kMomThreads, 0, as_stream(s)>>>(
      (const uint32_t*)img, w, y0, h_local, (const int2*)pts, class_off,
      (unsigned long long*)moments);
  B200_LAUNCH_OK();
  return 0;
}

extern "C" int b200_lab3_finalize(const long long* moments, int nc, double* params,
                                  b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kMaxC, "lab3: class count %d outside 1..%d", nc, kMaxC);
  lab3_finalize_kernel<<<1, 32, 0, as_stream(s)>>>(moments, nc, params);
  B200_LAUNCH_OK();
  return 0;
}

extern "C" size_t b200_lab3_classify_scratch_bytes(int64_t npix, int variant) {
  // full table + the coarse table of the two-level apply (kLutBytes alone still works:
  // the image pass then uses the one-level kernel)
  if (variant == B200_CLASSIFY_LUT) return kLutBytes + kCoarseBytes;
  if (variant == B200_CLASSIFY_AUTO && npix >= kLutMinPixels) return kLutBytes + kCoarseBytes;
  return 0;
}

extern "C" int b200_lab3_classify(void* img, int64_t npix, int nc, const double* params,
                                  int variant, void* scratch, size_t scratch_bytes,
                                  b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kMaxC, "lab3: class count %d outside 1..%d", nc, kMaxC);
  if (npix <= 0) return 0;
  cudaStream_t st = as_stream(s);
  if (!scratch) scratch_bytes = 0;
  variant = pick_variant(npix, variant, scratch_bytes);
  const int sms = device_info().sm_count;
  const bool aligned = (((uintptr_t)img) & 15u) == 0;
  const int64_t n_vec = aligned ? npix / 4 : 0;
  const int64_t tail = n_vec * 4;
  auto grid_for = [&](int64_t items, int per_sm) {
    int64_t want = (items + kClsThreads - 1) / kClsThreads;
    int64_t cap = (int64_t)sms * per_sm;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
  };
  if (variant == B200_CLASSIFY_LUT) {
    B200_REQUIRE(scratch_bytes >= kLutBytes, "lab3: LUT variant needs %zu bytes of scratch",
                 kLutBytes);
    // B200LABS_LAB3_LUT_BUILD=plain | hoisted select the older builds (table only, cells derived
    // by a separate pass); default: table and cell table from one kernel
    static const int build_mode = [] {
      const char* e = getenv("B200LABS_LAB3_LUT_BUILD");
      return !e ? 2 : !strcmp(e, "plain") ? 0 : !strcmp(e, "hoisted") ? 1 : 2;
    }();
    uint8_t* coarse = scratch_bytes >= kLutBytes + kCoarseBytes ? (uint8_t*)scratch + kLutBytes : nullptr;
    static PerDeviceOnce attr_once;
    if (int rc = attr_once.run("lab3", [&]() -> int {
          B200_CUDA_OK(cudaFuncSetAttribute(lab3_lut_build_hoisted_kernel,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LutSmem)));
          B200_CUDA_OK(cudaFuncSetAttribute(lab3_lut_build2_kernel,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Lut2Smem)));
          return 0;
        }))
      return rc;
    if (build_mode == 0) {
      lab3_lut_build_kernel<<<(1 << 22) / kClsThreads, kClsThreads, 0, st>>>((uint32_t*)scratch, nc,
                                                                             params);
    } else if (build_mode == 1) {
      lab3_lut_build_hoisted_kernel<<<256 * (256 / kLutR), kLutThreads, sizeof(LutSmem), st>>>(
          (uint8_t*)scratch, nc, params);
    } else {
      lab3_lut_build2_kernel<<<(256 / kLut2B) * 64 * 4, kLutThreads, sizeof(Lut2Smem), st>>>(
          (uint8_t*)scratch, coarse, nc, params);
    }
    B200_LAUNCH_OK();
    return run_lut_apply((uint32_t*)img, npix, (const uint8_t*)scratch, coarse, st, build_mode == 2);
  }
  B200_REQUIRE(variant == B200_CLASSIFY_DIRECT, "lab3: unknown classify variant %d", variant);
  if (n_vec) {
    lab3_classify_direct_kernel<<<grid_for(n_vec, 4), kClsThreads, 0, st>>>((uint32_t*)img, n_vec,
                                                                            nc, params);
    B200_LAUNCH_OK();
  }
  if (tail < npix) {
    lab3_classify_scalar_kernel<<<grid_for(npix - tail, 4), kClsThreads, 0, st>>>(
        (uint32_t*)img, tail, npix, nc, params);
    B200_LAUNCH_OK();
  }
  return 0;
}

extern "C" int b200_lab3_classify_geom(void* img, int64_t npix, int nc, const double* params,
                                       int blocks, int threads, b200_stream_t s) {
  B200_REQUIRE(nc > 0 && nc <= kMaxC, "lab3: class count %d outside 1..%d", nc, kMaxC);
  B200_REQUIRE(blocks > 0 && threads > 0, "lab3: launch dims must be positive (%d, %d)", blocks,
               threads);
  lab3_classify_scalar_kernel<<<blocks, threads, 0, as_stream(s)>>>((uint32_t*)img, 0, npix, nc,
                                                                    params);
  B200_LAUNCH_OK();
  return 0;
}

// Image pass only, against a colour table that is already complete (built by
// b200_lab3_classify's LUT variant or cooperatively by b200_lab3_lut_build_allgather).
extern "C" int b200_lab3_apply_lut(void* img, int64_t npix, const void* lut, void* coarse_scratch,
                                   b200_stream_t s) {
  if (npix <= 0) return 0;
  return run_lut_apply((uint32_t*)img, npix, (const uint8_t*)lut, (uint8_t*)coarse_scratch,
                       as_stream(s));
}

extern "C" size_t b200_lab3_coarse_bytes(void) { return kCoarseBytes; }

// Image pass against a table buffer that already holds its cell table behind the colour
// table (b200_lab3_lut_build2_allgather): no derivation pass.
extern "C" int b200_lab3_apply_lut2(void* img, int64_t npix, const void* lut_and_cells, b200_stream_t s) {
  if (npix <= 0) return 0;
  const uint8_t* lut = (const uint8_t*)lut_and_cells;
  return run_lut_apply((uint32_t*)img, npix, lut, const_cast<uint8_t*>(lut) + kLutBytes, as_stream(s), true);
}

extern "C" int b200_preload_lab3(void) {
  if (int rc = preload_kernel(lab3_moments_kernel)) return rc;
  if (int rc = preload_kernel(lab3_finalize_kernel)) return rc;
  if (int rc = preload_kernel(lab3_classify_direct_kernel)) return rc;
  if (int rc = preload_kernel(lab3_classify_scalar_kernel)) return rc;
  if (int rc = preload_kernel(lab3_lut_build_hoisted_kernel)) return rc;
  if (int rc = preload_kernel(lab3_lut_build2_kernel)) return rc;
  if (int rc = preload_kernel(lab3_lut_apply_kernel)) return rc;
  if (int rc = preload_kernel(lab3_lut_coarsen_kernel)) return rc;
  if (int rc = preload_kernel(lab3_lut_apply_coarse_kernel<4>)) return rc;
  return preload_kernel(lab3_lut_apply_scalar_kernel);
}
