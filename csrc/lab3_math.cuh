// lab3 device math shared by the single-GPU and the fused multi-GPU kernels.
#pragma once
#include "common.cuh"

#include <cfloat>

namespace b200 {

constexpr int kNP = B200_LAB3_PARAMS;
constexpr int kNM = B200_LAB3_MOMENTS;
constexpr int kMaxC = B200_LAB3_MAX_CLASSES;

// moments -> mean + inverse covariance.  Same formulas as b200_cpu_lab3_finalize
// (oracle.cpp); every product/sum uses an explicit-rounding intrinsic so nothing
// is contracted and host and device agree bit for bit.
__device__ __forceinline__ void lab3_finalize_class(const long long* m, double* out) {
  const long long n = m[0];
  const double dn = (double)n;
  out[0] = __ddiv_rn((double)m[1], dn);
  out[1] = __ddiv_rn((double)m[2], dn);
  out[2] = __ddiv_rn((double)m[3], dn);
  const long long S[3] = {m[1], m[2], m[3]};
  const int pair[3][3] = {{4, 5, 6}, {5, 7, 8}, {6, 8, 9}};
  const double denom = (double)(n * (n - 1));
  double cov[3][3];
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) {
      long long num = n * m[pair[a][b]] - S[a] * S[b];
      cov[a][b] = __ddiv_rn((double)num, denom);
    }
  auto mul = [](double x, double y) { return __dmul_rn(x, y); };
  auto sub = [](double x, double y) { return __dsub_rn(x, y); };
  double det =
      __dadd_rn(sub(mul(cov[0][0], sub(mul(cov[1][1], cov[2][2]), mul(cov[2][1], cov[1][2]))),
                    mul(cov[0][1], sub(mul(cov[1][0], cov[2][2]), mul(cov[1][2], cov[2][0])))),
                mul(cov[0][2], sub(mul(cov[1][0], cov[2][1]), mul(cov[1][1], cov[2][0]))));
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) {
      int b1 = (b + 1) % 3, b2 = (b + 2) % 3, a1 = (a + 1) % 3, a2 = (a + 2) % 3;
      out[3 + a * 3 + b] =
          __ddiv_rn(sub(mul(cov[b1][a1], cov[b2][a2]), mul(cov[b1][a2], cov[b2][a1])), det);
    }
}


// distance of colour (r,g,b) to one class: the reference's operation chain
//   temp[i] = fma(d2,m[2][i], fma(d1,m[1][i], fma(d0,m[0][i], 0)))
//   dist    = fma(d2,temp[2], fma(d1,temp[1], fma(d0,temp[0], 0)))
__device__ __forceinline__ double maha(double r, double g, double b, const double* p) {
  const double d0 = __dsub_rn(r, p[0]), d1 = __dsub_rn(g, p[1]), d2 = __dsub_rn(b, p[2]);
  double t0 = __fma_rn(d0, p[3], 0.0), t1 = __fma_rn(d0, p[4], 0.0), t2 = __fma_rn(d0, p[5], 0.0);
  t0 = __fma_rn(d1, p[6], t0), t1 = __fma_rn(d1, p[7], t1), t2 = __fma_rn(d1, p[8], t2);
  t0 = __fma_rn(d2, p[9], t0), t1 = __fma_rn(d2, p[10], t1), t2 = __fma_rn(d2, p[11], t2);
  double dist = __fma_rn(d0, t0, 0.0);
  dist = __fma_rn(d1, t1, dist);
  return __fma_rn(d2, t2, dist);
}

template <int NPX>
__device__ __forceinline__ void classify_group(const uint32_t (&px)[NPX], int nc,
                                               const double* sp, int (&best)[NPX]) {
  double r[NPX], g[NPX], b[NPX], mind[NPX];
#pragma unroll
  for (int k = 0; k < NPX; ++k) {
    r[k] = (double)(px[k] & 0xff);
    g[k] = (double)((px[k] >> 8) & 0xff);
    b[k] = (double)((px[k] >> 16) & 0xff);
    mind[k] = DBL_MAX;
    best[k] = -1;
  }
  for (int c = 0; c < nc; ++c) {
    double p[kNP];
#pragma unroll
    for (int j = 0; j < kNP; ++j) p[j] = sp[c * kNP + j];
#pragma unroll
    for (int k = 0; k < NPX; ++k) {
      const double d = maha(r[k], g[k], b[k], p);
      if (d < mind[k]) {
        mind[k] = d;
        best[k] = c;
      }
    }
  }
}

__device__ __forceinline__ uint32_t set_alpha(uint32_t px, int cls) {
  return (px & 0x00ffffffu) | ((uint32_t)(cls & 0xff) << 24);
}

__device__ __forceinline__ void load_params_smem(double* sp, const double* params, int nc) {
  for (int i = threadIdx.x; i < nc * kNP; i += blockDim.x) sp[i] = params[i];
  __syncthreads();
}


// ---- colour-table builder with the class-invariant part of the chain hoisted ----
// For a fixed (R, G) the two inner FMAs of every temp[k] do not depend on B:
//   u_k = fma(d1, m[1][k], fma(d0, m[0][k], 0))           (6 FMA + 2 subtractions)
// so a CTA computes u_k once per (class, R) for its G, keeps them in shared memory,
// and each colour then costs d2 = B - mu2, temp[k] = fma(d2, m[2][k], u_k) and the
// three-FMA distance: 7 fp64 operations + compare instead of 16 — the SAME operation
// sequence per colour as maha(), hence bit-identical decisions.
constexpr int kLutR = 64;        // R values per CTA (32 with 4 CTAs/SM and 64 registers measured 5% slower)
constexpr int kLutBGroup = 8;    // B values per thread task
constexpr int kLutThreads = 256;
struct LutSmem {
  double u[kMaxC][3][kLutR];
  double d0[kMaxC][kLutR];
  double d1[kMaxC];
  double sp[kMaxC * kNP];
};

// Classifies the colours {R in [r0, r0+64), G = g, B in [b_begin, b_end)} (b range a
// multiple of 8) and hands every (colour id, class) to `emit`.
template <typename Emit>
__device__ __forceinline__ void lut_build_block(LutSmem& S, int nc, const double* __restrict__ params,
                                                int g, int r0, int b_begin, int b_end, Emit emit) {
  const int tid = threadIdx.x;
  for (int i = tid; i < nc * kNP; i += kLutThreads) S.sp[i] = params[i];
  __syncthreads();
  for (int i = tid; i < nc * kLutR; i += kLutThreads) {
    const int c = i / kLutR, r = i - c * kLutR;
    const double* p = S.sp + c * kNP;
    const double d0 = __dsub_rn((double)(r0 + r), p[0]);
    const double d1 = __dsub_rn((double)g, p[1]);
    S.d0[c][r] = d0;
    if (r == 0) S.d1[c] = d1;
#pragma unroll
    for (int k = 0; k < 3; ++k) S.u[c][k][r] = __fma_rn(d1, p[6 + k], __fma_rn(d0, p[3 + k], 0.0));
  }
  __syncthreads();
  const int n_groups = (b_end - b_begin) / kLutBGroup;
  for (int t = tid; t < n_groups * kLutR; t += kLutThreads) {
    const int r = t & (kLutR - 1), b0 = b_begin + (t / kLutR) * kLutBGroup;
    double mind[kLutBGroup];
    int best[kLutBGroup];
#pragma unroll
    for (int j = 0; j < kLutBGroup; ++j) mind[j] = DBL_MAX, best[j] = -1;
    for (int c = 0; c < nc; ++c) {
      const double* p = S.sp + c * kNP;
      const double u0 = S.u[c][0][r], u1 = S.u[c][1][r], u2 = S.u[c][2][r];
      const double d0 = S.d0[c][r], d1 = S.d1[c];
      const double mu2 = p[2], m20 = p[9], m21 = p[10], m22 = p[11];
#pragma unroll
      for (int j = 0; j < kLutBGroup; ++j) {
        const double d2 = __dsub_rn((double)(b0 + j), mu2);
        const double t0 = __fma_rn(d2, m20, u0), t1 = __fma_rn(d2, m21, u1), t2 = __fma_rn(d2, m22, u2);
        double dist = __fma_rn(d0, t0, 0.0);
        dist = __fma_rn(d1, t1, dist);
        dist = __fma_rn(d2, t2, dist);
        if (dist < mind[j]) mind[j] = dist, best[j] = c;
      }
    }
#pragma unroll
    for (int j = 0; j < kLutBGroup; ++j)
      emit(((uint32_t)(b0 + j) << 16) | ((uint32_t)g << 8) | (uint32_t)(r0 + r), (uint8_t)best[j]);
  }
}

// ---- colour table AND cell table in one pass ---------------------------------------------
// A CTA owns a block of colours made of WHOLE 4x4x8 cells: 64 R values x 4 G values (one G
// cell) x 32 B values.  It classifies them G by G with the hoisted chain above (the B - mu2
// terms of all classes come from a small shared table: 7 instead of 8 fp64-pipe operations
// per colour and class), keeps the 8192 class bytes in shared memory, derives the 64 cell
// bytes from them (uniform class, or kCellMixed) and hands everything out as 16-BYTE pieces:
// 512 pieces of 16 consecutive R values and 4 pieces of 16 cells.  One piece = one 128-bit
// store - or one multimem.st that the NVSwitch replicates into every GPU's table - where the
// older build wrote single bytes to every peer.  No coarsening pass over the finished table.
constexpr int kLut2B = 32;              // B values per CTA (four cells deep)
constexpr uint32_t kCellMixed = 0xffu;  // never a class index (classes <= 32)
struct Lut2Smem {
  double u[kMaxC][3][kLutR];  // hoisted terms of the current G
  double d2[kMaxC][kLut2B];   // B - mu2 per class
  double sp[kMaxC * kNP];
  alignas(16) uint8_t stage[4][kLut2B][kLutR];  // class bytes: 4 G x 32 B x 64 R
  alignas(16) uint8_t cells[kLut2B / 8][kLutR / 4];
};

// The 256 B values are eight chunks of kLut2B; rank r of `world` owns chunks r, r + world, ...
__host__ __device__ inline int lut2_chunks_of(int rank, int world) {
  return (256 / kLut2B - rank + world - 1) / world;
}

// emit_colours(colour id of the first of 16 consecutive R values, 16 class bytes)
// emit_cells(cell index of the first of 16 consecutive R cells, 16 cell bytes)
template <typename EmitColours, typename EmitCells>
__device__ __forceinline__ void lut_build_cells_cta(Lut2Smem& S, int nc, const double* __restrict__ params,
                                                    int gcell, int r0, int b_base, EmitColours emit_colours,
                                                    EmitCells emit_cells) {
  static_assert(kLutThreads == 256 && kLutR == 64 && kLut2B == 32 && kLutBGroup == 8,
                "task layout below: 256 threads = 64 R x 4 groups of 8 B");
  const int tid = threadIdx.x;
  for (int i = tid; i < nc * kNP; i += kLutThreads) S.sp[i] = params[i];
  __syncthreads();
  for (int i = tid; i < nc * kLut2B; i += kLutThreads) {
    const int c = i / kLut2B, b = i - c * kLut2B;
    S.d2[c][b] = __dsub_rn((double)(b_base + b), S.sp[c * kNP + 2]);
  }
  const int r = tid & (kLutR - 1), bg = tid / kLutR;  // this thread: one R, eight B values
  for (int gi = 0; gi < 4; ++gi) {
    const int g = gcell * 4 + gi;
    __syncthreads();  // the previous G's classify loop is done with S.u (and S.d2 is complete)
    for (int i = tid; i < nc * kLutR; i += kLutThreads) {
      const int c = i / kLutR, rr = i - c * kLutR;
      const double* p = S.sp + c * kNP;
      const double d0 = __dsub_rn((double)(r0 + rr), p[0]);
      const double d1 = __dsub_rn((double)g, p[1]);
#pragma unroll
      for (int k = 0; k < 3; ++k) S.u[c][k][rr] = __fma_rn(d1, p[6 + k], __fma_rn(d0, p[3 + k], 0.0));
    }
    __syncthreads();
    double mind[kLutBGroup];
    int best[kLutBGroup];
#pragma unroll
    for (int j = 0; j < kLutBGroup; ++j) mind[j] = DBL_MAX, best[j] = -1;
    for (int c = 0; c < nc; ++c) {
      const double* p = S.sp + c * kNP;
      const double u0 = S.u[c][0][r], u1 = S.u[c][1][r], u2 = S.u[c][2][r];
      const double d0 = __dsub_rn((double)(r0 + r), p[0]), d1 = __dsub_rn((double)g, p[1]);
      const double m20 = p[9], m21 = p[10], m22 = p[11];
#pragma unroll
      for (int j = 0; j < kLutBGroup; ++j) {
        const double d2 = S.d2[c][bg * kLutBGroup + j];
        const double t0 = __fma_rn(d2, m20, u0), t1 = __fma_rn(d2, m21, u1), t2 = __fma_rn(d2, m22, u2);
        double dist = __fma_rn(d0, t0, 0.0);
        dist = __fma_rn(d1, t1, dist);
        dist = __fma_rn(d2, t2, dist);
        if (dist < mind[j]) mind[j] = dist, best[j] = c;
      }
    }
#pragma unroll
    for (int j = 0; j < kLutBGroup; ++j) S.stage[gi][bg * kLutBGroup + j][r] = (uint8_t)best[j];
  }
  __syncthreads();
  // ---- 512 pieces of 16 class bytes (16 consecutive R values of one (G, B))
  for (int v = tid; v < 4 * kLut2B * (kLutR / 16); v += kLutThreads) {
    const int gi = v >> 7, b = (v >> 2) & (kLut2B - 1), rq = v & 3;
    const uint4 val = *reinterpret_cast<const uint4*>(&S.stage[gi][b][rq * 16]);
    emit_colours(((uint32_t)(b_base + b) << 16) | ((uint32_t)(gcell * 4 + gi) << 8) | (uint32_t)(r0 + rq * 16), val);
  }
  // ---- 64 cells: uniform class of their 4 R x 4 G x 8 B colours, or "mixed"
  if (tid < (kLut2B / 8) * (kLutR / 4)) {
    const int rc = tid & (kLutR / 4 - 1), bc = tid / (kLutR / 4);
    const uint32_t first = S.stage[0][bc * 8][rc * 4];
    const uint32_t want = first * 0x01010101u;
    bool same = true;
#pragma unroll
    for (int gi = 0; gi < 4; ++gi)
#pragma unroll
      for (int b = 0; b < 8; ++b)
        same &= *reinterpret_cast<const uint32_t*>(&S.stage[gi][bc * 8 + b][rc * 4]) == want;
    S.cells[bc][rc] = (uint8_t)(same ? first : kCellMixed);
  }
  __syncthreads();
  if (tid < kLut2B / 8)
    emit_cells(((uint32_t)(b_base / 8 + tid) << 12) | ((uint32_t)gcell << 6) | (uint32_t)(r0 >> 2),
               *reinterpret_cast<const uint4*>(S.cells[tid]));
}

}  // namespace b200
