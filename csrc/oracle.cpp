// CPU oracles: plain C++ restatements of every lab's arithmetic, in the exact
// operation order of the GPU paths.  Used by the CPU binaries and by the tests
// as the fp reference.  Compile WITHOUT -mfma/-ffast-math: fused operations are
// spelled out with fmaf()/fma() where (and only where) the GPU code fuses.
// Built twice: serial for libb200labs (it is loaded into Python processes that manage
// their own thread pools), and with -DB200_ORACLE_THREADS for the stand-alone CPU
// binaries, where the element-wise loops are split over std::thread workers — every
// output element is computed independently, so results do not depend on the thread
// count.  (std::thread rather than OpenMP: the image has no libgomp development files.)
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#ifdef B200_ORACLE_THREADS
#include <thread>
#endif

#include "b200labs.h"

namespace {

// f(begin, end) over [0, n) in at most hardware_concurrency() contiguous pieces of >= grain
template <typename F>
void parallel_for(int64_t n, int64_t grain, F f) {
#ifdef B200_ORACLE_THREADS
  int64_t pieces = n / std::max<int64_t>(grain, 1);
  if (pieces > 1)  // (hardware_concurrency() reads /sys: tens of microseconds, so only when it can pay)
    pieces = std::min<int64_t>(pieces, std::max<int64_t>(1, (int64_t)std::thread::hardware_concurrency()));
  if (pieces > 1) {
    std::vector<std::thread> workers;
    for (int64_t k = 1; k < pieces; ++k) workers.emplace_back(f, k * n / pieces, (k + 1) * n / pieces);
    f(0, n / pieces);
    for (auto& t : workers) t.join();
    return;
  }
#else
  (void)grain;
#endif
  f(0, n);
}

struct Px {
  uint8_t r, g, b, a;
};

// Luminance with the contraction nvcc emits for
//   0.299f*R + 0.587f*G + 0.114f*B      (ref: lab2/src/main.cu:30-33)
// namely  fma(.114,B, fma(.299,R, rn(.587*G)))  (checked in the sm_100a SASS).
inline float luma(const Px& p) {
  float t = 0.587f * (float)p.g;
  t = fmaf(0.299f, (float)p.r, t);
  return fmaf(0.114f, (float)p.b, t);
}

inline uint8_t roberts_px(float y00, float y10, float y01, float y11) {
  float gx = y11 - y00;
  float gy = y10 - y01;
  // Gx*Gx + Gy*Gy contracts to fma(Gy,Gy, rn(Gx*Gx))  (ref: lab2/src/main.cu:40)
  float s = fmaf(gy, gy, gx * gx);
  float g = sqrtf(s);
  g = fminf(fmaxf(g, 0.0f), 255.0f);
  return (uint8_t)g;
}

// dtype-generic radix-free reference sort
template <typename T>
void sort_t(void* keys, size_t n) {
  T* k = (T*)keys;
  std::sort(k, k + n);
}

inline uint32_t f32_order_bits(uint32_t u) { return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }

}  // namespace

extern "C" {

void b200_cpu_vecsub_f64(const double* a, const double* b, double* c, size_t n) {
  parallel_for((int64_t)n, 1 << 16, [=](int64_t i0, int64_t i1) {
    for (int64_t i = i0; i < i1; ++i) c[i] = a[i] - b[i];
  });
}

void b200_cpu_roberts(const void* in_, void* out_, int w, int h) {
  const Px* in = (const Px*)in_;
  Px* out = (Px*)out_;
  if (w <= 0 || h <= 0) return;
  // rows [y0, y1): luminance rows are computed once and handed down (cur <- nxt)
  auto band = [&](int y0, int y1) {
    std::vector<float> cur(w + 1), nxt(w + 1);
    auto fill = [&](std::vector<float>& row, int y) {
      const Px* src = in + (size_t)std::min(y, h - 1) * w;
      for (int x = 0; x < w; ++x) row[x] = luma(src[x]);
      row[w] = row[w - 1];  // clamp-to-edge on the right
    };
    fill(cur, y0);
    for (int y = y0; y < y1; ++y) {
      fill(nxt, y + 1);  // y+1 == h clamps to the last row
      const Px* src = in + (size_t)y * w;
      Px* dst = out + (size_t)y * w;
      for (int x = 0; x < w; ++x) {
        uint8_t v = roberts_px(cur[x], cur[x + 1], nxt[x], nxt[x + 1]);
        dst[x] = Px{v, v, v, src[x].a};
      }
      cur.swap(nxt);
    }
  };
  parallel_for(h, std::max<int64_t>(1, (1 << 16) / w), [&](int64_t y0, int64_t y1) { band((int)y0, (int)y1); });
}

void b200_cpu_lab3_stats_ref(const void* img_, int w, const int* pts, const int* class_off,
                             int nc, double* params) {
  const Px* img = (const Px*)img_;
  for (int c = 0; c < nc; ++c) {
    const int* p = pts + 2 * (size_t)class_off[c];
    int np = class_off[c + 1] - class_off[c];
    double sum[3] = {0, 0, 0};
    for (int i = 0; i < np; ++i) {
      Px px = img[(size_t)p[2 * i + 1] * w + p[2 * i]];
      sum[0] += px.r;
      sum[1] += px.g;
      sum[2] += px.b;
    }
    double avg[3] = {sum[0] / np, sum[1] / np, sum[2] / np};
    double cov[3][3] = {{0}};
    for (int i = 0; i < np; ++i) {
      Px px = img[(size_t)p[2 * i + 1] * w + p[2 * i]];
      double d[3] = {px.r - avg[0], px.g - avg[1], px.b - avg[2]};
      for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b) cov[a][b] += d[a] * d[b];
    }
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) cov[a][b] /= (np - 1);
    double det = cov[0][0] * (cov[1][1] * cov[2][2] - cov[2][1] * cov[1][2]) -
                 cov[0][1] * (cov[1][0] * cov[2][2] - cov[1][2] * cov[2][0]) +
                 cov[0][2] * (cov[1][0] * cov[2][1] - cov[1][1] * cov[2][0]);
    double* out = params + (size_t)c * B200_LAB3_PARAMS;
    out[0] = avg[0];
    out[1] = avg[1];
    out[2] = avg[2];
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        int b1 = (b + 1) % 3, b2 = (b + 2) % 3, a1 = (a + 1) % 3, a2 = (a + 2) % 3;
        out[3 + a * 3 + b] = (cov[b1][a1] * cov[b2][a2] - cov[b1][a2] * cov[b2][a1]) / det;
      }
  }
}

void b200_cpu_lab3_moments(const void* img_, int w, int64_t y0, int64_t h_local, const int* pts,
                           const int* class_off, int nc, long long* moments) {
  const Px* img = (const Px*)img_;
  for (int c = 0; c < nc; ++c) {
    long long* m = moments + (size_t)c * B200_LAB3_MOMENTS;
    for (int i = class_off[c]; i < class_off[c + 1]; ++i) {
      int64_t x = pts[2 * (size_t)i], y = pts[2 * (size_t)i + 1];
      if (y < y0 || y >= y0 + h_local) continue;
      Px px = img[(size_t)(y - y0) * w + x];
      long long r = px.r, g = px.g, b = px.b;
      m[0] += 1;
      m[1] += r;
      m[2] += g;
      m[3] += b;
      m[4] += r * r;
      m[5] += r * g;
      m[6] += r * b;
      m[7] += g * g;
      m[8] += g * b;
      m[9] += b * b;
    }
  }
}

// Mean = S/n (bit-identical to the reference's sum/np: the sums are exact
// integers).  Covariance from exact integers:  (n*Sab - Sa*Sb) / (n*(n-1)),
// rounded once on conversion and once on division; the reference's two-pass
// sum rounds at every sample, so the two agree to ~1e-15 relative, not bitwise.
void b200_cpu_lab3_finalize(const long long* moments, int nc, double* params) {
  for (int c = 0; c < nc; ++c) {
    const long long* m = moments + (size_t)c * B200_LAB3_MOMENTS;
    double* out = params + (size_t)c * B200_LAB3_PARAMS;
    long long n = m[0];
    double dn = (double)n;
    out[0] = (double)m[1] / dn;
    out[1] = (double)m[2] / dn;
    out[2] = (double)m[3] / dn;
    const long long S[3] = {m[1], m[2], m[3]};
    const int pair[3][3] = {{4, 5, 6}, {5, 7, 8}, {6, 8, 9}};
    double denom = (double)(n * (n - 1));
    double cov[3][3];
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        long long num = n * m[pair[a][b]] - S[a] * S[b];
        cov[a][b] = (double)num / denom;
      }
    double det = cov[0][0] * (cov[1][1] * cov[2][2] - cov[2][1] * cov[1][2]) -
                 cov[0][1] * (cov[1][0] * cov[2][2] - cov[1][2] * cov[2][0]) +
                 cov[0][2] * (cov[1][0] * cov[2][1] - cov[1][1] * cov[2][0]);
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        int b1 = (b + 1) % 3, b2 = (b + 2) % 3, a1 = (a + 1) % 3, a2 = (a + 2) % 3;
        out[3 + a * 3 + b] = (cov[b1][a1] * cov[b2][a2] - cov[b1][a2] * cov[b2][a1]) / det;
      }
  }
}

// Quadratic form with the fused chain the GPU executes
// (ref: lab3/src/main.cu:49-71; DFMA chain checked in SASS).
void b200_cpu_lab3_classify(void* img_, int64_t npix, int nc, const double* params) {
  Px* img = (Px*)img_;
  parallel_for(npix, 1 << 12, [=](int64_t i0, int64_t i1) {
  for (int64_t i = i0; i < i1; ++i) {
    Px px = img[i];
    double best = DBL_MAX;
    int best_c = -1;
    for (int c = 0; c < nc; ++c) {
      const double* p = params + (size_t)c * B200_LAB3_PARAMS;
      double d[3] = {(double)px.r - p[0], (double)px.g - p[1], (double)px.b - p[2]};
      double dist = 0.0;
      for (int k = 0; k < 3; ++k) {
        double t = 0.0;
        for (int j = 0; j < 3; ++j) t = fma(d[j], p[3 + j * 3 + k], t);
        dist = fma(t, d[k], dist);
      }
      if (dist < best) {
        best = dist;
        best_c = c;
      }
    }
    img[i].a = (uint8_t)best_c;
  }
  });
}

void b200_cpu_sort(void* keys, size_t n, int key_type) {
  switch (key_type) {
    case B200_KEY_I32:
      sort_t<int32_t>(keys, n);
      break;
    case B200_KEY_U32:
      sort_t<uint32_t>(keys, n);
      break;
    case B200_KEY_U8:
      sort_t<uint8_t>(keys, n);
      break;
    case B200_KEY_F32: {
      // total order on the bit pattern (the order the GPU radix sort realises):
      // -NaN < -inf < ... < -0 < +0 < ... < +inf < +NaN
      uint32_t* k = (uint32_t*)keys;
      std::sort(k, k + n,
                [](uint32_t a, uint32_t b) { return f32_order_bits(a) < f32_order_bits(b); });
      break;
    }
    default:
      break;
  }
}

}  // extern "C"
