// lab3 device math shared by the single-GPU and the fused multi-GPU kernels.
#pragma once
#include "common.cuh"

namespace b200 {

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


}  // namespace b200
