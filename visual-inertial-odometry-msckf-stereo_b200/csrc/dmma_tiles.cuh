// Small building blocks of the tensor-pipe (mma.sync.m8n8k4.f64 -> DMMA.8x8x4) factorisation kernels, shared by the
// one-CTA update (k_update_dmma.cu) and the long-track gate (k_jacobian.cu): the 8 x 8 diagonal-block Cholesky with
// the inverse of its factor, and the walk over the tiles of a lower triangle.
#pragma once
#include <cuda_runtime.h>

namespace msckf {

constexpr unsigned FULL = 0xffffffffu;
constexpr int LDL = 12;            // leading dimension of the inverted diagonal block

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Cholesky of the 8x8 diagonal block at (p0, p0) of X (lower part) and the inverse of its factor, by one warp.
// Lane r (mod 8) owns row r; entries travel by shuffle.  Linv (lower triangular) goes to sLinv [8][LDL].
__device__ __forceinline__ bool chol8_inverse(const double* sX, int LDX, int p0, double* sLinv, int lane) {
  const int r = lane & 7;
  double a[8], il[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) a[c] = sX[(p0 + r) * LDX + p0 + c];
  bool bad = false;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const double djj = __shfl_sync(FULL, a[j], j);
    if (!(djj > 0.0)) bad = true;                      // warp-uniform
    il[j] = rsqrt(djj);                                // one reciprocal square root instead of a square root and a divide
    const double lrj = (r == j) ? djj * il[j] : a[j] * il[j];
    a[j] = lrj;
#pragma unroll
    for (int c = j + 1; c < 8; ++c) {
      const double lcj = __shfl_sync(FULL, lrj, c);
      if (r >= c) a[c] -= lrj * lcj;
    }
  }
  // column c = lane & 7 of L^-1 by forward substitution; L[r2][m] is broadcast from its owner lane
  double x[8];
#pragma unroll
  for (int r2 = 0; r2 < 8; ++r2) {
    double s = (r2 == r) ? 1.0 : 0.0;
#pragma unroll
    for (int m = 0; m < r2; ++m) {
      const double lrm = __shfl_sync(FULL, a[m], r2);
      s -= lrm * x[m];
    }
    x[r2] = s * il[r2];
  }
  if (lane < 8) {
#pragma unroll
    for (int r2 = 0; r2 < 8; ++r2) sLinv[r2 * LDL + r] = x[r2];
  }
  return bad;
}

// triangular index t -> (i, j), j <= i, row by row
__device__ __forceinline__ void tri_decode(int t, int& i, int& j) {
  i = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
  while ((i + 1) * (i + 2) / 2 <= t) ++i;
  while (i * (i + 1) / 2 > t) --i;
  j = t - i * (i + 1) / 2;
}
// the tile `step` places further on in the same row-by-row order
__device__ __forceinline__ void tri_advance(int& i, int& j, int step) {
  j += step;
  while (j > i) { j -= i + 1; i += 1; }
}

}  // namespace msckf
