// Stacked-Jacobian compression of measurement_update (msckf.py:549-553): blocked Householder QR of the m x (n_a + 1)
// matrix [H | r] on the FP64 tensor pipe, one CTA per filter, for compressed updates (m > n_a) with n_a + 1 <= 64.
//
// The stacked matrix is never materialised in HBM.  Rows are expanded from the compact per-landmark records of
// k_jacobian_gate into a shared-memory chunk C (128 rows) and folded into the running factor R (shared memory,
// Q^T r riding along as column n_a) with the triangular-on-top-of-pentagonal scheme of LAPACK's tpqrt, 8 columns at
// a time:
//   1. panel: warp 0 holds the 128 x 8 panel of C in registers (4 rows per lane) and applies up to 8 Householder
//      reflectors v = [e_j ; C(:,j) / (alpha - beta)] with one batched warp reduction per reflector, then forms the
//      8 x 8 triangular factor T of the compact-WY form I - V T V^T from V^T V;
//   2. trailing columns, one 8-column tile per warp:  W = R(panel rows, tile) + V^T C(:, tile)   (DMMA, K = rows)
//                                                   Y = T^T W                                   (DMMA)
//                                                   R(panel rows, tile) -= Y,  C(:, tile) -= V Y (DMMA, K = 8).
// Rows of C and R have a leading dimension of 4 (mod 16) doubles: every fragment load is bank-conflict free.
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int QD_T = 256;          // threads per CTA
constexpr int QD_CR = 128;         // rows per chunk (4 per lane of the panel warp)
constexpr int QD_LDL = 12;         // leading dimension of T and of the per-warp W / Y staging tiles
constexpr unsigned QFULL = 0xffffffffu;

__device__ __forceinline__ void qd_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double qd_allsum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(QFULL, v, o);
  return v;
}

size_t qr_dmma_smem_bytes(int cap) {
  const size_t ldr = cap + 4;
  const size_t dbl = (size_t)(cap + QD_CR) * ldr + 8 * QD_LDL + 8 * 8 * QD_LDL + 8;
  return dbl * sizeof(double) + 64 * sizeof(int) + 16;
}

// Panel factorisation by one warp.  x[t][c] = C(lane + 32 t, j0 + c); reflector jj is applied to the panel columns
// c > jj only (the other tiles get the blocked update).  Returns V in x, writes R(j0 + jj, j0 + jj ..) and T.
__device__ __forceinline__ void qd_panel(double* sR, double* sC, double* sTm, int ldr, int j0, int nrefl, int lane) {
  double x[4][8];
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const double* row = sC + (lane + 32 * t) * ldr + j0;
#pragma unroll
    for (int c = 0; c < 8; c += 2) {
      const double2 v = *(const double2*)(row + c);
      x[t][c] = v.x;
      x[t][c + 1] = v.y;
    }
  }
  double tau[8];
#pragma unroll
  for (int jj = 0; jj < 8; ++jj) {
    tau[jj] = 0.0;
    if (jj < nrefl) {                                     // warp-uniform
      double s[8];
#pragma unroll
      for (int c = jj; c < 8; ++c) {
        double a = 0.0;
#pragma unroll
        for (int t = 0; t < 4; ++t) a += x[t][jj] * x[t][c];
        s[c] = a;
      }
      double rrow[8];
#pragma unroll
      for (int c = jj; c < 8; ++c) rrow[c] = sR[(j0 + jj) * ldr + j0 + c];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int c = jj; c < 8; ++c) s[c] += __shfl_xor_sync(QFULL, s[c], o);
      }
      if (s[jj] != 0.0) {
        const double alpha = rrow[jj];
        const double nrm = sqrt(alpha * alpha + s[jj]);
        const double beta = (alpha >= 0) ? -nrm : nrm;
        const double tj = (beta - alpha) / beta, sc = 1.0 / (alpha - beta);
        tau[jj] = tj;
#pragma unroll
        for (int t = 0; t < 4; ++t) x[t][jj] *= sc;
#pragma unroll
        for (int c = jj + 1; c < 8; ++c) {
          const double w = tj * (rrow[c] + s[c] * sc);
          if (lane == 0) sR[(j0 + jj) * ldr + j0 + c] = rrow[c] - w;
#pragma unroll
          for (int t = 0; t < 4; ++t) x[t][c] -= w * x[t][jj];
        }
        if (lane == 0) sR[(j0 + jj) * ldr + j0 + jj] = beta;
      }
    } else {
      // not a reflector column (residual / padding): it carries data of the last panel, which has no trailing tiles
    }
  }
  // V back to the chunk (only the reflector columns changed meaning; the others hold the updated data)
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    double* row = sC + (lane + 32 * t) * ldr + j0;
#pragma unroll
    for (int c = 0; c < 8; c += 2) *(double2*)(row + c) = make_double2(x[t][c], x[t][c + 1]);
  }
  // T (8 x 8 upper): T(i,i) = tau_i, T(i,j) = -tau_j sum_{l=i}^{j-1} T(i,l) (v_l . v_j); lane i < 8 owns row i
  double trow[8];
  const int ri = lane & 7;
#pragma unroll
  for (int l = 0; l < 8; ++l) trow[l] = 0.0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    double gcol[8];
#pragma unroll
    for (int l = 0; l < j; ++l) {
      double a = 0.0;
#pragma unroll
      for (int t = 0; t < 4; ++t) a += x[t][l] * x[t][j];
      gcol[l] = a;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int l = 0; l < j; ++l) gcol[l] += __shfl_xor_sync(QFULL, gcol[l], o);
    }
    double acc = 0.0;
#pragma unroll
    for (int l = 0; l < j; ++l) acc += trow[l] * gcol[l];
    trow[j] = (ri == j) ? tau[j] : ((ri < j && j < nrefl) ? -tau[j] * acc : 0.0);
  }
  if (lane < 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) sTm[ri * QD_LDL + j] = trow[j];
  }
}

}  // namespace msckf
