// Panel factorisation shared by the stacked-Jacobian compression kernels (k_qr_dmma.cu, k_qr_wide.cu): one warp folds
// the 8 panel columns of a chunk of 32 RPL stacked rows into the 8 x 8 diagonal block of the running triangular factor
// (the triangle-on-top-of-pentagon step of LAPACK's tpqrt) and forms the 8 x 8 factor T of the compact-WY form
// I - V T V^T.  Lane l holds rows l, l + 32, ... of the panel in registers; every reflector costs one batched warp
// reduction (its norm and its dots with the columns to the right), T one more (all of V^T V).
#pragma once
#include "warp_reduce.cuh"

namespace msckf {

// Rin(jj, c) = Rin[jj * ldin + c]: the diagonal block of the factor as it is (shared memory); Rout(jj, c) receives the new
// values (may alias Rin).  sC: the chunk, leading dimension ldc, panel columns at sC[row * ldc + 0 .. 7]; V replaces them.
// sTm: [8][ldt].  nrefl < 8 on the last panel: the remaining columns carry data (residual / padding), not reflectors.
template <int RPL, bool FOLD = true>
__device__ __forceinline__ void qr_panel(const double* Rin, int ldin, double* Rout, int ldout, double* sC, int ldc, double* sTm, int ldt,
                                         int nrefl, bool need_T, int lane) {
  double x[RPL][8];
#pragma unroll
  for (int t = 0; t < RPL; ++t) {
    const double* row = sC + (size_t)(lane + 32 * t) * ldc;
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
      for (int c = 0; c < 8; ++c) {
        double a = 0.0;
        if (c >= jj) {
#pragma unroll
          for (int t = 0; t < RPL; ++t) a += x[t][jj] * x[t][c];
        }
        s[c] = a;
      }
      double rrow[8];
#pragma unroll
      for (int c = jj; c < 8; ++c) rrow[c] = Rin[jj * ldin + c];
      if (FOLD && jj < 4) {
        warp_allsum8(s, lane);                            // 5 .. 8 values: folded while they travel
      } else {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int c = jj; c < 8; ++c) s[c] += __shfl_xor_sync(WR_FULL, s[c], o);
        }
      }
      if (s[jj] != 0.0) {
        const double alpha = rrow[jj];
        const double n2 = alpha * alpha + s[jj];
        const double rn = rsqrt(n2);                       // 1 / |beta|: one reciprocal square root instead of a square root and a divide
        const double nrm = n2 * rn;
        const double beta = (alpha >= 0) ? -nrm : nrm;
        const double tj = (beta - alpha) * ((alpha >= 0) ? -rn : rn), sc = 1.0 / (alpha - beta);
        tau[jj] = tj;
#pragma unroll
        for (int t = 0; t < RPL; ++t) x[t][jj] *= sc;
#pragma unroll
        for (int c = jj + 1; c < 8; ++c) {
          const double w = tj * (rrow[c] + s[c] * sc);
          if (lane == 0) Rout[jj * ldout + c] = rrow[c] - w;
#pragma unroll
          for (int t = 0; t < RPL; ++t) x[t][c] -= w * x[t][jj];
        }
        if (lane == 0) Rout[jj * ldout + jj] = beta;
      }
    }
  }
  if (!need_T) return;       // last panel of the chunk: nothing reads V, T or the chunk again
  // V back to the chunk (only the reflector columns changed meaning; the others hold the updated data)
#pragma unroll
  for (int t = 0; t < RPL; ++t) {
    double* row = sC + (size_t)(lane + 32 * t) * ldc;
#pragma unroll
    for (int c = 0; c < 8; c += 2) *(double2*)(row + c) = make_double2(x[t][c], x[t][c + 1]);
  }
  // T (8 x 8 upper): T(i,i) = tau_i, T(i,j) = -tau_j sum_{l=i}^{j-1} T(i,l) (v_l . v_j); lane i < 8 owns row i.
  // All 28 dot products of V^T V go through one folded warp reduction (slot 4 j + ... : pairs (l, j), l < j, packed).
  double gv[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) gv[i] = 0.0;
#pragma unroll
  for (int j = 1; j < 8; ++j)
#pragma unroll
    for (int l = 0; l < j; ++l) {
      double a = 0.0;
#pragma unroll
      for (int t = 0; t < RPL; ++t) a += x[t][l] * x[t][j];
      gv[j * (j - 1) / 2 + l] = a;
    }
  if (FOLD) {
    warp_allsum32(gv, lane);
  } else {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int i = 0; i < 28; ++i) gv[i] += __shfl_xor_sync(WR_FULL, gv[i], o);
    }
  }
  double trow[8];
  const int ri = lane & 7;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    double acc = 0.0;
#pragma unroll
    for (int l = 0; l < j; ++l) acc += trow[l] * gv[j * (j - 1) / 2 + l];
    trow[j] = (ri == j) ? tau[j] : ((ri < j) ? -tau[j] * acc : 0.0);
  }
  if (lane < 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) sTm[ri * ldt + j] = trow[j];
  }
}

}  // namespace msckf
