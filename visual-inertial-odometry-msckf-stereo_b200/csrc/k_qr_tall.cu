// Stacked-Jacobian compression of the pruning update (msckf.py:549-553 reached from prune_cam_state_buffer, :730-818).
//
// Shape of that stack: every landmark involves exactly the two camera states being removed, so n_a = 12 active
// columns (+ the residual) and m = 5 rows per landmark, about a thousand rows per filter at 300 landmarks.  A matrix
// that tall and that narrow leaves the blocked tensor-pipe kernel (k_qr_dmma: one panel warp per filter, chunks folded
// one after the other) waiting on a single serial chain; here the row chunks are independent until the end (TSQR):
//   * every warp folds its own 128-row chunks into a private 13 x 13 factor: 4 rows per lane, all 13 columns of a row in
//     registers, 12 Householder reflectors per chunk with one batched warp reduction each (the in-panel scheme of
//     k_qr_dmma without the trailing tiles, there are none);
//   * the warps' factors are stacked and folded once more by warp 0; R and Q^T r go out as H_thin / r_thin.
// Row signs of R may differ from a sequential fold: the update is invariant to them (they flip H_thin and r_thin together).
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"
#include "warp_reduce.cuh"

namespace msckf {

constexpr int TQ_NC = 13;          // 12 active columns + residual
constexpr int TQ_LDR = 14;
constexpr int TQ_WARPS = 4;
constexpr unsigned TQ_FULL = 0xffffffffu;

// Fold the rows held in x (4 per lane) into the upper-triangular factor R (shared memory, row j holds R(j, j..12)).
__device__ __forceinline__ void tq_fold(double (&x)[4][TQ_NC], double* R, int lane) {
#pragma unroll
  for (int jj = 0; jj < TQ_NC - 1; ++jj) {
    // dots of column jj with itself and the columns right of it, compacted to s[0 .. 13 - jj): folded warp reductions
    // for the wide ones (warp_reduce.cuh), a plain butterfly once only a few values are left
    double s[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      double a = 0.0;
      if (c + jj < TQ_NC) {
#pragma unroll
        for (int t = 0; t < 4; ++t) a += x[t][jj] * x[t][c + jj];
      }
      s[c] = a;
    }
    if (TQ_NC - jj > 8) {
      warp_allsum16(s, TQ_NC - jj, lane);
    } else if (TQ_NC - jj > 4) {
      warp_allsum8(s, lane);
    } else {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int c = 0; c < TQ_NC - jj; ++c) s[c] += __shfl_xor_sync(TQ_FULL, s[c], o);
      }
    }
    if (s[0] != 0.0) {                                     // warp-uniform (the reduction leaves identical sums in every lane)
      const double alpha = R[jj * TQ_LDR + jj];
      const double n2 = alpha * alpha + s[0];
      const double rn = rsqrt(n2);                         // 1 / |beta|: one reciprocal square root instead of a square root and a divide
      const double nrm = n2 * rn;
      const double beta = (alpha >= 0) ? -nrm : nrm;
      const double tj = (beta - alpha) * ((alpha >= 0) ? -rn : rn), sc = 1.0 / (alpha - beta);
#pragma unroll
      for (int t = 0; t < 4; ++t) x[t][jj] *= sc;
#pragma unroll
      for (int c = jj + 1; c < TQ_NC; ++c) {
        const double rjc = R[jj * TQ_LDR + c];
        const double w = tj * (rjc + s[c - jj] * sc);
        if (lane == 0) R[jj * TQ_LDR + c] = rjc - w;
#pragma unroll
        for (int t = 0; t < 4; ++t) x[t][c] -= w * x[t][jj];
      }
      if (lane == 0) R[jj * TQ_LDR + jj] = beta;
    }
  }
}

__global__ void __launch_bounds__(32 * TQ_WARPS, 3) k_qr_tall(Dev d) {
  __shared__ double sR[TQ_WARPS][TQ_NC * TQ_LDR];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nlist = (int)d.stage_ctr[SC_NQR_TALL];
  double* R = sR[warp];
  for (int w = blockIdx.x; w < nlist; w += gridDim.x) {
    const int f = d.wl_qrt[w];
    const int m = d.m_rows[f], mp = d.m_pad[f];
    const size_t fb = (size_t)f * d.Fcap;
    const int* rmap = d.rowmap + (size_t)f * d.rowmap_cap;
    for (int e = lane; e < TQ_NC * TQ_LDR; e += 32) R[e] = 0.0;
    __syncwarp();
    const int nchunk = (m + 127) >> 7;
    for (int ch = warp; ch < nchunk; ch += TQ_WARPS) {
      double x[4][TQ_NC];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int i = 128 * ch + 32 * t + lane;
        if (i < m) {
          const int rm = rmap[i];
          const int kk = rm >> 8, li = rm & 255;
          const double* rec = d.pool + d.item_off[fb + kk] + (size_t)li * TQ_NC;     // M = 2: 12 Jacobian entries + residual
#pragma unroll
          for (int c = 0; c < TQ_NC; ++c) x[t][c] = rec[c];
        } else {
#pragma unroll
          for (int c = 0; c < TQ_NC; ++c) x[t][c] = 0.0;
        }
      }
      tq_fold(x, R, lane);
      __syncwarp();
    }
    __syncthreads();
    if (warp == 0) {
      // stack the other warps' factors (3 x 13 rows, upper triangles) and fold them into this warp's
      double x[4][TQ_NC];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int i = 32 * t + lane;
        const bool in = i < (TQ_WARPS - 1) * TQ_NC;
        const int ww = in ? 1 + i / TQ_NC : 1, rj = in ? i % TQ_NC : 0;
#pragma unroll
        for (int c = 0; c < TQ_NC; ++c) x[t][c] = (in && c >= rj) ? sR[ww][rj * TQ_LDR + c] : 0.0;
      }
      tq_fold(x, R, lane);
      __syncwarp();
      // H_thin = R (12 rows), r_thin = first 12 entries of Q^T r
      double* A = d.A + (size_t)f * d.A_elems;
      for (int e = lane; e < (TQ_NC - 1) * TQ_NC; e += 32) {
        const int i = e % (TQ_NC - 1), c = e / (TQ_NC - 1);
        A[(size_t)c * mp + i] = (c >= i) ? R[i * TQ_LDR + c] : 0.0;
      }
    }
    __syncthreads();
  }
}

void launch_qr_tall(const Dev& d, cudaStream_t st) {
  const int sms = sm_count();
  KLAUNCH(d, KID_QR_TALL, st, (k_qr_tall<<<std::min(d.F, sms * 3), 32 * TQ_WARPS, 0, st>>>(d)));
}

}  // namespace msckf
