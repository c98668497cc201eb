// measurement_update (msckf.py:543-604): Householder QR compression of the stacked Jacobian,
// Kalman gain and state / covariance correction.
//
// Algebra (all FP64).  With H_thin (k x n_a on the active columns) and r_thin:
//     T = H_thin P[act,:]            (k x n)
//     S = T[:,act] H_thin^T + sigma^2 I,   S = L L^T
//     W = L^-1 T,  y = L^-1 r_thin
//     dx = W^T y            ( = K r_thin )
//     P <- P - W^T W        ( = sym((I - K H_thin) P), exactly symmetric by construction )
// which is the reference's update up to rounding (SURVEY.md section 7 hard-part 7: <= 4e-14 per update;
// the reference's own form is (I-KH)P followed by an explicit symmetrisation).  Only the active
// columns of H are factorised: the 21 IMU columns of every Jacobian row are structurally zero
// (msckf.py:517-528) and in the pruning update only the two removed camera states appear.
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int UT = 256;

// ------------------------------------------------------------------ uncompressed stacks (msckf.py:549: m <= n_a skips the QR)
// The stacked Jacobian of a stage never exists as a matrix: the compression kernels (k_qr_dmma, k_qr_tall, k_qr_wide)
// expand the compact per-landmark records of the Jacobian kernels (H' = E - V Z, r' = Q^T r) chunk by chunk in shared
// memory.  Only when the reference would skip the QR are the rows written out as they are: H_thin = H (k = m rows).
constexpr int QT = 256;

// element (projected row li, active column of camera `cam`, component e6) of a landmark record [4M-3][6M+1]
__device__ __forceinline__ double record_elem(const double* rec, int M, unsigned long long lmask, int li, int cam, int e6, bool is_r) {
  const int ldr = 6 * M + 1;
  if (is_r) return rec[li * ldr + 6 * M];
  if (!((lmask >> cam) & 1ull)) return 0.0;
  const int a = __popcll(lmask & ((1ull << cam) - 1ull));
  return rec[li * ldr + 6 * a + e6];
}

// Persistent grid over the filters k_plan_stack queued in wl_qrb (wide or uncompressed stacks); the compressed ones are
// left to k_qr_wide.
__global__ void __launch_bounds__(QT) k_qr_build(Dev d) {
  __shared__ int s_cam[64];          // logical camera of active column block
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nlist = (int)d.stage_ctr[SC_NQR_BUILD];
  for (int w = blockIdx.x; w < nlist; w += gridDim.x) {
    const int f = d.wl_qrb[w];
    const int m = d.m_rows[f], na = d.n_act[f], nc = na + 1, mp = d.m_pad[f], ni = d.nitem[f];
    if (m == 0 || m > na) continue;
    double* A = d.A + (size_t)f * d.A_elems;
    const int* cm = d.colmap + (size_t)f * d.nmax;
    const size_t fb = (size_t)f * d.Fcap;
    __syncthreads();
    for (int b = tid; b < na / 6; b += QT) s_cam[b] = (cm[6 * b] - NI) / 6;
    __syncthreads();
    for (int kk = warp; kk < ni; kk += QT / 32) {
      const int ro = d.item_rowoff[fb + kk];
      if (ro < 0) continue;
      const int rows = d.item_rows[fb + kk], M = (rows + 3) / 4;
      const unsigned long long lmask = d.item_lmask[fb + kk];
      const double* rec = d.pool + d.item_off[fb + kk];
      for (int e = lane; e < rows * nc; e += 32) {
        const int li = e % rows, c = e / rows;
        A[(size_t)c * mp + ro + li] = record_elem(rec, M, lmask, li, c < na ? s_cam[c / 6] : 0, c % 6, c == na);
      }
    }
  }
}

// Dense caller-supplied H (measurement_update API): one CTA factorises Adense (column-major, ld = mp) in place
// with LAPACK-style Householder vectors, then copies H_thin | r_thin into the filter's compressed slot.
__global__ void __launch_bounds__(UT) k_qr_dense(Dev d, int f, int mp) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m = d.m_rows[f], na = d.n_act[f], k = d.k_rows[f], mpo = d.m_pad[f];
  if (m == 0) return;
  double* A = d.Adense;
  __shared__ double s_red[UT / 32];
  __shared__ double s_tau, s_v0inv;
  if (m > na) {
    for (int j = 0; j < na; ++j) {
      double* colj = A + (size_t)j * mp;
      double s = 0;
      for (int i = j + 1 + tid; i < m; i += UT) { const double x = colj[i]; s += x * x; }
      s = warp_sum(s);
      if (lane == 0) s_red[warp] = s;
      __syncthreads();
      if (tid == 0) {
        double xn2 = 0;
        for (int w = 0; w < UT / 32; ++w) xn2 += s_red[w];
        const double alpha = colj[j];
        if (xn2 == 0.0) {
          s_tau = 0.0; s_v0inv = 0.0;
        } else {
          const double nrm = sqrt(alpha * alpha + xn2);
          const double beta = (alpha >= 0) ? -nrm : nrm;
          s_tau = (beta - alpha) / beta;
          s_v0inv = 1.0 / (alpha - beta);
          colj[j] = beta;
        }
      }
      __syncthreads();
      const double tau = s_tau, sc = s_v0inv;
      if (tau != 0.0) {
        for (int i = j + 1 + tid; i < m; i += UT) colj[i] *= sc;
        __syncthreads();
        for (int c = j + 1 + warp; c <= na; c += UT / 32) {
          double* col = A + (size_t)c * mp;
          double w = (lane == 0) ? col[j] : 0.0;
          for (int i = j + 1 + lane; i < m; i += 32) w += colj[i] * col[i];
          w = warp_sum(w) * tau;
          if (lane == 0) col[j] -= w;
          for (int i = j + 1 + lane; i < m; i += 32) col[i] -= w * colj[i];
        }
      }
      __syncthreads();
    }
  }
  double* out = d.A + (size_t)f * d.A_elems;
  for (int e = tid; e < k * (na + 1); e += UT) {
    const int i = e % k, c = e / k;
    out[(size_t)c * mpo + i] = (m > na && i > c) ? 0.0 : A[(size_t)c * mp + i];
  }
}

// measurement_update(H, r) with a caller-supplied dense H: every state column is active.
__global__ void __launch_bounds__(UT) k_load_dense_H(Dev d, int f, const double* __restrict__ H, const double* __restrict__ r, int m,
                                                      int n) {
  const int tid = blockIdx.x * UT + threadIdx.x, stride = gridDim.x * UT;
  const int mp = max(4, ((m + 3) / 4) * 4);
  double* A = d.Adense;
  for (long long e = tid; e < (long long)m * (n + 1); e += stride) {
    const int c = (int)(e / m), i = (int)(e - (long long)c * m);
    A[(size_t)c * mp + i] = (c < n) ? H[(size_t)i * n + c] : r[i];
  }
  if (tid < n) d.colmap[(size_t)f * d.nmax + tid] = tid;
  if (tid == 0) {
    const int kr = (m > n) ? n : m;
    d.m_rows[f] = m; d.m_pad[f] = max(4, ((kr + 3) / 4) * 4); d.n_act[f] = n; d.k_rows[f] = kr;
  }
}

__global__ void k_zero_krows_except(Dev d, int keep) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f < d.F && f != keep) { d.k_rows[f] = 0; d.m_rows[f] = 0; d.upd_path[f] = 0; }
  if (f == 0) {                       // the caller-supplied H goes through the tiled path
    d.stage_ctr[SC_NUPD_BIG] = 1ull;
    d.wl_upd_big[0] = keep;
    d.upd_path[keep] = 2;
  }
}

// wide / tall / uncompressed stacks of a stage (work list wl_qrb).  The pruning stage only ever has n_a = 12: nothing wide,
// but when the batch leaves SMs idle (qr_leaf_cap > 1) k_plan_stack sends its tall stacks (m >= 2048: thousands of live
// landmarks per filter) to the grid-wide TSQR as well.
void launch_qr_build(const Dev& d, int stage, cudaStream_t st) {
  const int sms = sm_count();
  KLAUNCH(d, KID_QR, st, (k_qr_build<<<std::min(d.F, 2 * sms), QT, 0, st>>>(d)));
  if (stage != 1 || d.qr_leaf_cap > 1) launch_qr_wide(d, st);
}
void launch_compress_update(const Dev& d, int stage, cudaStream_t st) {
  bool all_ok = true;
  for (int c = 0; c < UPD_DMMA_CLASSES; ++c) all_ok = all_ok && d.upd_dmma_ok[c];
  // the pruning stage only ever stacks landmarks on the two cameras being removed: n_a = 12, class 0 (or uncompressed)
  const int ncls = (stage == 1) ? 1 : QR_DMMA_CLASSES;
  {
    Fork fk(d, st);
    // wide / uncompressed stacks on their own lane, next to the tensor-pipe compression classes; their updates may sit in
    // any class, so every class's update waits for this lane (event `built`)
    cudaStream_t sb = fk.lane(4);
    launch_qr_build(d, stage, sb);
    if (fk.on) cudaEventRecord(fk.L->built, sb);
    for (int c = 0; c < ncls; ++c) {              // class c: compression, then its update, back to back on one stream
      cudaStream_t s = fk.lane(c);
      if (c == 0 && stage == 1 && d.qr_tall_ok) launch_qr_tall(d, s);
      launch_qr_dmma_class(d, c, s);
      if (fk.on && s != sb) cudaStreamWaitEvent(s, fk.L->built, 0);
      launch_upd_dmma_class(d, c, s, stage);
    }
    // the tiled update only takes stacks k_qr_build compressed when every one-CTA class is available ...
    if (all_ok && stage != 1) launch_update_tiled(d, sb);
  }
  // ... otherwise it also takes filters of a k_qr_dmma class (window too wide for the class's shared memory): after the join
  if (!all_ok) launch_update_tiled(d, st);
}
void launch_qr_dense(const Dev& d, int filter, int mp_dense, cudaStream_t st) {
  KLAUNCH(d, KID_QR, st, (k_qr_dense<<<1, UT, 0, st>>>(d, filter, mp_dense)));
}

void launch_update(const Dev& d, int stage, cudaStream_t st) {
  (void)stage;
  Fork fk(d, st);
  launch_update_dmma(d, fk);
  launch_update_tiled(d, fk.lane(4));
}
void launch_load_dense_H(const Dev& d, int filter, const double* H, const double* r, int m, int n_state, cudaStream_t st) {
  KLAUNCH(d, KID_MISC, st, (k_zero_krows_except<<<(d.F + 255) / 256, 256, 0, st>>>(d, filter)));
  const long long tot = (long long)m * (n_state + 1);
  int blocks = (int)((tot + UT - 1) / UT);
  blocks = blocks < 1 ? 1 : (blocks > 1024 ? 1024 : blocks);
  if (blocks * UT < n_state) blocks = (n_state + UT - 1) / UT;
  KLAUNCH(d, KID_MISC, st, (k_load_dense_H<<<blocks, UT, 0, st>>>(d, filter, H, r, m, n_state)));
}

}  // namespace msckf
