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

// ------------------------------------------------------------------ stacking + QR (msckf.py:549-553, 672-674, 790-792)
// One CTA per filter.  The stacked Jacobian is never materialised: rows are expanded chunk by chunk from the
// compact per-landmark records of k_jacobian_gate (H' = E - V Z, r' = Q^T r) straight into shared memory and
// folded into the running triangular factor R (packed, with Q^T r riding along as column n_a) by Householder
// reflectors on [R(j,j) ; chunk(:,j)].  Landmarks arrive oldest first, so a chunk of short tracks only touches
// the trailing camera columns: elimination starts at the chunk's first non-zero column.  When m <= n_a the
// reference skips the QR (msckf.py:549): rows are expanded straight into the output.
constexpr int QT = 256;
constexpr int QR_MAX_CR = 256;      // rows per chunk (8 per lane)
constexpr int QR_HDR = 2 * QR_MAX_CR + QR_MAX_CR / 2;   // doubles of per-row metadata in front of the buffers

__host__ __device__ inline int qr_packed_size(int na) { return na * (na + 1) / 2 + na; }   // row j holds columns j..na
__host__ __device__ inline int qr_row_off(int j, int na) { return j * (na + 1) - j * (j - 1) / 2; }

// element (projected row li, active column of camera `cam`, component e6) of a landmark record [4M-3][6M+1]
__device__ __forceinline__ double record_elem(const double* rec, int M, unsigned long long lmask, int li, int cam, int e6, bool is_r) {
  const int ldr = 6 * M + 1;
  if (is_r) return rec[li * ldr + 6 * M];
  if (!((lmask >> cam) & 1ull)) return 0.0;
  const int a = __popcll(lmask & ((1ull << cam) - 1ull));
  return rec[li * ldr + 6 * a + e6];
}

__device__ void qr_build_body(const Dev& d, int f, int smem_doubles, double* sm) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m = d.m_rows[f];
  if (m == 0 || d.qr_cls[f] != QR_DMMA_CLASSES) return;   // compressed updates of up to 64 columns go through k_qr_dmma
  const int na = d.n_act[f], nc = na + 1, k = d.k_rows[f], mp = d.m_pad[f], ni = d.nitem[f];
  const bool do_qr = m > na;
  double* A = d.A + (size_t)f * d.A_elems;
  const int* cm = d.colmap + (size_t)f * d.nmax;
  const size_t fb = (size_t)f * d.Fcap;
  __shared__ int s_c0;
  __shared__ int s_cam[64];          // logical camera of active column block
  for (int b = tid; b < na / 6; b += QT) s_cam[b] = (cm[6 * b] - NI) / 6;

  if (!do_qr) {
    __syncthreads();
    // rows go out as they are: H_thin = H (k = m rows)
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
    return;
  }

  // ---- buffers: packed R (+ rhs), row -> item map, chunk (column-major, ld = CR)
  const int Rsize = qr_packed_size(na);
  double* Rp;
  double* Ch;
  int CR;
  // per-row metadata of the current chunk: record offset, logical observation mask, (local row | M << 16)
  long long* row_rec = (long long*)sm;                  // [QR_MAX_CR]
  unsigned long long* row_lmask = (unsigned long long*)(sm + QR_MAX_CR);
  int* row_meta = (int*)(sm + 2 * QR_MAX_CR);
  double* sm_d = sm + QR_HDR;
  {
    const int avail = smem_doubles - QR_HDR - Rsize;
    CR = avail > 0 ? min(QR_MAX_CR, (avail / nc) & ~31) : 0;
    if (CR >= 32) {
      Rp = sm_d;
      Ch = sm_d + Rsize;
    } else {                                            // does not fit: global scratch (rare: very wide windows)
      Rp = d.qr_scratch + (size_t)f * d.qr_scratch_stride;
      Ch = Rp + Rsize;
      CR = QR_MAX_CR;
    }
  }
  for (int e = tid; e < Rsize; e += QT) Rp[e] = 0.0;
  __syncthreads();

  int it0 = 0;                                          // first item that may still have rows at or after r0
  for (int r0 = 0; r0 < m; r0 += CR) {
    const int cr = min(CR, m - r0);
    if (tid == 0) s_c0 = na;
    __syncthreads();
    // row -> item map of this chunk, first non-zero column
    for (int kk = it0 + tid; kk < ni; kk += QT) {
      const int ro = d.item_rowoff[fb + kk];
      if (ro < 0) continue;
      const int rows = d.item_rows[fb + kk];
      if (ro + rows <= r0 || ro >= r0 + cr) continue;
      const int lo = max(ro, r0), hi = min(ro + rows, r0 + cr);
      const unsigned long long lmask = d.item_lmask[fb + kk];
      const long long off = d.item_off[fb + kk];
      const int M = (rows + 3) / 4;
      for (int g = lo; g < hi; ++g) {
        row_rec[g - r0] = off;
        row_lmask[g - r0] = lmask;
        row_meta[g - r0] = (g - ro) | (M << 16);
      }
      const int first_cam = __ffsll((long long)lmask) - 1;
      int b0 = 0;
      while (b0 < na / 6 && s_cam[b0] < first_cam) ++b0;
      atomicMin(&s_c0, 6 * b0);
    }
    __syncthreads();
    const int c0 = s_c0;
    // expand the rows (columns c0 .. na)
    for (int e = tid; e < cr * (nc - c0); e += QT) {
      const int i = e % cr, c = c0 + e / cr;
      const int meta = row_meta[i];
      Ch[(size_t)c * CR + i] = record_elem(d.pool + row_rec[i], meta >> 16, row_lmask[i], meta & 0xffff, c < na ? s_cam[c / 6] : 0, c % 6,
                                           c == na);
    }
    __syncthreads();
    // ---- fold the chunk into R
    for (int j = c0; j < na; ++j) {
      double v[QR_MAX_CR / 32];
      double s = 0;
      const double* colj = Ch + (size_t)j * CR;
#pragma unroll
      for (int t = 0; t < QR_MAX_CR / 32; ++t) {
        const int i = lane + 32 * t;
        const double x = (i < cr) ? colj[i] : 0.0;
        v[t] = x;
        s += x * x;
      }
      s = warp_sum(s);
      if (s == 0.0) continue;                           // uniform over the CTA: nothing below the diagonal
      const int ro = qr_row_off(j, na);
      const double alpha = Rp[ro];
      const double nrm = sqrt(alpha * alpha + s);
      const double beta = (alpha >= 0) ? -nrm : nrm;
      const double tau = (beta - alpha) / beta, sc = 1.0 / (alpha - beta);
#pragma unroll
      for (int t = 0; t < QR_MAX_CR / 32; ++t) v[t] *= sc;
      for (int c = j + 1 + warp; c <= na; c += QT / 32) {
        double* col = Ch + (size_t)c * CR;
        double dot = 0;
#pragma unroll
        for (int t = 0; t < QR_MAX_CR / 32; ++t) {
          const int i = lane + 32 * t;
          if (i < cr) dot += v[t] * col[i];
        }
        dot = warp_sum(dot);
        const double rjc = Rp[ro + c - j];
        const double w = tau * (rjc + dot);
        if (lane == 0) Rp[ro + c - j] = rjc - w;
#pragma unroll
        for (int t = 0; t < QR_MAX_CR / 32; ++t) {
          const int i = lane + 32 * t;
          if (i < cr) col[i] -= w * v[t];
        }
      }
      __syncthreads();
      if (tid == 0) Rp[ro] = beta;
    }
    __syncthreads();
    // items wholly before the next chunk are done
    while (it0 < ni) {
      const int ro = d.item_rowoff[fb + it0];
      if (ro >= 0 && ro + d.item_rows[fb + it0] > r0 + cr) break;
      ++it0;
    }
  }
  // ---- H_thin = R (k = n_a rows), r_thin = first n_a entries of Q^T r
  for (int e = tid; e < na * nc; e += QT) {
    const int i = e % na, c = e / na;
    A[(size_t)c * mp + i] = (c >= i) ? Rp[qr_row_off(i, na) + c - i] : 0.0;
  }
}
// Persistent grid over the filters k_plan_stack queued in wl_qrb (a few per cent of the updates at the headline workload).
__global__ void __launch_bounds__(QT) k_qr_build(Dev d, int smem_doubles) {
  extern __shared__ double sm[];
  const int nlist = (int)d.stage_ctr[SC_NQR_BUILD];
  for (int w = blockIdx.x; w < nlist; w += gridDim.x) {
    qr_build_body(d, d.wl_qrb[w], smem_doubles, sm);
    __syncthreads();
  }
}

int qr_smem_doubles_for(int Nmax) {
  const int na = 6 * Nmax;
  long long want = (long long)QR_HDR + qr_packed_size(na) + 32LL * (na + 1) + 64;
  const long long lo = 110 * 1024 / 8, hi = 220 * 1024 / 8;
  if (want < lo) want = lo;
  if (want > hi) want = hi;
  return (int)want;
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

// ------------------------------------------------------------------ tiled FP64 GEMM building block
// CTA computes a 32x32 tile; 16x16 threads, 2x2 outputs each; K staged 32 at a time through smem.
template <class LA, class LB>
__device__ __forceinline__ void gemm_tile(int K, LA loadA, LB loadB, double (&acc)[2][2], double (*As)[33], double (*Bs)[33]) {
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  acc[0][0] = acc[0][1] = acc[1][0] = acc[1][1] = 0.0;
  for (int k0 = 0; k0 < K; k0 += 32) {
    for (int e = tid; e < 1024; e += UT) {
      const int a = e >> 5, b = e & 31;
      As[a][b] = loadA(a, k0 + b);     // A(i = a, k = k0 + b)
      Bs[a][b] = loadB(k0 + a, b);     // B(k = k0 + a, j = b)
    }
    __syncthreads();
    const int kk_end = min(32, K - k0);
    for (int kk = 0; kk < kk_end; ++kk) {
      const double a0 = As[ty][kk], a1 = As[ty + 16][kk];
      const double b0 = Bs[kk][tx], b1 = Bs[kk][tx + 16];
      acc[0][0] += a0 * b0; acc[0][1] += a0 * b1;
      acc[1][0] += a1 * b0; acc[1][1] += a1 * b1;
    }
    __syncthreads();
  }
}

// T = H_thin P[act,:]  (k x n), extra column n = r_thin.   grid (ceil((nmax+1)/32), ceil(Tk/32), F)
__device__ void k_upd_T_body(const Dev& d, int f, int bx, int by) {
  const int k = d.k_rows[f];

  const int i0 = by * 32, j0 = bx * 32;
  if (k == 0 || i0 >= k) return;
  const int n = NI + 6 * d.ncam[f];
  if (j0 > n) return;
  __shared__ double As[32][33], Bs[32][33];
  const int na = d.n_act[f], mp = d.m_pad[f], ld = d.ld;
  const double* A = d.A + (size_t)f * d.A_elems;
  const double* P = d.P + (size_t)f * ld * ld;
  const int* cm = d.colmap + (size_t)f * d.nmax;
  double acc[2][2];
  gemm_tile(na,
            [&](int i, int c) { return (i0 + i < k && c < na) ? A[(size_t)c * mp + i0 + i] : 0.0; },
            [&](int c, int j) { return (c < na && j0 + j < n) ? P[(size_t)cm[c] * ld + j0 + j] : 0.0; }, acc, As, Bs);
  double* T = d.T + (size_t)f * d.Tk * Tld(d);
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      const int i = i0 + ty + 16 * a, j = j0 + tx + 16 * b;
      if (i < k) {
        if (j < n) T[(size_t)i * Tld(d) + j] = acc[a][b];
        else if (j == n) T[(size_t)i * Tld(d) + j] = A[(size_t)na * mp + i];
      }
    }
}
__global__ void __launch_bounds__(UT) k_upd_T(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  for (int w = blockIdx.z; w < nbig; w += gridDim.z) {
    k_upd_T_body(d, d.wl_upd_big[w], blockIdx.x, blockIdx.y);
    __syncthreads();
  }
}

// S = T[:,act] H_thin^T + sigma^2 I  (k x k).   grid (ceil(Tk/32), ceil(Tk/32), F)
__device__ void k_upd_S_body(const Dev& d, int f, int bx, int by) {
  const int k = d.k_rows[f];

  const int i0 = by * 32, j0 = bx * 32;
  if (k == 0 || i0 >= k || j0 >= k || j0 > i0) return;     // lower triangle tiles only
  __shared__ double As[32][33], Bs[32][33];
  const int na = d.n_act[f], mp = d.m_pad[f];
  const double* A = d.A + (size_t)f * d.A_elems;
  const double* T = d.T + (size_t)f * d.Tk * Tld(d);
  const int* cm = d.colmap + (size_t)f * d.nmax;
  const int tld = Tld(d);
  double acc[2][2];
  gemm_tile(na,
            [&](int i, int c) { return (i0 + i < k && c < na) ? T[(size_t)(i0 + i) * tld + cm[c]] : 0.0; },
            [&](int c, int j) { return (c < na && j0 + j < k) ? A[(size_t)c * mp + j0 + j] : 0.0; }, acc, As, Bs);
  double* S = d.S + (size_t)f * d.Tk * d.Tk;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      const int i = i0 + ty + 16 * a, j = j0 + tx + 16 * b;
      if (i < k && j < k) S[(size_t)i * d.Tk + j] = acc[a][b] + (i == j ? d.prm->obs_noise : 0.0);
    }
}
__global__ void __launch_bounds__(UT) k_upd_S(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  for (int w = blockIdx.z; w < nbig; w += gridDim.z) {
    k_upd_S_body(d, d.wl_upd_big[w], blockIdx.x, blockIdx.y);
    __syncthreads();
  }
}

// Cholesky of S (lower, in place, row-major), one CTA per filter.  Right-looking, one column at
// a time; rows are distributed over threads.  (np.linalg.solve in the reference is an LU; S is
// SPD so the factorisations agree to rounding.)
__device__ void k_upd_chol_body(const Dev& d, int f, int bx, int by) {
  const int tid = threadIdx.x, k = d.k_rows[f];

  if (k == 0) return;
  double* S = d.S + (size_t)f * d.Tk * d.Tk;
  const int ldS = d.Tk;
  __shared__ int s_bad;
  if (tid == 0) s_bad = 0;
  __syncthreads();
  for (int j = 0; j < k; ++j) {
    const double dj = S[(size_t)j * ldS + j];
    if (!(dj > 0.0)) { if (tid == 0) s_bad = 1; break; }
    const double l = sqrt(dj);
    for (int i = j + 1 + tid; i < k; i += UT) S[(size_t)i * ldS + j] /= l;
    __syncthreads();
    if (tid == 0) S[(size_t)j * ldS + j] = l;
    // trailing update, element-parallel over the lower triangle rows j+1..k-1
    const int rem = k - j - 1;
    const int npairs = rem * (rem + 1) / 2;
    for (int e = tid; e < npairs; e += UT) {
      // row-major enumeration of (i, c), c <= i, both relative to j+1
      int i = (int)((sqrt(8.0 * e + 1.0) - 1.0) * 0.5);
      while ((i + 1) * (i + 2) / 2 <= e) ++i;
      while (i * (i + 1) / 2 > e) --i;
      const int c = e - i * (i + 1) / 2;
      const int gi = j + 1 + i, gc = j + 1 + c;
      S[(size_t)gi * ldS + gc] -= S[(size_t)gi * ldS + j] * S[(size_t)gc * ldS + j];
    }
    __syncthreads();
  }
  __syncthreads();
  if (s_bad && tid == 0) {
    raise_not_spd(d, 3, f, k, d.n_act[f], d.m_rows[f]);
    d.k_rows[f] = 0;       // later stages skip this filter
  }
}
__global__ void __launch_bounds__(UT) k_upd_chol(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
    k_upd_chol_body(d, d.wl_upd_big[w], 0, 0);
    __syncthreads();
  }
}

// W = L^-1 T (forward substitution, in place), 32 columns per CTA.   grid (ceil((nmax+1)/32), F)
__device__ void k_upd_trsm_body(const Dev& d, int f, int bx, int by) {
  const int k = d.k_rows[f];

  if (k == 0) return;
  const int n = NI + 6 * d.ncam[f];
  const int j0 = bx * 32;
  if (j0 > n) return;
  __shared__ double As[32][33], Bs[32][33];
  __shared__ double Lbb[32][33], Xb[32][33];
  const double* S = d.S + (size_t)f * d.Tk * d.Tk;
  const int ldS = d.Tk, tld = Tld(d);
  double* T = d.T + (size_t)f * d.Tk * tld;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  for (int rb = 0; rb < k; rb += 32) {
    double acc[2][2];
    gemm_tile(rb,
              [&](int i, int l) { return (rb + i < k && l < rb) ? S[(size_t)(rb + i) * ldS + l] : 0.0; },
              [&](int l, int j) { return (l < rb && j0 + j <= n) ? T[(size_t)l * tld + j0 + j] : 0.0; }, acc, As, Bs);
    for (int a = 0; a < 2; ++a)
      for (int b = 0; b < 2; ++b) {
        const int i = ty + 16 * a, j = tx + 16 * b;
        Xb[i][j] = (rb + i < k && j0 + j <= n) ? T[(size_t)(rb + i) * tld + j0 + j] - acc[a][b] : 0.0;
      }
    for (int e = tid; e < 1024; e += UT) {
      const int i = e >> 5, l = e & 31;
      Lbb[i][l] = (rb + i < k && l <= i) ? S[(size_t)(rb + i) * ldS + rb + l] : (i == l ? 1.0 : 0.0);
    }
    __syncthreads();
    if (tid < 32) {
      const int j = tid;
      for (int i = 0; i < 32; ++i) {
        double x = Xb[i][j];
        for (int l = 0; l < i; ++l) x -= Lbb[i][l] * Xb[l][j];
        Xb[i][j] = x / Lbb[i][i];
      }
    }
    __syncthreads();
    for (int e = tid; e < 1024; e += UT) {
      const int i = e >> 5, j = e & 31;
      if (rb + i < k && j0 + j <= n) T[(size_t)(rb + i) * tld + j0 + j] = Xb[i][j];
    }
    __syncthreads();
  }
}
__global__ void __launch_bounds__(UT) k_upd_trsm(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  for (int w = blockIdx.y; w < nbig; w += gridDim.y) {
    k_upd_trsm_body(d, d.wl_upd_big[w], blockIdx.x, 0);
    __syncthreads();
  }
}

// dx = W^T y, state injection (msckf.py:565-595): inject_state lives in msckf_dev.cuh.  One CTA per filter.
__global__ void __launch_bounds__(UT) k_upd_inject(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int tid = threadIdx.x;
  for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
    const int f = d.wl_upd_big[w], k = d.k_rows[f];
    if (k == 0) continue;
    const int n = NI + 6 * d.ncam[f], tld = Tld(d);
    const double* T = d.T + (size_t)f * d.Tk * tld;
    double* dx = d.dx + (size_t)f * d.nmax;
    for (int j = tid; j < n; j += UT) {
      double s = 0;
      for (int i = 0; i < k; ++i) s += T[(size_t)i * tld + j] * T[(size_t)i * tld + n];
      dx[j] = s;
    }
    __syncthreads();
    inject_state(d, f, dx, tid, UT);
    __syncthreads();
  }
}

// P <- P - W^T W.   grid (ceil(nmax/32), ceil(nmax/32), F); upper tiles only, mirrored.
__device__ void k_upd_downdate_body(const Dev& d, int f, int bx, int by) {
  const int k = d.k_rows[f];

  if (k == 0) return;
  const int n = NI + 6 * d.ncam[f];
  const int i0 = by * 32, j0 = bx * 32;
  if (i0 >= n || j0 >= n || j0 < i0) return;
  __shared__ double As[32][33], Bs[32][33];
  const int tld = Tld(d), ld = d.ld;
  const double* T = d.T + (size_t)f * d.Tk * tld;
  double acc[2][2];
  gemm_tile(k,
            [&](int i, int r) { return (i0 + i < n && r < k) ? T[(size_t)r * tld + i0 + i] : 0.0; },
            [&](int r, int j) { return (r < k && j0 + j < n) ? T[(size_t)r * tld + j0 + j] : 0.0; }, acc, As, Bs);
  double* P = d.P + (size_t)f * ld * ld;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      const int i = i0 + ty + 16 * a, j = j0 + tx + 16 * b;
      if (i < n && j < n && j >= i) {
        const double v = P[(size_t)i * ld + j] - acc[a][b];
        P[(size_t)i * ld + j] = v;
        P[(size_t)j * ld + i] = v;
      }
    }
}
__global__ void __launch_bounds__(UT) k_upd_downdate(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  for (int w = blockIdx.z; w < nbig; w += gridDim.z) {
    k_upd_downdate_body(d, d.wl_upd_big[w], blockIdx.x, blockIdx.y);
    __syncthreads();
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

void launch_qr_build(const Dev& d, cudaStream_t st) {
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_qr_build, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
  });
  const int sms = sm_count();
  KLAUNCH(d, KID_QR, st, (k_qr_build<<<std::min(d.F, 2 * sms), QT, (size_t)d.qr_smem_doubles * sizeof(double), st>>>(d, d.qr_smem_doubles)));
}
void launch_qr_all(const Dev& d, cudaStream_t st) {
  Fork fk(d, st);
  launch_qr_dmma(d, fk);
  launch_qr_build(d, fk.lane(4));
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
    launch_qr_build(d, sb);
    if (fk.on) cudaEventRecord(fk.L->built, sb);
    for (int c = 0; c < ncls; ++c) {              // class c: compression, then its update, back to back on one stream
      cudaStream_t s = fk.lane(c);
      if (c == 0 && stage == 1 && d.qr_tall_ok) launch_qr_tall(d, s);
      launch_qr_dmma_class(d, c, s);
      if (fk.on && s != sb) cudaStreamWaitEvent(s, fk.L->built, 0);
      launch_upd_dmma_class(d, c, s);
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
void launch_update_tiled(const Dev& d, cudaStream_t st) {
  // tiled path for the filters k_plan_stack queued in wl_upd_big (none in the common case: the grids exit at once)
  const int tn = (d.nmax + 1 + 31) / 32, tk = (d.Tk + 31) / 32, tp = (d.nmax + 31) / 32;
  const int fz = d.F < 32 ? d.F : 32;
  KLAUNCH(d, KID_UPD_T, st, (k_upd_T<<<dim3(tn, tk, fz), UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_S, st, (k_upd_S<<<dim3(tk, tk, fz), UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_CHOL, st, (k_upd_chol<<<fz, UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_TRSM, st, (k_upd_trsm<<<dim3(tn, fz), UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_INJECT, st, (k_upd_inject<<<fz, UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_DOWNDATE, st, (k_upd_downdate<<<dim3(tp, tp, fz), UT, 0, st>>>(d)));
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
