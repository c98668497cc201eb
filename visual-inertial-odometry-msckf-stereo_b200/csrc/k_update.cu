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
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int UT = 256;

// ------------------------------------------------------------------ QR (jit_utils.py:173-175 / msckf.py:549-553)
// One CTA per filter; LAPACK-style Householder vectors stored below the diagonal, then zeroed
// in the top n_a rows so that H_thin = R is a clean upper triangle.  The residual rides along as
// the last column, which yields r_thin = Q^T r without forming Q.
__global__ void __launch_bounds__(UT) k_qr(Dev d) {
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m = d.m_rows[f], na = d.n_act[f], mp = d.m_pad[f];
  if (m <= na || m == 0) return;
  double* A = d.A + (size_t)f * d.A_elems;
  __shared__ double s_red[UT / 32];
  __shared__ double s_tau, s_v0inv;
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
      // apply H_j = I - tau v v^T to the remaining columns (incl. the residual column): warp per column
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
  // clean lower triangle of the top n_a x n_a block
  for (int e = tid; e < na * na; e += UT) {
    const int c = e / na, i = e - c * na;
    if (i > c) A[(size_t)c * mp + i] = 0.0;
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
__global__ void __launch_bounds__(UT) k_upd_T(Dev d) {
  const int f = blockIdx.z, k = d.k_rows[f];
  const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
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

// S = T[:,act] H_thin^T + sigma^2 I  (k x k).   grid (ceil(Tk/32), ceil(Tk/32), F)
__global__ void __launch_bounds__(UT) k_upd_S(Dev d) {
  const int f = blockIdx.z, k = d.k_rows[f];
  const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
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

// Cholesky of S (lower, in place, row-major), one CTA per filter.  Right-looking, one column at
// a time; rows are distributed over threads.  (np.linalg.solve in the reference is an LU; S is
// SPD so the factorisations agree to rounding.)
__global__ void __launch_bounds__(UT) k_upd_chol(Dev d) {
  const int f = blockIdx.x, tid = threadIdx.x, k = d.k_rows[f];
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
    atomicOr(d.err, ERRBIT_NOT_SPD);
    d.k_rows[f] = 0;       // later stages skip this filter
  }
}

// W = L^-1 T (forward substitution, in place), 32 columns per CTA.   grid (ceil((nmax+1)/32), F)
__global__ void __launch_bounds__(UT) k_upd_trsm(Dev d) {
  const int f = blockIdx.y, k = d.k_rows[f];
  if (k == 0) return;
  const int n = NI + 6 * d.ncam[f];
  const int j0 = blockIdx.x * 32;
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

// dx = W^T y, state injection (msckf.py:565-595).  One CTA per filter.
__global__ void __launch_bounds__(UT) k_upd_inject(Dev d) {
  const int f = blockIdx.x, tid = threadIdx.x, k = d.k_rows[f];
  if (k == 0) return;
  const int N = d.ncam[f], n = NI + 6 * N, tld = Tld(d);
  const double* T = d.T + (size_t)f * d.Tk * tld;
  double* dx = d.dx + (size_t)f * d.nmax;
  for (int j = tid; j < n; j += UT) {
    double s = 0;
    for (int i = 0; i < k; ++i) s += T[(size_t)i * tld + j] * T[(size_t)i * tld + n];
    dx[j] = s;
  }
  __syncthreads();
  if (tid == 0) {
    double* imu = d.imu + (size_t)f * IMU_STRIDE;
    double dq[4], q[4];
    small_angle_quat(dx, dq);
    quat_mul(dq, imu + MSCKF_IMU_Q, q);
    for (int i = 0; i < 4; ++i) imu[MSCKF_IMU_Q + i] = q[i];
    for (int i = 0; i < 3; ++i) {
      imu[MSCKF_IMU_BG + i] += dx[3 + i];
      imu[MSCKF_IMU_V + i] += dx[6 + i];
      imu[MSCKF_IMU_BA + i] += dx[9 + i];
      imu[MSCKF_IMU_P + i] += dx[12 + i];
    }
    if (imu[MSCKF_IMU_NULL_ALIASED] != 0.0) {       // the reference's "+=" moves the aliased null members too
      for (int i = 0; i < 3; ++i) {
        imu[MSCKF_IMU_V_NULL + i] = imu[MSCKF_IMU_V + i];
        imu[MSCKF_IMU_P_NULL + i] = imu[MSCKF_IMU_P + i];
      }
    }
    double dqe[4], Re[9], Rn[9];
    small_angle_quat(dx + 15, dqe);
    to_rotation(dqe, Re);
    mat3_mul(Re, imu + MSCKF_IMU_R_IC, Rn);
    for (int i = 0; i < 9; ++i) imu[MSCKF_IMU_R_IC + i] = Rn[i];
    for (int i = 0; i < 3; ++i) imu[MSCKF_IMU_T_CI + i] += dx[18 + i];
  }
  for (int c = tid; c < N; c += UT) {
    double* cs = d.cam + ((size_t)f * d.Nmax + c) * CAM_STRIDE;
    const double* dc = dx + NI + 6 * c;
    double dq[4], q[4];
    small_angle_quat(dc, dq);
    quat_mul(dq, cs, q);
    for (int i = 0; i < 4; ++i) cs[i] = q[i];
    for (int i = 0; i < 3; ++i) {
      cs[4 + i] += dc[3 + i];
      cs[11 + i] = cs[4 + i];      // position_null aliases _vio_position__ (msckf.py:400, 595)
    }
  }
}

// P <- P - W^T W.   grid (ceil(nmax/32), ceil(nmax/32), F); upper tiles only, mirrored.
__global__ void __launch_bounds__(UT) k_upd_downdate(Dev d) {
  const int f = blockIdx.z, k = d.k_rows[f];
  if (k == 0) return;
  const int n = NI + 6 * d.ncam[f];
  const int i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
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

// measurement_update(H, r) with a caller-supplied dense H: every state column is active.
__global__ void __launch_bounds__(UT) k_load_dense_H(Dev d, int f, const double* __restrict__ H, const double* __restrict__ r, int m,
                                                      int n) {
  const int tid = blockIdx.x * UT + threadIdx.x, stride = gridDim.x * UT;
  const int mp = max(4, ((m + 3) / 4) * 4);
  double* A = d.A + (size_t)f * d.A_elems;
  for (long long e = tid; e < (long long)m * (n + 1); e += stride) {
    const int c = (int)(e / m), i = (int)(e - (long long)c * m);
    A[(size_t)c * mp + i] = (c < n) ? H[(size_t)i * n + c] : r[i];
  }
  if (tid < n) d.colmap[(size_t)f * d.nmax + tid] = tid;
  if (tid == 0) {
    d.m_rows[f] = m; d.m_pad[f] = mp; d.n_act[f] = n; d.k_rows[f] = (m > n) ? n : m;
  }
}

__global__ void k_zero_krows_except(Dev d, int keep) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f < d.F && f != keep) { d.k_rows[f] = 0; d.m_rows[f] = 0; }
}

void launch_qr(const Dev& d, cudaStream_t st) { KLAUNCH(d, KID_QR, st, (k_qr<<<d.F, UT, 0, st>>>(d))); }

void launch_update(const Dev& d, int stage, cudaStream_t st) {
  (void)stage;
  const int tn = (d.nmax + 1 + 31) / 32, tk = (d.Tk + 31) / 32, tp = (d.nmax + 31) / 32;
  KLAUNCH(d, KID_UPD_T, st, (k_upd_T<<<dim3(tn, tk, d.F), UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_S, st, (k_upd_S<<<dim3(tk, tk, d.F), UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_CHOL, st, (k_upd_chol<<<d.F, UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_TRSM, st, (k_upd_trsm<<<dim3(tn, d.F), UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_INJECT, st, (k_upd_inject<<<d.F, UT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_DOWNDATE, st, (k_upd_downdate<<<dim3(tp, tp, d.F), UT, 0, st>>>(d)));
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
