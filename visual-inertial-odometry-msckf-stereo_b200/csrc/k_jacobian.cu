// Per-landmark measurement Jacobian, left-null-space projection and chi-square gate:
// measurement_jacobian (msckf.py:429-498), feature_jacobian (:500-541), gating_test (:606-614).
//
// One warp per landmark.  The work matrix W = [H_f | H_x(compact, 6 columns per observing
// camera) | r] (4M rows, column-major) is built lane-per-observation, H_f is annihilated with
// three Householder reflectors applied to the whole of W (instead of the reference's SVD: any
// orthonormal basis of the left null space gives the same H^T H, H^T r and gamma), and the gate
// gamma = r^T (H P H^T + sigma^2 I)^-1 r is evaluated on the compact columns with a Cholesky
// factorisation (S is SPD by construction; the reference's LU gives the same gamma).
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int JAC_WPB = 4;

// msckf.py:429-498 for one (camera state, landmark).  cs = CAM_STRIDE block of the camera state.
__device__ void measurement_jacobian(const double* cs, const double* p_w, const double* z, const double* g,
                                     const double* R01, const double* t01, double* Hx /*[4][6]*/, double* Hf /*[4][3]*/,
                                     double* r /*[4]*/) {
  double R0[9], R1[9];
  to_rotation(cs, R0);
  const double* t0 = cs + 4;
  mat3_mul(R01, R0, R1);
  double tmp[3], t1[3];
  mat3t_vec(R1, t01, tmp);
  for (int i = 0; i < 3; ++i) t1[i] = t0[i] - tmp[i];
  const double d0[3] = {p_w[0] - t0[0], p_w[1] - t0[1], p_w[2] - t0[2]};
  const double d1[3] = {p_w[0] - t1[0], p_w[1] - t1[1], p_w[2] - t1[2]};
  double pc0[3], pc1[3];
  mat3_vec(R0, d0, pc0);
  mat3_vec(R1, d1, pc1);
  // dz/dp rows (only the non-zero 2x3 halves)
  const double a0[3] = {1 / pc0[2], 0.0, -pc0[0] / (pc0[2] * pc0[2])};
  const double a1[3] = {0.0, 1 / pc0[2], -pc0[1] / (pc0[2] * pc0[2])};
  const double b0[3] = {1 / pc1[2], 0.0, -pc1[0] / (pc1[2] * pc1[2])};
  const double b1[3] = {0.0, 1 / pc1[2], -pc1[1] / (pc1[2] * pc1[2])};
  double S0[9], D1[9];
  skew3(pc0, S0);
  mat3_mul(R01, S0, D1);
  double A[24];
  for (int j = 0; j < 3; ++j) {
    A[0 * 6 + j] = a0[0] * S0[j] + a0[1] * S0[3 + j] + a0[2] * S0[6 + j];
    A[1 * 6 + j] = a1[0] * S0[j] + a1[1] * S0[3 + j] + a1[2] * S0[6 + j];
    A[2 * 6 + j] = b0[0] * D1[j] + b0[1] * D1[3 + j] + b0[2] * D1[6 + j];
    A[3 * 6 + j] = b1[0] * D1[j] + b1[1] * D1[3 + j] + b1[2] * D1[6 + j];
    A[0 * 6 + 3 + j] = a0[0] * -R0[j] + a0[1] * -R0[3 + j] + a0[2] * -R0[6 + j];
    A[1 * 6 + 3 + j] = a1[0] * -R0[j] + a1[1] * -R0[3 + j] + a1[2] * -R0[6 + j];
    A[2 * 6 + 3 + j] = b0[0] * -R1[j] + b0[1] * -R1[3 + j] + b0[2] * -R1[6 + j];
    A[3 * 6 + 3 + j] = b1[0] * -R1[j] + b1[1] * -R1[3 + j] + b1[2] * -R1[6 + j];
  }
  // observability constraint (msckf.py:483-490)
  double u[6], Rn[9];
  to_rotation(cs + 7, Rn);
  mat3_vec(Rn, g, u);
  const double dn[3] = {p_w[0] - cs[11], p_w[1] - cs[12], p_w[2] - cs[13]};
  cross3(dn, g, u + 3);
  double uu = 0;
  for (int i = 0; i < 6; ++i) uu += u[i] * u[i];
  for (int i = 0; i < 4; ++i) {
    double au = 0;
    for (int j = 0; j < 6; ++j) au += A[6 * i + j] * u[j];
    for (int j = 0; j < 6; ++j) Hx[6 * i + j] = A[6 * i + j] - au * u[j] / uu;
    for (int j = 0; j < 3; ++j) Hf[3 * i + j] = -Hx[6 * i + 3 + j];
  }
  r[0] = z[0] - pc0[0] / pc0[2];
  r[1] = z[1] - pc0[1] / pc0[2];
  r[2] = z[2] - pc1[0] / pc1[2];
  r[3] = z[3] - pc1[1] / pc1[2];
}

// grid (ceil(Fcap/JAC_WPB), F); warp w of filter f takes item w.
__global__ void __launch_bounds__(JAC_WPB * 32) k_jacobian_gate(Dev d) {
  __shared__ int s_cl[JAC_WPB][64];     // logical camera index of observation a
  const int f = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k = blockIdx.x * JAC_WPB + warp;
  if (k >= d.nitem[f]) return;
  const size_t ik = (size_t)f * d.Fcap + k;
  const int slot = d.cand[(size_t)f * d.Fcap + d.item_cand[ik]];
  const size_t fs = (size_t)f * d.Fcap + slot;
  const unsigned long long mask = d.item_mask[ik];
  const int N = d.ncam[f];
  const Params& prm = *d.prm;
  const double* g = d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY;
  unsigned long long lmask = 0;
  for (int c0 = 0; c0 < N; c0 += 32) {
    const int c = c0 + lane;
    const int bit = (c < N) ? (int)((mask >> d.cam_phys[(size_t)f * d.Nmax + c]) & 1ull) : 0;
    lmask |= (unsigned long long)__ballot_sync(0xffffffffu, bit) << c0;
  }
  const int M = __popcll(lmask);
  const int R4 = 4 * M, NC = 6 * M + 4, rows = R4 - 3, nc6 = 6 * M;
  double* W = d.pool + d.item_off[ik];
  double* T1 = W + (size_t)R4 * NC;          // [rows][6M] column-major
  double* S = T1 + (size_t)rows * nc6;       // [rows][rows] column-major, lower triangle
  for (int e = lane; e < R4 * NC; e += 32) W[e] = 0.0;
  __syncwarp();
  const double p_w[3] = {d.f_pos[fs * 3], d.f_pos[fs * 3 + 1], d.f_pos[fs * 3 + 2]};
  for (int c = lane; c < N; c += 32) {
    if ((lmask >> c) & 1ull) {
      const int a = __popcll(lmask & ((1ull << c) - 1ull));
      s_cl[warp][a] = c;
      const double* cs = d.cam + ((size_t)f * d.Nmax + c) * CAM_STRIDE;
      const double* z = d.f_obs + (fs * d.Nmax + d.cam_phys[(size_t)f * d.Nmax + c]) * 4;
      double Hx[24], Hf[12], r[4];
      measurement_jacobian(cs, p_w, z, g, prm.R01, prm.t01, Hx, Hf, r);
      for (int i = 0; i < 4; ++i) {
        const int row = 4 * a + i;
        for (int j = 0; j < 3; ++j) W[row + (size_t)j * R4] = Hf[3 * i + j];
        for (int j = 0; j < 6; ++j) W[row + (size_t)(3 + 6 * a + j) * R4] = Hx[6 * i + j];
        W[row + (size_t)(3 + nc6) * R4] = r[i];
      }
    }
  }
  __syncwarp();
  // ---- three Householder reflectors on the H_f columns, applied to every other column
  for (int kk = 0; kk < 3; ++kk) {
    double s = 0;
    for (int i = kk + lane; i < R4; i += 32) { const double x = W[i + (size_t)kk * R4]; s += x * x; }
    s = warp_sum(s);
    if (s == 0.0) continue;
    const double alpha = W[kk + (size_t)kk * R4];
    const double beta = (alpha >= 0) ? -sqrt(s) : sqrt(s);
    const double v0 = alpha - beta;
    const double vtv = s - alpha * alpha + v0 * v0;
    const double scale = 2.0 / vtv;
    const double* vcol = W + (size_t)kk * R4;
    for (int c = kk + 1 + lane; c < NC; c += 32) {
      double* col = W + (size_t)c * R4;
      double dot = v0 * col[kk];
      for (int i = kk + 1; i < R4; ++i) dot += vcol[i] * col[i];
      dot *= scale;
      col[kk] -= dot * v0;
      for (int i = kk + 1; i < R4; ++i) col[i] -= dot * vcol[i];
    }
    __syncwarp();
  }
  // ---- gate: T1 = Hc Pcc ; S = T1 Hc^T + sigma^2 I ; gamma = |L^-1 r|^2
  const double* P = d.P + (size_t)f * d.ld * d.ld;
  const int ld = d.ld;
  const double* Hc = W + 3 + (size_t)3 * R4;         // Hc(i,c) = Hc[i + c*R4]
  for (int e = lane; e < rows * nc6; e += 32) {
    const int i = e % rows, c = e / rows;
    const int pc = NI + 6 * s_cl[warp][c / 6] + (c % 6);
    double acc = 0;
    for (int a = 0; a < M; ++a) {
      const double* Prow = P + (size_t)(NI + 6 * s_cl[warp][a]) * ld + pc;
#pragma unroll
      for (int e2 = 0; e2 < 6; ++e2) acc += Hc[i + (size_t)(6 * a + e2) * R4] * Prow[(size_t)e2 * ld];
    }
    T1[e] = acc;
  }
  __syncwarp();
  const int npairs = rows * (rows + 1) / 2;
  for (int e = lane; e < npairs; e += 32) {
    // (i, j) with i >= j, enumerated column by column
    int j = 0, rem = e;
    while (rem >= rows - j) { rem -= rows - j; ++j; }
    const int i = j + rem;
    double acc = 0;
    for (int c = 0; c < nc6; ++c) acc += T1[i + (size_t)c * rows] * Hc[j + (size_t)c * R4];
    if (i == j) acc += prm.obs_noise;
    S[i + (size_t)j * rows] = acc;
  }
  __syncwarp();
  double* rr = W + 3 + (size_t)(3 + nc6) * R4;       // projected residual (also the stacked r)
  // y overwrites a scratch copy in T1's first column (T1 is no longer needed)
  double* y = T1;
  for (int i = lane; i < rows; i += 32) y[i] = rr[i];
  __syncwarp();
  bool spd = true;
  for (int j = 0; j < rows; ++j) {
    const double dj = S[j + (size_t)j * rows];
    if (!(dj > 0.0)) { spd = false; break; }
    const double l = sqrt(dj);
    for (int i = j + 1 + lane; i < rows; i += 32) S[i + (size_t)j * rows] /= l;
    __syncwarp();
    if (lane == 0) { S[j + (size_t)j * rows] = l; y[j] = y[j] / l; }
    __syncwarp();
    const double yj = y[j];
    for (int i = j + 1 + lane; i < rows; i += 32) {
      const double lij = S[i + (size_t)j * rows];
      y[i] -= lij * yj;
      for (int c = j + 1; c <= i; ++c) S[i + (size_t)c * rows] -= lij * S[c + (size_t)j * rows];
    }
    __syncwarp();
  }
  double gamma = 0;
  for (int i = lane; i < rows; i += 32) gamma += y[i] * y[i];
  gamma = warp_sum(gamma);
  if (!spd) { gamma = INFINITY; if (lane == 0) atomicOr(d.err, ERRBIT_NOT_SPD); }
  if (lane == 0) {
    d.item_gamma[ik] = gamma;
    d.item_pass[ik] = (gamma < prm.chi2[d.item_dof[ik]]) ? 1 : 0;
  }
}

// Copies the projected rows of every stacked landmark into the filter's stacked matrix
// A[m_pad][n_act+1] (column-major): msckf.py:672-674 / 790-792, on the active columns only.
__global__ void __launch_bounds__(JAC_WPB * 32) k_stack(Dev d) {
  const int f = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int k = blockIdx.x * JAC_WPB + warp;
  if (k >= d.nitem[f]) return;
  const size_t ik = (size_t)f * d.Fcap + k;
  const int ro = d.item_rowoff[ik];
  if (ro < 0 || d.m_rows[f] == 0) return;
  const unsigned long long mask = d.item_mask[ik];
  const int rows = d.item_rows[ik], M = (rows + 3) / 4, R4 = 4 * M, nc6 = 6 * M;
  const int na = d.n_act[f], mp = d.m_pad[f];
  const double* W = d.pool + d.item_off[ik];
  double* A = d.A + (size_t)f * d.A_elems;
  const int* cm = d.colmap + (size_t)f * d.nmax;
  const int* phys = d.cam_phys + (size_t)f * d.Nmax;
  // logical mask of this landmark's observations
  const int N = d.ncam[f];
  unsigned long long lmask = 0;
  for (int c0 = 0; c0 < N; c0 += 32) {
    const int c = c0 + lane;
    const int bit = (c < N) ? (int)((mask >> phys[c]) & 1ull) : 0;
    lmask |= (unsigned long long)__ballot_sync(0xffffffffu, bit) << c0;
  }
  for (int e = lane; e < rows * (na + 1); e += 32) {
    const int i = e % rows, ac = e / rows;
    double v;
    if (ac == na) {
      v = W[3 + i + (size_t)(3 + nc6) * R4];
    } else {
      const int sidx = cm[ac] - NI, c = sidx / 6, e6 = sidx - 6 * c;
      if ((lmask >> c) & 1ull) {
        const int a = __popcll(lmask & ((1ull << c) - 1ull));
        v = W[3 + i + (size_t)(3 + 6 * a + e6) * R4];
      } else {
        v = 0.0;
      }
    }
    A[(size_t)ro + i + (size_t)ac * mp] = v;
  }
}

// ------------------------------------------------------------------ API helpers (single filter)
__global__ void k_measurement_jacobian(Dev d, int f, int c, int slot, double* out) {
  if (threadIdx.x != 0) return;
  const size_t fs = (size_t)f * d.Fcap + slot;
  const double* cs = d.cam + ((size_t)f * d.Nmax + c) * CAM_STRIDE;
  const double* z = d.f_obs + (fs * d.Nmax + d.cam_phys[(size_t)f * d.Nmax + c]) * 4;
  measurement_jacobian(cs, d.f_pos + fs * 3, z, d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY, d.prm->R01, d.prm->t01,
                       out, out + 24, out + 36);
}

// dense (rows x n_state) view of item 0 of filter f (feature_jacobian's return value)
__global__ void k_export_item_dense(Dev d, int f, double* H, double* r, int n_state) {
  const size_t ik = (size_t)f * d.Fcap;
  const unsigned long long mask = d.item_mask[ik];
  const int rows = d.item_rows[ik], M = (rows + 3) / 4, R4 = 4 * M, nc6 = 6 * M;
  const double* W = d.pool + d.item_off[ik];
  const int N = d.ncam[f];
  for (int e = threadIdx.x; e < rows * n_state; e += blockDim.x) H[e] = 0.0;
  __syncthreads();
  int a = 0;
  for (int c = 0; c < N; ++c) {
    if ((mask >> d.cam_phys[(size_t)f * d.Nmax + c]) & 1ull) {
      for (int e = threadIdx.x; e < rows * 6; e += blockDim.x) {
        const int i = e / 6, j = e - 6 * i;
        H[(size_t)i * n_state + NI + 6 * c + j] = W[3 + i + (size_t)(3 + 6 * a + j) * R4];
      }
      ++a;
    }
  }
  for (int i = threadIdx.x; i < rows; i += blockDim.x) r[i] = W[3 + i + (size_t)(3 + nc6) * R4];
}

// gating_test on a dense caller-supplied H (m x n), r (m).  scratch: m*n + m*m + m doubles.
__global__ void __launch_bounds__(256) k_gate_dense(Dev d, int f, const double* __restrict__ H, const double* __restrict__ r, int m,
                                                     int n, int dof, double* scratch, double* out_gamma, int* out_pass) {
  const int tid = threadIdx.x;
  const double* P = d.P + (size_t)f * d.ld * d.ld;
  double* T1 = scratch;                 // m x n row-major
  double* S = T1 + (size_t)m * n;       // m x m row-major
  double* y = S + (size_t)m * m;
  for (int e = tid; e < m * n; e += 256) {
    const int i = e / n, j = e - i * n;
    double acc = 0;
    for (int c = 0; c < n; ++c) acc += H[(size_t)i * n + c] * P[(size_t)c * d.ld + j];
    T1[e] = acc;
  }
  __syncthreads();
  for (int e = tid; e < m * m; e += 256) {
    const int i = e / m, j = e - i * m;
    double acc = 0;
    for (int c = 0; c < n; ++c) acc += T1[(size_t)i * n + c] * H[(size_t)j * n + c];
    if (i == j) acc += d.prm->obs_noise;
    S[e] = acc;
  }
  for (int i = tid; i < m; i += 256) y[i] = r[i];
  __syncthreads();
  __shared__ int s_bad;
  if (tid == 0) s_bad = 0;
  __syncthreads();
  for (int j = 0; j < m; ++j) {
    const double dj = S[(size_t)j * m + j];
    if (!(dj > 0.0)) { if (tid == 0) s_bad = 1; break; }
    const double l = sqrt(dj);
    for (int i = j + 1 + tid; i < m; i += 256) S[(size_t)i * m + j] /= l;
    __syncthreads();
    if (tid == 0) { S[(size_t)j * m + j] = l; y[j] /= l; }
    __syncthreads();
    const double yj = y[j];
    for (int i = j + 1 + tid; i < m; i += 256) {
      const double lij = S[(size_t)i * m + j];
      y[i] -= lij * yj;
      for (int c = j + 1; c <= i; ++c) S[(size_t)i * m + c] -= lij * S[(size_t)c * m + j];
    }
    __syncthreads();
  }
  __syncthreads();
  if (tid == 0) {
    double gamma = 0;
    for (int i = 0; i < m; ++i) gamma += y[i] * y[i];
    if (s_bad) { gamma = INFINITY; atomicOr(d.err, ERRBIT_NOT_SPD); }
    *out_gamma = gamma;
    *out_pass = (gamma < d.prm->chi2[dof]) ? 1 : 0;
  }
}

void launch_jacobian_gate(const Dev& d, cudaStream_t st) {
  dim3 grid((d.Fcap + JAC_WPB - 1) / JAC_WPB, d.F);
  KLAUNCH(d, KID_JACOBIAN_GATE, st, (k_jacobian_gate<<<grid, JAC_WPB * 32, 0, st>>>(d)));
}
void launch_stack(const Dev& d, cudaStream_t st) {
  dim3 grid((d.Fcap + JAC_WPB - 1) / JAC_WPB, d.F);
  KLAUNCH(d, KID_STACK, st, (k_stack<<<grid, JAC_WPB * 32, 0, st>>>(d)));
}
void launch_measurement_jacobian(const Dev& d, int filter, int cam_logical, int slot, double* out, cudaStream_t st) {
  KLAUNCH(d, KID_MISC, st, (k_measurement_jacobian<<<1, 32, 0, st>>>(d, filter, cam_logical, slot, out)));
}
void launch_export_item_dense(const Dev& d, int filter, double* H, double* r, int n_state, cudaStream_t st) {
  KLAUNCH(d, KID_MISC, st, (k_export_item_dense<<<1, 256, 0, st>>>(d, filter, H, r, n_state)));
}
void launch_gate_dense(const Dev& d, int filter, const double* H, const double* r, int m, int n_state, int dof, double* scratch,
                       double* out_gamma, int* out_pass, cudaStream_t st) {
  KLAUNCH(d, KID_MISC, st, (k_gate_dense<<<1, 256, 0, st>>>(d, filter, H, r, m, n_state, dof, scratch, out_gamma, out_pass)));
}

}  // namespace msckf
