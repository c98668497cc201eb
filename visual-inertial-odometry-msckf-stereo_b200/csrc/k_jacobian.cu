// Per-landmark measurement Jacobian, left-null-space projection and chi-square gate:
// measurement_jacobian (msckf.py:429-498), feature_jacobian (:500-541), gating_test (:606-614).
//
// Work comes from a global list of (filter, item) pairs built on the device (k_plan_items); a
// persistent grid walks it: one WARP per landmark with M <= JAC_SMALL_M observations (everything in
// shared memory; the pruning update is always M = 2), one CTA per landmark for longer tracks.
//
// Algebra.  E = blockdiag(H_x,a) (4M x 6M, one 4x6 block per observation), H_f (4M x 3), r (4M).
// The left null space of H_f is spanned by the last 4M-3 columns of Q = H1 H2 H3 (three Householder
// reflectors, compact WY form Q = I - V T V^T) instead of the reference's SVD: any orthonormal basis
// gives the same H^T H, H^T r and gamma.  With Z = T^T V^T E (3 x 6M):
//     Q^T E = E - V Z,   Q^T r = r - V T^T V^T r,          (rows 3.. are the projected H, r)
// so nothing dense of size 4M x 6M is ever formed for the stacked Jacobian: the compact record
// {H_x,a ; V ; Z ; Q^T r} (58 M doubles) goes to the pool and the QR kernel expands rows on the fly.
// The gate gamma = r'^T (H' P H'^T + sigma^2 I)^-1 r' uses the same structure,
//     T1 = (E - V Z) Pcc = E Pcc - V (Z Pcc),     S = T1 E^T - (T1 Z^T) V^T + sigma^2 I,
// O(M^2) block work instead of the dense O(M^3) products, then a Cholesky factorisation (S is SPD by
// construction; the reference's LU gives the same gamma).
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"
#include "dmma_tiles.cuh"

namespace msckf {

constexpr int JAC_WARPS = 4;          // warps per CTA of the small-M kernel
constexpr int JB_T = 256;             // threads per CTA of the long-track kernel
constexpr int JB_YLD = 12;            // leading dimension of the residual tile that rides along the Cholesky factorisation

// Per-stage derived quantities of every camera state, shared by all landmarks it observes (msckf.py:436-447,
// 483-485): R0 = R(q) world -> cam0, R1 = R_01 R0, t0, t1 = t0 - R1^T t_01, R(q_null) g, p_null.
__global__ void __launch_bounds__(64) k_cam_prepare(Dev d) {
  const int f = blockIdx.x;
  const int N = d.ncam[f];
  const Params& prm = *d.prm;
  const double* g = d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY;
  for (int c = threadIdx.x; c < N; c += 64) {
    const double* cs = d.cam + ((size_t)f * d.Nmax + c) * CAM_STRIDE;
    double* out = d.cam_pre + ((size_t)f * d.Nmax + c) * CAM_PRE;
    double R0[9], R1[9], Rn[9], tmp[3], u[3];
    to_rotation(cs, R0);
    mat3_mul(prm.R01, R0, R1);
    mat3t_vec(R1, prm.t01, tmp);
    to_rotation(cs + 7, Rn);
    mat3_vec(Rn, g, u);
    for (int i = 0; i < 9; ++i) { out[i] = R0[i]; out[9 + i] = R1[i]; }
    for (int i = 0; i < 3; ++i) {
      out[18 + i] = cs[4 + i];
      out[21 + i] = cs[4 + i] - tmp[i];
      out[24 + i] = u[i];
      out[27 + i] = cs[11 + i];
    }
  }
}

// msckf.py:429-498 for one (camera state, landmark), one camera of the stereo pair per call: h = 0 gives rows 0-1
// (cam0), h = 1 rows 2-3 (cam1) of H_x (4x6), H_f (4x3), r (4).  pre = CAM_PRE block of the camera state.
__device__ __forceinline__ void measurement_jacobian_half(const double* pre, const double* p_w, const double* z, const double* g,
                                                          const double* R01, int h, double* Hx /*[2][6]*/, double* Hf /*[2][3]*/,
                                                          double* r /*[2]*/) {
  double R0[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) R0[i] = pre[i];
  const double d0[3] = {p_w[0] - pre[18], p_w[1] - pre[19], p_w[2] - pre[20]};
  double pc0[3];
  mat3_vec(R0, d0, pc0);
  double S0[9];
  skew3(pc0, S0);
  double A[12];
  double pcx, pcy, pcz, iz;
  if (h == 0) {
    pcx = pc0[0]; pcy = pc0[1]; pcz = pc0[2];
    // dz/dp rows (only the non-zero 2x3 halves)
    iz = 1 / pcz;                                               // one reciprocal of the depth for the row pair and the residual
    const double a0[3] = {iz, 0.0, -(pcx * iz) * iz};
    const double a1[3] = {0.0, iz, -(pcy * iz) * iz};
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      A[j] = a0[0] * S0[j] + a0[1] * S0[3 + j] + a0[2] * S0[6 + j];
      A[6 + j] = a1[0] * S0[j] + a1[1] * S0[3 + j] + a1[2] * S0[6 + j];
      A[3 + j] = a0[0] * -R0[j] + a0[1] * -R0[3 + j] + a0[2] * -R0[6 + j];
      A[9 + j] = a1[0] * -R0[j] + a1[1] * -R0[3 + j] + a1[2] * -R0[6 + j];
    }
  } else {
    double R1[9], D1[9], pc1[3];
#pragma unroll
    for (int i = 0; i < 9; ++i) R1[i] = pre[9 + i];
    const double d1[3] = {p_w[0] - pre[21], p_w[1] - pre[22], p_w[2] - pre[23]};
    mat3_vec(R1, d1, pc1);
    mat3_mul(R01, S0, D1);
    pcx = pc1[0]; pcy = pc1[1]; pcz = pc1[2];
    iz = 1 / pcz;
    const double b0[3] = {iz, 0.0, -(pcx * iz) * iz};
    const double b1[3] = {0.0, iz, -(pcy * iz) * iz};
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      A[j] = b0[0] * D1[j] + b0[1] * D1[3 + j] + b0[2] * D1[6 + j];
      A[6 + j] = b1[0] * D1[j] + b1[1] * D1[3 + j] + b1[2] * D1[6 + j];
      A[3 + j] = b0[0] * -R1[j] + b0[1] * -R1[3 + j] + b0[2] * -R1[6 + j];
      A[9 + j] = b1[0] * -R1[j] + b1[1] * -R1[3 + j] + b1[2] * -R1[6 + j];
    }
  }
  // observability constraint (msckf.py:483-490)
  double u[6];
  u[0] = pre[24]; u[1] = pre[25]; u[2] = pre[26];
  const double dn[3] = {p_w[0] - pre[27], p_w[1] - pre[28], p_w[2] - pre[29]};
  cross3(dn, g, u + 3);
  double uu = 0;
#pragma unroll
  for (int i = 0; i < 6; ++i) uu += u[i] * u[i];
  const double iuu = 1.0 / uu;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    double au = 0;
#pragma unroll
    for (int j = 0; j < 6; ++j) au += A[6 * i + j] * u[j];
    const double sc = au * iuu;          // (reference: A - (A u) u^T / (u^T u); one reciprocal for both rows)
#pragma unroll
    for (int j = 0; j < 6; ++j) Hx[6 * i + j] = A[6 * i + j] - sc * u[j];
#pragma unroll
    for (int j = 0; j < 3; ++j) Hf[3 * i + j] = -Hx[6 * i + 3 + j];
  }
  r[0] = z[2 * h] - pcx * iz;
  r[1] = z[2 * h + 1] - pcy * iz;
}

// Work buffers of one landmark (shared memory, or global scratch for T1 / S of very long tracks).
struct JacBuf {
  double *Hx;    // [M][24]
  double *V;     // [4M][3]   H_f, overwritten by the Householder vectors
  double *Zt;    // [6M][3]   Z transposed
  double *r;     // [4M]      r, overwritten by Q^T r
  double *PZt;   // [6M][3]   Pcc Z^T
  double *TZ;    // [4M][3]   T1 Z^T
  double *Tm;    // [16]      T (3x3 upper, row-major) and scratch
  double *T1;    // [4M][ldT]
  double *S;     // [R][R]    lower triangle, row-major; y rides in Sy
  double *y;     // [R]
  int *cl;       // [M]       logical camera index of observation a
};
__host__ __device__ inline int jac_small_doubles(int M) {   // everything, for the warp kernel
  const int R = 4 * M - 3;
  return 24 * M + 12 * M + 18 * M + 4 * M + 18 * M + 12 * M + 16 + 4 * M * (6 * M + 1) + R * R + R + (M + 1) / 2 + 1;
}
__host__ __device__ inline int jac_fixed_doubles(int M) {   // without T1 and S
  return 24 * M + 12 * M + 18 * M + 4 * M + 18 * M + 12 * M + 16 + (4 * M - 3) + (M + 1) / 2 + 1;
}
__device__ __forceinline__ void jac_carve(JacBuf& B, double* base, int M, double* big /*nullptr: T1,S follow*/) {
  double* p = base;
  B.Hx = p; p += 24 * M;
  B.V = p; p += 12 * M;
  B.Zt = p; p += 18 * M;
  B.r = p; p += 4 * M;
  B.PZt = p; p += 18 * M;
  B.TZ = p; p += 12 * M;
  B.Tm = p; p += 16;
  B.y = p; p += 4 * M - 3;
  B.cl = (int*)p; p += (M + 1) / 2 + 1;
  double* q = big ? big : p;
  B.T1 = q; q += 4 * M * (6 * M + 1);
  B.S = q;
}

// A group of NT threads works on one landmark: NT = 8 / 16 / 32 lanes of a warp (gmask = their lane mask), or a
// whole CTA (NT > 32).  The leader part (Householder on H_f) runs on the first LW = min(NT, 32) threads.
template <int NT>
__device__ __forceinline__ void gsync_m(unsigned gmask) {
  if (NT <= 32) __syncwarp(gmask); else __syncthreads();
}
template <int LW>
__device__ __forceinline__ double gsum(double v, unsigned m) {
#pragma unroll
  for (int o = LW / 2; o > 0; o >>= 1) v += __shfl_xor_sync(m, v, o);
  return v;
}
#define gsync gsync_m

// Steps 1-4 of a landmark by a group of NT threads: per-observation blocks (msckf.py:429-498), the three Householder
// reflectors on H_f with T and Q^T r, Z = T^T V^T E and Pcc Z^T.
template <int NT>
__device__ __forceinline__ void jac_front(const Dev& d, int f, int k, int M, unsigned long long lmask, const JacBuf& B, int tid,
                                          unsigned gmask_in) {
  constexpr int LW = NT < 32 ? NT : 32;
  const unsigned gmask = NT <= 32 ? gmask_in : 0xffffffffu;
  const int lane = tid;                 // index inside the leader part (only used when tid < LW)
  const size_t ik = (size_t)f * d.Fcap + k;
  const int slot = d.item_slot[ik];
  const size_t fs = (size_t)f * d.Fcap + slot;
  const int N = d.ncam[f];
  const Params& prm = *d.prm;
  const double* g = d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY;
  const int R4 = 4 * M, C6 = 6 * M;
  const int* phys = d.cam_phys + (size_t)f * d.Nmax;

  for (int c = tid; c < N; c += NT)
    if ((lmask >> c) & 1ull) B.cl[__popcll(lmask & ((1ull << c) - 1ull))] = c;
  gsync<NT>(gmask);
  // ---- step 1: per-observation blocks
  {
    const double p_w[3] = {d.f_pos[fs * 3], d.f_pos[fs * 3 + 1], d.f_pos[fs * 3 + 2]};
    for (int t = tid; t < 2 * M; t += NT) {            // (observation, camera of the stereo pair)
      const int a = t >> 1, h = t & 1;
      const int c = B.cl[a];
      const double* pre = d.cam_pre + ((size_t)f * d.Nmax + c) * CAM_PRE;
      const double* z = d.f_obs + (fs * d.Nmax + phys[c]) * 4;
      double Hx[12], Hf[6], r[2];
      measurement_jacobian_half(pre, p_w, z, g, prm.R01, h, Hx, Hf, r);
#pragma unroll
      for (int e = 0; e < 12; ++e) B.Hx[24 * a + 12 * h + e] = Hx[e];
#pragma unroll
      for (int e = 0; e < 6; ++e) B.V[12 * a + 6 * h + e] = Hf[e];
      B.r[4 * a + 2 * h] = r[0];
      B.r[4 * a + 2 * h + 1] = r[1];
    }
  }
  gsync<NT>(gmask);
  // ---- step 2: three Householder reflectors on H_f (warp 0), T of the compact WY form, Q^T r
  if (tid < LW) {
    double* V = B.V;
    double tau[3];
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      double xn2 = 0;
      for (int i = kk + 1 + lane; i < R4; i += LW) { const double x = V[3 * i + kk]; xn2 += x * x; }
      xn2 = gsum<LW>(xn2, gmask);
      const double alpha = V[3 * kk + kk];
      __syncwarp(gmask);
      if (xn2 == 0.0) {
        tau[kk] = 0.0;
      } else {
        const double n2 = alpha * alpha + xn2;
        const double rn = rsqrt(n2);                 // 1 / |beta|
        const double nrm = n2 * rn;
        const double beta = (alpha >= 0) ? -nrm : nrm;
        tau[kk] = (beta - alpha) * ((alpha >= 0) ? -rn : rn);
        const double sc = 1.0 / (alpha - beta);
        for (int i = kk + 1 + lane; i < R4; i += LW) V[3 * i + kk] *= sc;
      }
      if (lane == 0) V[3 * kk + kk] = 1.0;
      __syncwarp(gmask);
      for (int c = kk + 1; c < 3; ++c) {
        double dot = 0;
        for (int i = kk + lane; i < R4; i += LW) dot += V[3 * i + kk] * V[3 * i + c];
        dot = gsum<LW>(dot, gmask) * tau[kk];
        for (int i = kk + lane; i < R4; i += LW) V[3 * i + c] -= dot * V[3 * i + kk];
      }
      __syncwarp(gmask);
    }
    if (lane == 0) { V[1] = 0.0; V[2] = 0.0; V[5] = 0.0; }
    __syncwarp(gmask);
    double d01 = 0, d02 = 0, d12 = 0, y0 = 0, y1 = 0, y2 = 0;
    for (int i = lane; i < R4; i += LW) {
      const double v0 = V[3 * i], v1 = V[3 * i + 1], v2 = V[3 * i + 2], ri = B.r[i];
      d01 += v0 * v1; d02 += v0 * v2; d12 += v1 * v2;
      y0 += v0 * ri; y1 += v1 * ri; y2 += v2 * ri;
    }
    d01 = gsum<LW>(d01, gmask); d02 = gsum<LW>(d02, gmask); d12 = gsum<LW>(d12, gmask);
    y0 = gsum<LW>(y0, gmask); y1 = gsum<LW>(y1, gmask); y2 = gsum<LW>(y2, gmask);
    const double T00 = tau[0], T11 = tau[1], T22 = tau[2];
    const double T01 = -tau[1] * (T00 * d01);
    const double T02 = -tau[2] * (T00 * d02 + T01 * d12);
    const double T12 = -tau[2] * (T11 * d12);
    if (lane == 0) {
      B.Tm[0] = T00; B.Tm[1] = T01; B.Tm[2] = T02; B.Tm[3] = T11; B.Tm[4] = T12; B.Tm[5] = T22;
    }
    const double z0 = T00 * y0, z1 = T01 * y0 + T11 * y1, z2 = T02 * y0 + T12 * y1 + T22 * y2;
    for (int i = lane; i < R4; i += LW) B.r[i] -= V[3 * i] * z0 + V[3 * i + 1] * z1 + V[3 * i + 2] * z2;
  }
  gsync<NT>(gmask);
  // ---- step 3: Z = T^T V^T E, stored transposed
  {
    const double T00 = B.Tm[0], T01 = B.Tm[1], T02 = B.Tm[2], T11 = B.Tm[3], T12 = B.Tm[4], T22 = B.Tm[5];
    for (int c = tid; c < C6; c += NT) {
      const int a = c / 6, e = c - 6 * a;
      double Y0 = 0, Y1 = 0, Y2 = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const double h = B.Hx[24 * a + 6 * i + e];
        const double* v = B.V + 3 * (4 * a + i);
        Y0 += v[0] * h; Y1 += v[1] * h; Y2 += v[2] * h;
      }
      B.Zt[3 * c + 0] = T00 * Y0;
      B.Zt[3 * c + 1] = T01 * Y0 + T11 * Y1;
      B.Zt[3 * c + 2] = T02 * Y0 + T12 * Y1 + T22 * Y2;
    }
  }
  gsync<NT>(gmask);
  // ---- step 4: PZt = Pcc Z^T  (6M x 3)
  const double* P = cov_of(d, f);
  const int ld = d.ld;
  for (int c = tid; c < C6; c += NT) {
    const double* prow = P + (size_t)(NI + 6 * B.cl[c / 6] + (c % 6)) * ld + NI;
    double s0 = 0, s1 = 0, s2 = 0;
    for (int b = 0; b < M; ++b) {
      const double* pp = prow + 6 * B.cl[b];
#pragma unroll
      for (int e = 0; e < 6; ++e) {
        const double pv = pp[e];
        const double* zz = B.Zt + 3 * (6 * b + e);
        s0 += pv * zz[0]; s1 += pv * zz[1]; s2 += pv * zz[2];
      }
    }
    B.PZt[3 * c] = s0; B.PZt[3 * c + 1] = s1; B.PZt[3 * c + 2] = s2;
  }
  gsync<NT>(gmask);
}

// Gate result and the record for the stacking / QR kernel: projected rows (Q^T E - rows 3..) and residual, row-major
// [R][6M+1].  `gamma` is valid on thread 0.
template <int NT>
__device__ __forceinline__ void jac_finish(const Dev& d, int f, int k, int M, unsigned long long lmask, const JacBuf& B, int tid,
                                           int write_all, unsigned gmask, double gamma, bool spd) {
  const size_t ik = (size_t)f * d.Fcap + k;
  const int C6 = 6 * M, R = 4 * M - 3;
  if (tid == 0) {
    // S is positive definite in exact arithmetic; a landmark whose position is wildly off (a gross outlier) can make
    // H P H^T so large that rounding produces a non-positive pivot.  The reference's LU (msckf.py:609) still returns
    // some gamma there: the landmark is counted and its gate is redone by k_gate_lu from the record written below.
    if (!spd) {
      atomicAdd((unsigned long long*)(d.counters + (size_t)NCNT * f + MSCKF_CNT_GATE_NOT_SPD), 1ull);
      queue_gate_lu(d.wl_lu, d.stage_ctr, f, k);
    }
    const int pass = (gamma < d.prm->chi2[d.item_dof[ik]]) ? 1 : 0;
    d.item_gamma[ik] = gamma;
    d.item_pass[ik] = pass;
    B.Tm[8] = (double)(pass || !spd);
  }
  gsync<NT>(gmask);
  const int pass = (int)B.Tm[8];
  if ((pass || write_all) && d.item_off[ik] + jac_record_doubles(M) <= d.pool_cap) {     // (pool overflow is reported by k_plan_items)
    double* out = d.pool + d.item_off[ik];
    const int ldr = C6 + 1;
    for (int e = tid; e < R * ldr; e += NT) {
      const int ii = e / ldr, c = e - ii * ldr, i = ii + 3;
      double v;
      if (c == C6) {
        v = B.r[i];
      } else {
        const int a = c / 6;
        v = (i >= 4 * a && i < 4 * a + 4) ? B.Hx[24 * a + 6 * (i - 4 * a) + (c - 6 * a)] : 0.0;
        v -= B.V[3 * i] * B.Zt[3 * c] + B.V[3 * i + 1] * B.Zt[3 * c + 1] + B.V[3 * i + 2] * B.Zt[3 * c + 2];
      }
      out[e] = v;
    }
  }
  gsync<NT>(gmask);
}

// One landmark through Jacobian, projection and gate by a group of NT <= 32 lanes (short tracks, everything in shared memory).
template <int NT>
__device__ void jac_core(const Dev& d, int f, int k, int M, unsigned long long lmask, const JacBuf& B, int tid, int write_all,
                         unsigned gmask_in) {
  constexpr int LW = NT < 32 ? NT : 32;
  const unsigned gmask = NT <= 32 ? gmask_in : 0xffffffffu;
  const int lane = tid;
  const Params& prm = *d.prm;
  const int R4 = 4 * M, C6 = 6 * M, R = R4 - 3, ldT = C6 + 1;
  jac_front<NT>(d, f, k, M, lmask, B, tid, gmask_in);
  const double* P = cov_of(d, f);
  const int ld = d.ld;
  // ---- step 5: T1 = E Pcc - V PZt^T   (one (block a, column c) 4-vector per thread; P is exactly symmetric)
  for (int e2 = tid; e2 < M * C6; e2 += NT) {
    const int a = e2 / C6, c = e2 - a * C6;
    const double* pp = P + (size_t)(NI + 6 * B.cl[c / 6] + (c % 6)) * ld + NI + 6 * B.cl[a];
    double p6[6];
#pragma unroll
    for (int e = 0; e < 6; ++e) p6[e] = pp[e];
    const double q0 = B.PZt[3 * c], q1 = B.PZt[3 * c + 1], q2 = B.PZt[3 * c + 2];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const double* h = B.Hx + 24 * a + 6 * i;
      const double* v = B.V + 3 * (4 * a + i);
      double s = 0;
#pragma unroll
      for (int e = 0; e < 6; ++e) s += h[e] * p6[e];
      B.T1[(4 * a + i) * ldT + c] = s - (v[0] * q0 + v[1] * q1 + v[2] * q2);
    }
  }
  gsync<NT>(gmask);
  // ---- step 6: TZ = T1 Z^T  (4M x 3)
  for (int e2 = tid; e2 < R4 * 3; e2 += NT) {
    const int i = e2 / 3, kk = e2 - 3 * i;
    double s = 0;
    for (int c = 0; c < C6; ++c) s += B.T1[i * ldT + c] * B.Zt[3 * c + kk];
    B.TZ[3 * i + kk] = s;
  }
  gsync<NT>(gmask);
  // ---- step 7: S = T1 E^T - TZ V^T + sigma^2 I on rows / columns 3.. (lower triangle), y = projected residual
  for (int e2 = tid; e2 < R * M; e2 += NT) {          // (row i, observation block b): up to 4 columns
    const int ii = e2 / M, b = e2 - ii * M;
    const int i = ii + 3;
    double t6[6];
#pragma unroll
    for (int e = 0; e < 6; ++e) t6[e] = B.T1[i * ldT + 6 * b + e];
    const double z0 = B.TZ[3 * i], z1 = B.TZ[3 * i + 1], z2 = B.TZ[3 * i + 2];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = 4 * b + jj;
      if (j >= 3 && j <= i) {
        const double* h = B.Hx + 24 * b + 6 * jj;
        const double* v = B.V + 3 * j;
        double s = 0;
#pragma unroll
        for (int e = 0; e < 6; ++e) s += t6[e] * h[e];
        s -= z0 * v[0] + z1 * v[1] + z2 * v[2];
        if (i == j) s += prm.obs_noise;
        B.S[ii * R + (j - 3)] = s;
      }
    }
  }
  for (int i = tid; i < R; i += NT) B.y[i] = B.r[i + 3];
  gsync<NT>(gmask);
  // ---- step 8: Cholesky (right-looking), forward substitution riding along
  bool spd = true;
  for (int j = 0; j < R; ++j) {
    const double dj = B.S[j * R + j];
    if (!(dj > 0.0)) { spd = false; break; }
    const double il = rsqrt(dj);                    // only 1 / L(j,j) is needed: one reciprocal square root, no divides
    const double yj = B.y[j] * il;
    gsync<NT>(gmask);
    for (int i = j + 1 + tid; i < R; i += NT) B.S[i * R + j] *= il;
    if (tid == 0) B.y[j] = yj;
    gsync<NT>(gmask);
    for (int i = j + 1 + tid; i < R; i += NT) {
      const double lij = B.S[i * R + j];
      B.y[i] -= lij * yj;
      for (int c = j + 1; c <= i; ++c) B.S[i * R + c] -= lij * B.S[c * R + j];
    }
    gsync<NT>(gmask);
  }
  double gamma = 0;
  if (tid < LW) {
    for (int i = lane; i < R; i += LW) gamma += B.y[i] * B.y[i];
    gamma = gsum<LW>(gamma, gmask);
  }
  if (!spd) gamma = INFINITY;
  jac_finish<NT>(d, f, k, M, lmask, B, tid, write_all, gmask, gamma, spd);
}

// ------------------------------------------------------------------ gate by LU (msckf.py:606-614 as written)
// gamma = r^T S^-1 r through Gaussian elimination with partial pivoting (what np.linalg.solve / dgesv does) for an
// n x n system in global memory (row-major, ld = n); S and y are overwritten, x receives the solution.  One CTA.
__device__ double lu_gamma(double* S, int n, double* y, double* x, const double* r0, int tid, int nthr, double* s_val, int* s_idx) {
  for (int j = 0; j < n; ++j) {
    double best = -1.0;
    int bi = j;
    for (int i = j + tid; i < n; i += nthr) {
      const double a = fabs(S[(size_t)i * n + j]);
      if (a > best) { best = a; bi = i; }
    }
    s_val[tid] = best;
    s_idx[tid] = bi;
    __syncthreads();
    if (tid == 0) {
      for (int t = 1; t < nthr; ++t)
        if (s_val[t] > s_val[0] || (s_val[t] == s_val[0] && s_idx[t] < s_idx[0])) { s_val[0] = s_val[t]; s_idx[0] = s_idx[t]; }
    }
    __syncthreads();
    const int p = s_idx[0];
    if (p != j) {
      for (int c = tid; c < n; c += nthr) {
        const double a = S[(size_t)j * n + c];
        S[(size_t)j * n + c] = S[(size_t)p * n + c];
        S[(size_t)p * n + c] = a;
      }
      if (tid == 0) { const double a = y[j]; y[j] = y[p]; y[p] = a; }
    }
    __syncthreads();
    const double piv = S[(size_t)j * n + j], yj = y[j];
    for (int i = j + 1 + tid; i < n; i += nthr) {
      const double l = S[(size_t)i * n + j] / piv;
      y[i] -= l * yj;
      for (int c = j + 1; c < n; ++c) S[(size_t)i * n + c] -= l * S[(size_t)j * n + c];
    }
    __syncthreads();
  }
  for (int i = n - 1; i >= 0; --i) {            // back substitution, one row at a time
    double s = 0.0;
    for (int c = i + 1 + tid; c < n; c += nthr) s += S[(size_t)i * n + c] * x[c];
    s_val[tid] = s;
    __syncthreads();
    if (tid == 0) {
      double t = 0.0;
      for (int q = 0; q < nthr; ++q) t += s_val[q];
      x[i] = (y[i] - t) / S[(size_t)i * n + i];
    }
    __syncthreads();
  }
  double g = 0.0;
  if (tid == 0)
    for (int i = 0; i < n; ++i) g += r0[i] * x[i];
  return g;                                      // valid on thread 0
}

// Redoes the gate of the landmarks queued in wl_lu from their records (projected rows H' [R][6M] and residual r'):
// T1 = H' Pcc, S = T1 H'^T + sigma^2 I, gamma by LU.  Rare (a gross outlier among ~1e8 gates): one CTA walks the list.
__global__ void __launch_bounds__(256) k_gate_lu(Dev d) {
  __shared__ double s_val[256];
  __shared__ int s_idx[256];
  __shared__ int s_cl[64];
  const int tid = threadIdx.x;
  const int n = (int)d.stage_ctr[SC_NLU];
  for (int w = 0; w < n; ++w) {
    const int2 it = d.wl_lu[w];
    const int f = it.x, k = it.y;
    const size_t ik = (size_t)f * d.Fcap + k;
    const int R = d.item_rows[ik], M = (R + 3) / 4, C6 = 6 * M, ldr = C6 + 1, ld = d.ld;
    const unsigned long long lmask = d.item_lmask[ik];
    if (d.item_off[ik] + jac_record_doubles(M) > d.pool_cap) continue;
    const double* rec = d.pool + d.item_off[ik];
    const double* P = cov_of(d, f);
    double* T1 = d.lu_scratch;                  // [R][C6]
    double* S = T1 + (size_t)R * C6;            // [R][R]
    double* y = S + (size_t)R * R;              // [R]
    double* x = y + R;                          // [R]
    __syncthreads();
    for (int c = tid; c < 64; c += 256)
      if ((lmask >> c) & 1ull) s_cl[__popcll(lmask & ((1ull << c) - 1ull))] = c;
    __syncthreads();
    for (int e = tid; e < R * C6; e += 256) {
      const int i = e / C6, c = e - i * C6;
      const double* pcol = P + (size_t)(NI + 6 * s_cl[c / 6] + c % 6) * ld + NI;     // row of P = column (P is symmetric)
      double acc = 0.0;
      for (int b = 0; b < M; ++b)
        for (int e6 = 0; e6 < 6; ++e6) acc += rec[(size_t)i * ldr + 6 * b + e6] * pcol[6 * s_cl[b] + e6];
      T1[e] = acc;
    }
    __syncthreads();
    for (int e = tid; e < R * R; e += 256) {
      const int i = e / R, j = e - i * R;
      double acc = (i == j) ? d.prm->obs_noise : 0.0;
      for (int c = 0; c < C6; ++c) acc += T1[(size_t)i * C6 + c] * rec[(size_t)j * ldr + c];
      S[e] = acc;
    }
    for (int i = tid; i < R; i += 256) { y[i] = rec[(size_t)i * ldr + C6]; x[i] = 0.0; }
    __syncthreads();
    // r0 = the residual column of the record (strided): gather it behind x
    double* r0 = x + R;
    for (int i = tid; i < R; i += 256) r0[i] = rec[(size_t)i * ldr + C6];
    __syncthreads();
    const double gamma = lu_gamma(S, R, y, x, r0, tid, 256, s_val, s_idx);
    if (tid == 0) {
      d.item_gamma[ik] = gamma;
      d.item_pass[ik] = (gamma < d.prm->chi2[d.item_dof[ik]]) ? 1 : 0;    // (NaN compares false: rejected)
    }
  }
}

// Small tracks: one group of G lanes per landmark, persistent grid over the work list wl_jac_small[0 .. n).
template <int G>
__global__ void __launch_bounds__(JAC_WARPS * 32, G == 32 ? 3 : 8) k_jacobian_gate_small(Dev d, int cls, int per_group_doubles, int write_all) {
  extern __shared__ double smem[];
  constexpr int GPW = 32 / G;                       // groups per warp
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = lane / G, tid = lane - grp * G, gbase = grp * G;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << gbase);
  const int n = (int)d.stage_ctr[SC_NJAC_SMALL + cls];
  const int2* wl = d.wl_jac_small + (size_t)cls * d.F * d.Fcap;
  double* base = smem + (size_t)(warp * GPW + grp) * per_group_doubles;
  const int stride = gridDim.x * JAC_WARPS * GPW;
  for (int w = (blockIdx.x * JAC_WARPS + warp) * GPW + grp; w < n; w += stride) {
    const int2 it = wl[w];
    const size_t ik = (size_t)it.x * d.Fcap + it.y;
    const unsigned long long lmask = d.item_lmask[ik];          // logical camera order (k_plan_items)
    const int M = __popcll(lmask);
    JacBuf B;
    jac_carve(B, base, M, nullptr);
    jac_core<G>(d, it.x, it.y, M, lmask, B, tid, write_all, gmask);
  }
}

// ------------------------------------------------------------------ long tracks (M > JAC_SMALL_M): one CTA per landmark
// The same algebra as jac_core, organised for tracks of up to 64 observations (R = 4M - 3 up to 253 gate rows):
//   * T1 = (E - V Z) Pcc is never stored: a group of lanes owns a row block a (the 4 rows of one observation), each lane a
//     column block b; the lane forms the 4 x 6 block T1_ab = Hx_a Pcc[a,b] - V_a (Pcc Z^T)_b^T in registers, adds its share
//     of (T1 Z^T)_a (reduced over the group by shuffles) and, for b <= a, the 4 x 4 block S_ab = T1_ab Hx_b^T - (T1 Z^T)_a V_b^T;
//   * S = L L^T by a right-looking blocked Cholesky on the FP64 tensor pipe (8-row panels: the diagonal block is factored and
//     inverted by one warp, panel solves and trailing updates are DMMA tiles), the projected residual riding along as one
//     more column tile, gamma = |L^-1 r'|^2;
//   * S lives in shared memory when it fits (tracks of up to ~38 observations), else in a per-CTA global scratch (L2).
// Two launches share the code: tracks of up to JAC_BIG_SPLIT_M observations (small S, several CTAs per SM) and longer ones.
__host__ __device__ inline int jb_rp(int M) { return (4 * M - 3 + 7) & ~7; }                       // gate rows padded to 8
__host__ __device__ inline int jb_lds(int M) { return jb_rp(M) + 4; }                               // 4 (mod 8): conflict-free fragments
__host__ __device__ inline long long jb_S_doubles(int M) { return (long long)jb_rp(M) * jb_lds(M); }
__host__ __device__ inline int jb_fixed_doubles(int M) {                                           // everything but S
  return ((jac_fixed_doubles(M) + 1) & ~1) + jb_rp(M) * JB_YLD + 8 * LDL;
}

template <int GS, int MAXR>
__device__ __forceinline__ void jb_build_S(const double* __restrict__ P, int ld, int M, const JacBuf& B, double* S, int ldS,
                                           double sigma2, int tid) {
  constexpr int NG = JB_T / GS;
  const int grp = tid / GS, gl = tid - grp * GS;
  const unsigned gmask = GS == 32 ? 0xffffffffu : (0xffffu << (16 * ((tid >> 4) & 1)));
  for (int a = grp; a < M; a += NG) {
    double hxa[24], va[12], tz[12], s16[MAXR][16];
#pragma unroll
    for (int e = 0; e < 24; ++e) hxa[e] = B.Hx[24 * a + e];
#pragma unroll
    for (int e = 0; e < 12; ++e) { va[e] = B.V[12 * a + e]; tz[e] = 0.0; }
    const double* prow = P + (size_t)(NI + 6 * B.cl[a]) * ld + NI;
#pragma unroll
    for (int rr = 0; rr < MAXR; ++rr) {
      const int b = rr * GS + gl;
      if (b < M) {
        const double* pp = prow + 6 * B.cl[b];
        double t1[24];
#pragma unroll
        for (int e = 0; e < 24; ++e) t1[e] = 0.0;
#pragma unroll
        for (int e = 0; e < 6; ++e) {
          double p6[6];
#pragma unroll
          for (int c = 0; c < 6; ++c) p6[c] = pp[(size_t)e * ld + c];
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int c = 0; c < 6; ++c) t1[6 * i + c] += hxa[6 * i + e] * p6[c];
        }
        const double* pz = B.PZt + 18 * b;
        const double* zt = B.Zt + 18 * b;
#pragma unroll
        for (int c = 0; c < 6; ++c) {
          const double q0 = pz[3 * c], q1 = pz[3 * c + 1], q2 = pz[3 * c + 2];
          const double z0 = zt[3 * c], z1 = zt[3 * c + 1], z2 = zt[3 * c + 2];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const double t = t1[6 * i + c] - (va[3 * i] * q0 + va[3 * i + 1] * q1 + va[3 * i + 2] * q2);
            t1[6 * i + c] = t;
            tz[3 * i] += t * z0; tz[3 * i + 1] += t * z1; tz[3 * i + 2] += t * z2;
          }
        }
        if (b <= a) {
          const double* hxb = B.Hx + 24 * b;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            double h6[6];
#pragma unroll
            for (int c = 0; c < 6; ++c) h6[c] = hxb[6 * j + c];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              double acc = 0.0;
#pragma unroll
              for (int c = 0; c < 6; ++c) acc += t1[6 * i + c] * h6[c];
              s16[rr][4 * i + j] = acc;
            }
          }
        }
      }
    }
#pragma unroll
    for (int o = GS / 2; o > 0; o >>= 1) {
#pragma unroll
      for (int e = 0; e < 12; ++e) tz[e] += __shfl_xor_sync(gmask, tz[e], o);
    }
#pragma unroll
    for (int rr = 0; rr < MAXR; ++rr) {
      const int b = rr * GS + gl;
      if (b < M && b <= a) {
        const double* vb = B.V + 12 * b;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const double v0 = vb[3 * j], v1 = vb[3 * j + 1], v2 = vb[3 * j + 2];
          const int col = 4 * b + j - 3;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int row = 4 * a + i - 3;
            if (row >= 0 && col >= 0) {
              double v = s16[rr][4 * i + j] - (tz[3 * i] * v0 + tz[3 * i + 1] * v1 + tz[3 * i + 2] * v2);
              if (row == col) v += sigma2;
              S[(size_t)row * ldS + col] = v;
            }
          }
        }
      }
    }
  }
}

// S (lower tiles, nkt x nkt tiles of 8) = L L^T in place, Y <- L^-1 Y.  Returns true when a pivot is not positive.
__device__ __forceinline__ bool jb_cholesky(double* S, int ldS, int nkt, double* sY, double* sLinv, int* s_bad, int tid) {
  constexpr int NW = JB_T / 32;
  const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3;
  for (int p = 0; p < nkt; ++p) {
    const int p0 = 8 * p;
    if (warp == 0) {
      if (chol8_inverse(S, ldS, p0, sLinv, lane) && lane == 0) *s_bad = 1;
    }
    __syncthreads();
    if (*s_bad) return true;
    const double li0 = sLinv[g * LDL + tg], li1 = sLinv[g * LDL + 4 + tg];   // A fragment of L^-1 == B fragment of L^-T
    const int nbelow = nkt - p - 1;
    // panel: Y_p = L_pp^-1 Y_p, L_ip = S_ip L_pp^-T (row tiles below)
    for (int item = warp; item < 1 + nbelow; item += NW) {
      double c0 = 0.0, c1 = 0.0;
      if (item == 0) {
        double* base = sY + p0 * JB_YLD;
        const double b0 = base[tg * JB_YLD + g], b1 = base[(4 + tg) * JB_YLD + g];
        dmma884(c0, c1, li0, b0);
        dmma884(c0, c1, li1, b1);
        base[g * JB_YLD + 2 * tg] = c0;
        base[g * JB_YLD + 2 * tg + 1] = c1;
      } else {
        double* base = S + (size_t)(8 * (p + item)) * ldS + p0;
        const double a0 = base[(size_t)g * ldS + tg], a1 = base[(size_t)g * ldS + 4 + tg];
        dmma884(c0, c1, a0, li0);
        dmma884(c0, c1, a1, li1);
        *(double2*)(base + (size_t)g * ldS + 2 * tg) = make_double2(c0, c1);
      }
    }
    __syncthreads();
    // trailing update: Y_i -= L_ip Y_p ;  S_ij -= L_ip L_jp^T  (p < j <= i)
    const int nS = nbelow * (nbelow + 1) / 2;
    for (int item = warp; item < nbelow; item += NW) {
      const int it = p + 1 + item;
      const double* lrow = S + (size_t)(8 * it + g) * ldS + p0 + tg;
      const double a0 = -lrow[0], a1 = -lrow[4];
      const double* wb = sY + p0 * JB_YLD;
      const double b0 = wb[tg * JB_YLD + g], b1 = wb[(4 + tg) * JB_YLD + g];
      double* cp = sY + (8 * it + g) * JB_YLD + 2 * tg;
      double c0 = cp[0], c1 = cp[1];
      dmma884(c0, c1, a0, b0);
      dmma884(c0, c1, a1, b1);
      cp[0] = c0;
      cp[1] = c1;
    }
    if (warp < nS) {
      int ib, jb;
      tri_decode(warp, ib, jb);
      for (int item = warp; item < nS; item += NW) {
        const int it = p + 1 + ib, jt = p + 1 + jb;
        const double* lrow = S + (size_t)(8 * it + g) * ldS + p0 + tg;
        const double a0 = -lrow[0], a1 = -lrow[4];
        const double* lcol = S + (size_t)(8 * jt + g) * ldS + p0 + tg;
        const double b0 = lcol[0], b1 = lcol[4];
        double2* cp = (double2*)(S + (size_t)(8 * it + g) * ldS + 8 * jt + 2 * tg);
        double2 c = *cp;
        dmma884(c.x, c.y, a0, b0);
        dmma884(c.x, c.y, a1, b1);
        *cp = c;
        tri_advance(ib, jb, NW);
      }
    }
    __syncthreads();
  }
  return false;
}

// cls 0: tracks of JAC_SMALL_M < M <= JAC_BIG_SPLIT_M (work list wl_jac_big), cls 1: longer tracks (wl_jac_big2)
template <int GS, int MAXR, int MINB>
__global__ void __launch_bounds__(JB_T, MINB) k_jacobian_gate_big(Dev d, int cls, int smem_doubles, int write_all) {
  extern __shared__ __align__(16) double smem[];
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31;
  const int n = (int)d.stage_ctr[cls == 0 ? SC_NJAC_BIG : SC_NJAC_BIG2];
  const int2* wl = cls == 0 ? d.wl_jac_big : d.wl_jac_big2;
  const double sigma2 = d.prm->obs_noise;
  for (int w = blockIdx.x; w < n; w += gridDim.x) {
    const int2 it = wl[w];
    const int f = it.x, k = it.y;
    const size_t ik = (size_t)f * d.Fcap + k;
    const unsigned long long lmask = d.item_lmask[ik];          // logical camera order (k_plan_items)
    const int M = __popcll(lmask);
    const int R = 4 * M - 3, Rp = jb_rp(M), nkt = Rp >> 3;
    JacBuf B;
    jac_carve(B, smem, M, smem);                       // (T1 / S pointers of the short-track layout are not used here)
    double* sY = smem + ((jac_fixed_doubles(M) + 1) & ~1);
    double* sLinv = sY + Rp * JB_YLD;
    const bool fits = jb_fixed_doubles(M) + jb_S_doubles(M) <= smem_doubles;
    double* S = fits ? sLinv + 8 * LDL : d.jac_scratch + (size_t)blockIdx.x * d.jac_scratch_stride;
    const int ldS = jb_lds(M);
    __syncthreads();                                    // the previous landmark's record has been written
    if (tid == 0) s_bad = 0;
    jac_front<JB_T>(d, f, k, M, lmask, B, tid, 0xffffffffu);
    // S: zero, unit diagonal on the padding rows; Y: the projected residual in column 0
    for (int e = tid; e < Rp * (ldS / 2); e += JB_T) ((double2*)S)[e] = make_double2(0.0, 0.0);
    for (int e = tid; e < Rp * JB_YLD; e += JB_T) {
      const int i = e / JB_YLD, c = e - i * JB_YLD;
      sY[e] = (c == 0 && i < R) ? B.r[i + 3] : 0.0;
    }
    __syncthreads();
    for (int i = R + tid; i < Rp; i += JB_T) S[(size_t)i * ldS + i] = 1.0;
    jb_build_S<GS, MAXR>(cov_of(d, f), d.ld, M, B, S, ldS, sigma2, tid);
    __syncthreads();
    const bool bad = jb_cholesky(S, ldS, nkt, sY, sLinv, &s_bad, tid);
    double gamma = 0.0;
    if (tid < 32) {
      for (int i = lane; i < R; i += 32) gamma += sY[i * JB_YLD] * sY[i * JB_YLD];
      gamma = gsum<32>(gamma, 0xffffffffu);
    }
    if (bad) gamma = INFINITY;
    jac_finish<JB_T>(d, f, k, M, lmask, B, tid, write_all, 0xffffffffu, gamma, !bad);
  }
}

// ------------------------------------------------------------------ API helpers (single filter)
__global__ void k_measurement_jacobian(Dev d, int f, int c, int slot, double* out) {
  if (threadIdx.x != 0) return;
  const size_t fs = (size_t)f * d.Fcap + slot;
  const double* pre = d.cam_pre + ((size_t)f * d.Nmax + c) * CAM_PRE;
  const double* z = d.f_obs + (fs * d.Nmax + d.cam_phys[(size_t)f * d.Nmax + c]) * 4;
  const double* g = d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY;
  for (int h = 0; h < 2; ++h)
    measurement_jacobian_half(pre, d.f_pos + fs * 3, z, g, d.prm->R01, h, out + 12 * h, out + 24 + 6 * h, out + 36 + 2 * h);
}

// dense (rows x n_state) view of item 0 of filter f (feature_jacobian's return value), scattered from its record
__global__ void k_export_item_dense(Dev d, int f, double* H, double* r, int n_state) {
  const size_t ik = (size_t)f * d.Fcap;
  const unsigned long long lmask = d.item_lmask[ik];
  const int rows = d.item_rows[ik], M = (rows + 3) / 4, C6 = 6 * M, ldr = C6 + 1;
  const double* rec = d.pool + d.item_off[ik];
  const int N = d.ncam[f];
  for (int e = threadIdx.x; e < rows * n_state; e += blockDim.x) H[e] = 0.0;
  __syncthreads();
  for (int e = threadIdx.x; e < rows * C6; e += blockDim.x) {
    const int ii = e / C6, c = e - ii * C6, a = c / 6, e6 = c - 6 * a;
    int cam = -1, seen = 0;           // logical camera of observation a
    for (int cc = 0; cc < N; ++cc)
      if ((lmask >> cc) & 1ull) { if (seen == a) { cam = cc; break; } ++seen; }
    H[(size_t)ii * n_state + NI + 6 * cam + e6] = rec[ii * ldr + c];
  }
  for (int i = threadIdx.x; i < rows; i += blockDim.x) r[i] = rec[i * ldr + C6];
}

// gating_test on a dense caller-supplied H (m x n), r (m).  scratch: m*n + m*m + m doubles.
__global__ void __launch_bounds__(256) k_gate_dense(Dev d, int f, const double* __restrict__ H, const double* __restrict__ r, int m,
                                                     int n, int dof, double* scratch, double* out_gamma, int* out_pass) {
  const int tid = threadIdx.x;
  const double* P = cov_of(d, f);
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
    if (s_bad) gamma = INFINITY;      // rejected, see jac_core
    *out_gamma = gamma;
    *out_pass = (gamma < d.prm->chi2[dof]) ? 1 : 0;
  }
}

int sm_count() {                        // of the current device (contexts of one process may sit on different GPUs)
  static std::atomic<int> cache[64];
  int dev = 0;
  cudaGetDevice(&dev);
  int n = cache[dev & 63].load(std::memory_order_relaxed);
  if (!n) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    cache[dev & 63].store(n, std::memory_order_relaxed);
  }
  return n;
}

template <int G>
static void launch_small(const Dev& d, int cls, int write_all, cudaStream_t st) {
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_jacobian_gate_small<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  int per_group = (jac_small_doubles(jac_class_maxM(cls)) + 15) & ~15;
  if (G < 32) per_group += 8;      // neighbouring groups of a warp land on different halves of the banks
  const size_t smem = (size_t)JAC_WARPS * (32 / G) * per_group * sizeof(double);
  int per_sm = (int)std::min<size_t>(8, (220 * 1024) / std::max<size_t>(smem + 1024, 1));
  per_sm = std::max(per_sm, 1);
  KLAUNCH(d, KID_JACOBIAN_GATE, st, (k_jacobian_gate_small<G><<<sm_count() * per_sm, JAC_WARPS * 32, smem, st>>>(d, cls, per_group, write_all)));
}

// class_mask: which size classes of the group-per-landmark kernel can occur in this stage (bit c = class c; the
// pruning stage only has M = 2).  Lanes per landmark grow with the class so that the shared-memory-limited number of
// landmarks in flight still fills the SM with warps.
void launch_jacobian_gate(const Dev& d, int class_mask, int with_big, int write_all, cudaStream_t st) {
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_jacobian_gate_big<16, 1, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(k_jacobian_gate_big<32, 2, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  cudaStream_t main_st = st;
  {
  Fork fk(d, st);
  if (class_mask & 1) {
    if (d.jac_rows_ok) launch_jacobian_rows(d, 0, write_all, fk.lane(0));
    else launch_small<8>(d, 0, write_all, fk.lane(0));
  }
  if (class_mask & 2) {
    if (d.jac_rows_ok) launch_jacobian_rows(d, 1, write_all, fk.lane(1));
    else launch_small<16>(d, 1, write_all, fk.lane(1));
  }
  if (class_mask & 4) {
    if (d.jac_rows_ok) launch_jacobian_rows(d, 2, write_all, fk.lane(2));
    else launch_small<32>(d, 2, write_all, fk.lane(2));
  }
  if (class_mask & 8) launch_small<32>(d, 3, write_all, fk.lane(3));
  if (with_big) {
    st = fk.lane(4);
    const int sms = sm_count();
    {   // tracks of up to JAC_BIG_SPLIT_M observations: S always fits, two CTAs per SM
      const int Mm = std::min(d.Nmax, JAC_BIG_SPLIT_M);
      const int sd = (int)(jb_fixed_doubles(Mm) + jb_S_doubles(Mm));
      KLAUNCH(d, KID_JACOBIAN_BIG, st,
              (k_jacobian_gate_big<16, 1, 2><<<std::min(d.jac_big_grid, 2 * sms), JB_T, (size_t)sd * sizeof(double), st>>>(d, 0, sd, write_all)));
    }
    if (d.Nmax > JAC_BIG_SPLIT_M) {   // longer tracks: one CTA per SM, S in shared memory up to ~38 observations, else in L2
      const long long want = jb_fixed_doubles(d.Nmax) + jb_S_doubles(d.Nmax);
      const int sd = (int)std::min<long long>(want, (200 * 1024) / 8);
      KLAUNCH(d, KID_JACOBIAN_BIG, st,
              (k_jacobian_gate_big<32, 2, 1><<<std::min(d.jac_big_grid, sms), JB_T, (size_t)sd * sizeof(double), st>>>(d, 1, sd, write_all)));
    }
  }
  }
  launch_gate_lu(d, main_st);          // after the join: gates whose innovation covariance was not positive definite to rounding
}
void launch_gate_lu(const Dev& d, cudaStream_t st) { KLAUNCH(d, KID_MISC, st, (k_gate_lu<<<1, 256, 0, st>>>(d))); }
void launch_cam_prepare(const Dev& d, cudaStream_t st) { KLAUNCH(d, KID_CAM_PREPARE, st, (k_cam_prepare<<<d.F, 64, 0, st>>>(d))); }
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
