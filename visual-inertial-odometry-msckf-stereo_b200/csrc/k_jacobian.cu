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

namespace msckf {

constexpr int JAC_WARPS = 4;          // warps per CTA of the small-M kernel
constexpr int JAC_BIG_T = 128;        // threads per CTA of the long-track kernel

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
    const double sc = au / uu;           // (reference: A - (A u) u^T / (u^T u); one division per row instead of six)
    for (int j = 0; j < 6; ++j) Hx[6 * i + j] = A[6 * i + j] - sc * u[j];
    for (int j = 0; j < 3; ++j) Hf[3 * i + j] = -Hx[6 * i + 3 + j];
  }
  r[0] = z[0] - pc0[0] / pc0[2];
  r[1] = z[1] - pc0[1] / pc0[2];
  r[2] = z[2] - pc1[0] / pc1[2];
  r[3] = z[3] - pc1[1] / pc1[2];
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

// One landmark through Jacobian, projection and gate by a group of NT threads.
template <int NT>
__device__ void jac_core(const Dev& d, int f, int k, int M, unsigned long long lmask, const JacBuf& B, int tid, int write_all,
                         unsigned gmask_in) {
  constexpr int LW = NT < 32 ? NT : 32;
  const unsigned gmask = NT <= 32 ? gmask_in : 0xffffffffu;
  const int lane = tid;                 // index inside the leader part (only used when tid < LW)
  const size_t ik = (size_t)f * d.Fcap + k;
  const int slot = d.cand[(size_t)f * d.Fcap + d.item_cand[ik]];
  const size_t fs = (size_t)f * d.Fcap + slot;
  const int N = d.ncam[f];
  const Params& prm = *d.prm;
  const double* g = d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY;
  const int R4 = 4 * M, C6 = 6 * M, R = R4 - 3, ldT = C6 + 1;
  const int* phys = d.cam_phys + (size_t)f * d.Nmax;

  for (int c = tid; c < N; c += NT)
    if ((lmask >> c) & 1ull) B.cl[__popcll(lmask & ((1ull << c) - 1ull))] = c;
  gsync<NT>(gmask);
  // ---- step 1: per-observation blocks
  {
    const double p_w[3] = {d.f_pos[fs * 3], d.f_pos[fs * 3 + 1], d.f_pos[fs * 3 + 2]};
    for (int a = tid; a < M; a += NT) {
      const int c = B.cl[a];
      const double* cs = d.cam + ((size_t)f * d.Nmax + c) * CAM_STRIDE;
      const double* z = d.f_obs + (fs * d.Nmax + phys[c]) * 4;
      double Hx[24], Hf[12], r[4];
      measurement_jacobian(cs, p_w, z, g, prm.R01, prm.t01, Hx, Hf, r);
#pragma unroll
      for (int e = 0; e < 24; ++e) B.Hx[24 * a + e] = Hx[e];
#pragma unroll
      for (int e = 0; e < 12; ++e) B.V[12 * a + e] = Hf[e];
#pragma unroll
      for (int e = 0; e < 4; ++e) B.r[4 * a + e] = r[e];
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
        const double nrm = sqrt(alpha * alpha + xn2);
        const double beta = (alpha >= 0) ? -nrm : nrm;
        tau[kk] = (beta - alpha) / beta;
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
  const double* P = d.P + (size_t)f * d.ld * d.ld;
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
    const double l = sqrt(dj);
    const double yj = B.y[j] / l;
    gsync<NT>(gmask);
    for (int i = j + 1 + tid; i < R; i += NT) B.S[i * R + j] /= l;
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
  int pass = 0;
  if (tid == 0) {
    // S is positive definite in exact arithmetic; a landmark whose position is wildly off (a gross outlier) can make
    // H P H^T so large that rounding produces a non-positive pivot.  The reference's LU returns noise there; the
    // landmark is rejected (gamma = inf) and counted, it is not an error of the filter.
    if (!spd) atomicAdd((unsigned long long*)(d.counters + (size_t)NCNT * f + MSCKF_CNT_GATE_NOT_SPD), 1ull);
    pass = (gamma < prm.chi2[d.item_dof[ik]]) ? 1 : 0;
    d.item_gamma[ik] = gamma;
    d.item_pass[ik] = pass;
    d.item_lmask[ik] = lmask;
    B.Tm[8] = (double)pass;
  }
  gsync<NT>(gmask);
  pass = (int)B.Tm[8];
  // ---- compact record for the stacking / QR kernel: Hx[M][24] V[4M][3] Zt[6M][3] Q^T r[4M]  (contiguous in B)
  if ((pass || write_all) && d.item_off[ik] + 58 * M <= d.pool_cap) {     // (pool overflow is reported by k_plan_items)
    double* out = d.pool + d.item_off[ik];
    for (int e = tid; e < 58 * M; e += NT) out[e] = B.Hx[e];
  }
  gsync<NT>(gmask);
}

// logical-order observation mask of a landmark, computed by a group of G lanes (gmask, first lane gbase)
template <int G>
__device__ __forceinline__ unsigned long long logical_mask(const Dev& d, int f, unsigned long long mask, int tid, unsigned gmask, int gbase) {
  const int N = d.ncam[f];
  unsigned long long lmask = 0;
  for (int c0 = 0; c0 < N; c0 += G) {
    const int c = c0 + tid;
    const int bit = (c < N) ? (int)((mask >> d.cam_phys[(size_t)f * d.Nmax + c]) & 1ull) : 0;
    unsigned b = __ballot_sync(gmask, bit) >> gbase;
    if (G < 32) b &= (1u << G) - 1u;
    lmask |= (unsigned long long)b << c0;
  }
  return lmask;
}

// Small tracks: one group of G lanes per landmark, persistent grid over the work list wl_jac_small[0 .. n).
template <int G>
__global__ void __launch_bounds__(JAC_WARPS * 32, 4) k_jacobian_gate_small(Dev d, int per_group_doubles, int write_all) {
  extern __shared__ double smem[];
  constexpr int GPW = 32 / G;                       // groups per warp
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = lane / G, tid = lane - grp * G, gbase = grp * G;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << gbase);
  const int n = (int)d.stage_ctr[SC_NJAC_SMALL];
  double* base = smem + (size_t)(warp * GPW + grp) * per_group_doubles;
  const int stride = gridDim.x * JAC_WARPS * GPW;
  for (int w = (blockIdx.x * JAC_WARPS + warp) * GPW + grp; w < n; w += stride) {
    const int2 it = d.wl_jac_small[w];
    const size_t ik = (size_t)it.x * d.Fcap + it.y;
    const unsigned long long lmask = logical_mask<G>(d, it.x, d.item_mask[ik], tid, gmask, gbase);
    const int M = __popcll(lmask);
    JacBuf B;
    jac_carve(B, base, M, nullptr);
    jac_core<G>(d, it.x, it.y, M, lmask, B, tid, write_all, gmask);
  }
}

// Long tracks: one CTA per landmark; T1 and S live in shared memory when they fit, else in global scratch.
__global__ void __launch_bounds__(JAC_BIG_T) k_jacobian_gate_big(Dev d, int smem_doubles, int write_all) {
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31;
  const int n = (int)d.stage_ctr[SC_NJAC_BIG];
  for (int w = blockIdx.x; w < n; w += gridDim.x) {
    const int2 it = d.wl_jac_big[w];
    const size_t ik = (size_t)it.x * d.Fcap + it.y;
    const unsigned long long lmask = logical_mask<32>(d, it.x, d.item_mask[ik], lane, 0xffffffffu, 0);
    const int M = __popcll(lmask);
    JacBuf B;
    const bool fits = jac_small_doubles(M) <= smem_doubles;
    jac_carve(B, smem, M, fits ? nullptr : d.jac_scratch + (size_t)blockIdx.x * d.jac_scratch_stride);
    jac_core<JAC_BIG_T>(d, it.x, it.y, M, lmask, B, threadIdx.x, write_all, 0xffffffffu);
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

// dense (rows x n_state) view of item 0 of filter f (feature_jacobian's return value), expanded from the compact record
__global__ void k_export_item_dense(Dev d, int f, double* H, double* r, int n_state) {
  const size_t ik = (size_t)f * d.Fcap;
  const unsigned long long lmask = d.item_lmask[ik];
  const int rows = d.item_rows[ik], M = (rows + 3) / 4, C6 = 6 * M;
  const double* rec = d.pool + d.item_off[ik];
  const double *Hx = rec, *V = rec + 24 * M, *Zt = rec + 36 * M, *rq = rec + 54 * M;
  const int N = d.ncam[f];
  for (int e = threadIdx.x; e < rows * n_state; e += blockDim.x) H[e] = 0.0;
  __syncthreads();
  for (int e = threadIdx.x; e < rows * C6; e += blockDim.x) {
    const int ii = e / C6, c = e - ii * C6, a = c / 6, e6 = c - 6 * a, i = ii + 3;
    // logical camera of observation a
    int cam = -1, seen = 0;
    for (int cc = 0; cc < N; ++cc)
      if ((lmask >> cc) & 1ull) { if (seen == a) { cam = cc; break; } ++seen; }
    double v = (i >= 4 * a && i < 4 * a + 4) ? Hx[24 * a + 6 * (i - 4 * a) + e6] : 0.0;
    v -= V[3 * i] * Zt[3 * c] + V[3 * i + 1] * Zt[3 * c + 1] + V[3 * i + 2] * Zt[3 * c + 2];
    H[(size_t)ii * n_state + NI + 6 * cam + e6] = v;
  }
  for (int i = threadIdx.x; i < rows; i += blockDim.x) r[i] = rq[i + 3];
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
    if (s_bad) gamma = INFINITY;      // rejected, see jac_core
    *out_gamma = gamma;
    *out_pass = (gamma < d.prm->chi2[dof]) ? 1 : 0;
  }
}

static int g_sm_count = 0;
int sm_count() {
  if (!g_sm_count) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

template <int G>
static void launch_small(const Dev& d, int max_small_M, int write_all, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_jacobian_gate_small<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set = true;
  }
  const int per_group = (jac_small_doubles(max_small_M) + 1) & ~1;
  const size_t smem = (size_t)JAC_WARPS * (32 / G) * per_group * sizeof(double);
  int per_sm = (int)std::min<size_t>(4, (200 * 1024) / std::max<size_t>(smem + 1024, 1));
  per_sm = std::max(per_sm, 1);
  KLAUNCH(d, KID_JACOBIAN_GATE, st, (k_jacobian_gate_small<G><<<sm_count() * per_sm, JAC_WARPS * 32, smem, st>>>(d, per_group, write_all)));
}

// max_small_M: largest M the group kernel can meet in this stage (2 for the pruning stage: 8 lanes per landmark).
void launch_jacobian_gate(const Dev& d, int max_small_M, int write_all, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_jacobian_gate_big, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr_set = true;
  }
  if (max_small_M <= 2) launch_small<8>(d, max_small_M, write_all, st);
  else launch_small<16>(d, max_small_M, write_all, st);
  if (max_small_M >= JAC_SMALL_M) {     // the stage can contain long tracks
    const int smem_doubles = d.jac_big_smem_doubles;
    KLAUNCH(d, KID_JACOBIAN_BIG, st,
            (k_jacobian_gate_big<<<d.jac_big_grid, JAC_BIG_T, (size_t)smem_doubles * sizeof(double), st>>>(d, smem_doubles, write_all)));
  }
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
