// IMU propagation (process_model / predict_new_state, msckf.py:291-380, jit_utils.py:6-135),
// state augmentation (msckf.py:385-407, jit_utils.py:137-167), camera-state removal
// (msckf.py:803-818) and online reset (msckf.py:820-886).  One CTA per filter.
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int PT = 256;   // threads per CTA for the per-filter kernels in this file

// ------------------------------------------------------------------ RK4 state prediction (thread 0)
// jit_utils.py:46-128, including the stale-skew quirk: the skew matrix built from the
// normalised dq(dt) (lines 80-84) is re-used for R(dq(dt/2)) (line 94) and R(q) (line 104).
// Coefficients of the two quaternion propagators of one gyro sample (jit_utils.py:62-78): dq(dt) = (cA I + sA Omega) q,
// dq(dt/2) = (cB I + sB Omega) q.  They depend on the sample only, so the samples of a chunk compute them in parallel
// and the sequential chain below starts from them.
__device__ __forceinline__ void quat_step_coeffs(double dt, const double* gyro, double* out /*[4]: cA sA cB sB*/) {
  const double gn = norm3(gyro);
  const double c1 = cos(gn * dt * 0.5), c2 = cos(gn * dt * 0.25);
  out[0] = c1;
  out[2] = c2;
  if (gn > 1e-5) {
    out[1] = sin(gn * dt * 0.5) / gn;
    out[3] = sin(gn * dt * 0.25) / gn;
  } else {                                              // first-order branch: cos(.) (I + Omega dt / 2)
    out[1] = c1 * (dt * 0.5);
    out[3] = c2 * (dt * 0.25);
  }
}
// 1 / |q| through one reciprocal square root (MUFU.RSQ64H + Newton) instead of a square root and a divide: the
// normalisations sit on the sequential part of the RK4 chain.  Differs from sqrt-then-divide by at most an ulp.
__device__ __forceinline__ double rnorm4(const double* q) { return rsqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]); }
__device__ __forceinline__ void to_rotation_rs(const double* q_in, double* R) {       // utils.py:14-27
  const double rn = rnorm4(q_in);
  const double q[4] = {q_in[0] * rn, q_in[1] * rn, q_in[2] * rn, q_in[3] * rn};
  rot_from_parts(q, q[3], q, R);
}
__device__ void predict_new_state(double dt, const double* gyro, const double* acc, const double* cf, const double* g,
                                  double* q, double* v, double* p) {
  const double Om[16] = {0, gyro[2], -gyro[1], gyro[0],
                         -gyro[2], 0, gyro[0], gyro[1],
                         gyro[1], -gyro[0], 0, gyro[2],
                         -gyro[0], -gyro[1], -gyro[2], 0};
  double dq[4], dq2[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double oq = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) oq += Om[4 * i + j] * q[j];
    dq[i] = cf[0] * q[i] + cf[1] * oq;
    dq2[i] = cf[2] * q[i] + cf[3] * oq;
  }
  double n = rnorm4(dq);
  for (int i = 0; i < 4; ++i) dq[i] *= n;
  const double stale[3] = {dq[0], dq[1], dq[2]};
  double RdT[9], Rd2T[9], R[9];
  rot_from_parts(dq, dq[3], stale, RdT);            // dR_dt (transposed when applied)
  n = rnorm4(dq2);
  for (int i = 0; i < 4; ++i) dq2[i] *= n;
  rot_from_parts(dq2, dq2[3], stale, Rd2T);
  n = rnorm4(q);
  const double qn[4] = {q[0] * n, q[1] * n, q[2] * n, q[3] * n};
  rot_from_parts(qn, qn[3], stale, R);

  double k1v[3], k2v[3], k4v[3], tmp[3];
  mat3t_vec(R, acc, tmp);                            // R^T acc  (R.T @ acc)
  for (int i = 0; i < 3; ++i) k1v[i] = tmp[i] + g[i];
  // dR_dt2_transpose @ acc with dR_dt2_transpose = (...)^T  ->  (Rd2T)^T acc
  mat3t_vec(Rd2T, acc, tmp);
  for (int i = 0; i < 3; ++i) k2v[i] = tmp[i] + g[i];   // k3_v_dot is the same expression
  mat3t_vec(RdT, acc, tmp);
  for (int i = 0; i < 3; ++i) k4v[i] = tmp[i] + g[i];
  double k1p[3], k2p[3], k3p[3], k4p[3];
  for (int i = 0; i < 3; ++i) {
    k1p[i] = v[i];
    k2p[i] = v[i] + k1v[i] * dt / 2.;
    k3p[i] = v[i] + k2v[i] * dt / 2;
    k4p[i] = v[i] + k2v[i] * dt;
  }
  n = rnorm4(dq);
  for (int i = 0; i < 4; ++i) q[i] = dq[i] * n;
  for (int i = 0; i < 3; ++i) {
    const double vi = v[i];
    v[i] = vi + (k1v[i] + 2 * k2v[i] + 2 * k2v[i] + k4v[i]) * dt / 6.;
    p[i] = p[i] + (k1p[i] + 2 * k2p[i] + 2 * k3p[i] + k4p[i]) * dt / 6.;
  }
}

// Sparsity of Phi = I + F dt + (F dt)^2/2 + (F dt)^3/6 (jit_utils.py:10-41): only the theta (0:3), v (6:9) and p (12:15)
// block rows differ from the identity, theta rows in columns 0:6, v and p rows in columns 0:15.  Products of such
// matrices keep the pattern, so every 21 x 21 product below only visits those entries (the skipped terms are exact
// zeros of the dense products the reference forms).
__device__ __forceinline__ bool phi_identity_row(int i) { return !(i < 3 || (i >= 6 && i < 9) || (i >= 12 && i < 15)); }
__device__ __forceinline__ int phi_row_width(int i) { return i < 3 ? 6 : 15; }

constexpr int PW = 2;                         // warps per CTA (one CTA per filter): small CTAs so that 7 fit on an SM and 1024 filters are
                                              // all resident at once - the kernel is as long as one filter's serial RK4 chain, not two waves of it
constexpr int PKC = 10;                       // IMU samples per chunk
constexpr int PHC = 84;                       // doubles of one compact Phi: 8 blocks, dt, sigma_a^2 R^T R
constexpr int PST = 24;                       // doubles of one staged sample / state: q v p | dt w a | cA sA cB sB
constexpr int PLD = 28;                       // leading dimension of the padded 24 x 24 matrices: 12 (mod 16) doubles, conflict-free DMMA fragments
constexpr int PMAT = 24 * PLD;
constexpr int PROP_DOUBLES = 4 * PMAT + PKC * PHC + (PKC + 1) * PST + IMU_STRIDE + 12;

__device__ __forceinline__ void pdmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Compact Phi of one IMU sample (lane-parallel over the samples of a chunk): jit_utils.py:6-43 followed by the
// observability-constrained blocks of msckf.py:328-344.  x0 = state before the sample (q v p), x1 = after,
// xn = null state (q v p) the constraint refers to.
__device__ void build_phi(const double* x0, const double* x1, const double* qn, const double* vn, const double* pn, const double* smp,
                          const double* g, double Qc_a, double* out) {
  const double dt = smp[0];
  const double* w = smp + 1;
  const double* a = smp + 4;
  double R[9], Sa[9], A00[9], A20[9], A23[9], RtS[9], Sw[9];
  to_rotation_rs(x0, R);
  skew3(a, Sa);
  mat3_mul_tn(R, Sa, RtS);
  skew3(w, Sw);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      A00[3 * i + j] = -Sw[3 * i + j] * dt;
      A20[3 * i + j] = -RtS[3 * i + j] * dt;
      A23[3 * i + j] = -R[3 * j + i] * dt;
    }
  const double mdt = -1.0 * dt, pdt = 1.0 * dt;
  constexpr double i6 = 1.0 / 6.0;            // the reference divides by 6 (jit_utils.py:41); a product differs by at most an ulp
  // (F dt)^2 and (F dt)^3, block by block (the skipped terms of the dense products are exact zeros)
  double B00[9], B20[9], C00[9], C20[9], C40[9], P00[9], P20[9], P40[9];
  mat3_mul(A00, A00, B00);                                   // sq[0,0] = A00 A00
  mat3_mul(A20, A00, B20);                                   // sq[2,0] = A20 A00;  sq[4,0] = dt A20, sq[4,3] = dt A23
  mat3_mul(B00, A00, C00);                                   // cube[0,0] = sq[0,0] A00
  mat3_mul(B20, A00, C20);                                   // cube[2,0] = sq[2,0] A00
  double B40[9];
  for (int e = 0; e < 9; ++e) B40[e] = pdt * A20[e];
  mat3_mul(B40, A00, C40);                                   // cube[4,0] = sq[4,0] A00
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const int e = 3 * i + j;
      const double I = (i == j) ? 1.0 : 0.0;
      P00[e] = (I + A00[e]) + B00[e] / 2. + C00[e] * i6;
      out[9 + e] = (0.0 + I * mdt) + (A00[e] * mdt) / 2. + (B00[e] * mdt) * i6;     // Phi01
      P20[e] = (0.0 + A20[e]) + B20[e] / 2. + C20[e] * i6;
      out[27 + e] = (A20[e] * mdt) / 2. + (B20[e] * mdt) * i6;                       // Phi21
      out[36 + e] = A23[e];                                                            // Phi23
      P40[e] = B40[e] / 2. + C40[e] * i6;
      out[54 + e] = (B40[e] * mdt) * i6;                                               // Phi41
      out[63 + e] = (pdt * A23[e]) / 2.;                                               // Phi43
    }
  out[72] = dt;
  // G Qc G^T, block (6:9,6:9) = (-R^T) sigma_a^2 (-R^T)^T
  {
    double nR[9], nRq[9];
    for (int i = 0; i < 9; ++i) nR[i] = -R[i];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) nRq[3 * i + j] = nR[3 * j + i] * Qc_a;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        out[73 + 3 * i + j] = nRq[3 * i] * nR[3 * 0 + j] + nRq[3 * i + 1] * nR[3 * 1 + j] + nRq[3 * i + 2] * nR[3 * 2 + j];
  }
  // observability-constrained blocks (msckf.py:328-344)
  double Rk[9], Rn[9];
  to_rotation_rs(qn, Rk);
  to_rotation_rs(x1, Rn);
  mat3_mul_nt(Rn, Rk, P00);
  double u[3], sv[3];
  mat3_vec(Rk, g, u);
  const double uu = u[0] * u[0] + u[1] * u[1] + u[2] * u[2];
  const double iuu = 1.0 / uu;
  for (int i = 0; i < 3; ++i) sv[i] = u[i] * iuu;
  double dv[3], dp[3], w1[3], w2[3];
  for (int i = 0; i < 3; ++i) {
    dv[i] = vn[i] - x1[4 + i];
    dp[i] = dt * vn[i] + pn[i] - x1[7 + i];
  }
  cross3(dv, g, w1);
  cross3(dp, g, w2);
  for (int i = 0; i < 3; ++i) {
    double* r2 = P20 + 3 * i;
    const double c2 = (r2[0] * u[0] + r2[1] * u[1] + r2[2] * u[2]) - w1[i];
    for (int j = 0; j < 3; ++j) r2[j] = r2[j] - c2 * sv[j];
    double* r4 = P40 + 3 * i;
    const double c4 = (r4[0] * u[0] + r4[1] * u[1] + r4[2] * u[2]) - w2[i];
    for (int j = 0; j < 3; ++j) r4[j] = r4[j] - c4 * sv[j];
  }
  for (int e = 0; e < 9; ++e) { out[e] = P00[e]; out[18 + e] = P20[e]; out[45 + e] = P40[e]; }
}

// One CTA of PW warps per filter integrates the filter's IMU samples of this frame, PKC samples at a time:
//   1. thread 0 runs the (sequential) RK4 state chain over the chunk,
//   2. thread k builds Phi_k of sample k from the states before / after it (samples in parallel),
//   3. the CTA advances the 21x21 IMU block of P sample by sample and accumulates the product of the Phis.
// The IMU-camera cross covariance only needs that product, applied once at the end (identical algebra:
// P_IC <- Phi_k ... Phi_1 P_IC; symmetrising it is exact).
__global__ void __launch_bounds__(PW * 32, 7) k_propagate(Dev d, const int* __restrict__ imu_off, const double* __restrict__ ts,
                                                        const double* __restrict__ gyro_m, const double* __restrict__ acc_m, const FrameIn* fin) {
  extern __shared__ double psm[];
  if (fin) { imu_off = fin->imu_off; ts = fin->ts; gyro_m = fin->gyro; acc_m = fin->acc; }
  constexpr int NT = PW * 32;
  const int lane = threadIdx.x;            // index inside the filter's CTA
  const int wrp = lane >> 5, g = (lane & 31) >> 2, tg = lane & 3;   // warp, DMMA fragment coordinates
  const int f = blockIdx.x;
  if (f >= d.F) return;
  const int k0 = imu_off[f], k1 = imu_off[f + 1];
  if (k1 <= k0) return;
  double* sPhi = psm;                      // Phi of the current sample, 24 x 24 zero padded, ld = PLD
  double* sPII = sPhi + PMAT;              // IMU block of P (+ dt G Qc G^T)
  double* sT = sPII + PMAT;                // Phi (P_II + dt G Qc G^T)
  double* sAcc = sT + PMAT;                // product of the Phis of this frame
  double* sPc = sAcc + PMAT;               // [PKC][PHC] compact Phis
  double* sSt = sPc + PKC * PHC;           // [PKC+1][PST]: state after sample c-1 (q v p) | sample c (dt w a)
  double* sx = sSt + (PKC + 1) * PST;
  const Params& prm = *d.prm;
  double* imu = d.imu + (size_t)f * IMU_STRIDE;
  double* P = cov_of(d, f);
  const int ld = d.ld;
  const int n = NI + 6 * d.ncam[f];

  for (int e = lane; e < IMU_STRIDE; e += NT) sx[e] = imu[e];
  for (int e = lane; e < PMAT; e += NT) {
    const int i = e / PLD, j = e - PLD * i;
    sPII[e] = (i < 21 && j < 21) ? P[(size_t)i * ld + j] : 0.0;
    sAcc[e] = (i == j && i < 21) ? 1.0 : 0.0;
    sPhi[e] = (i == j && i < 21) ? 1.0 : 0.0;           // every sample rewrites the same 8 blocks + the dt diagonal; the rest stays I
  }
  __syncthreads();

  for (int c0 = k0; c0 < k1; c0 += PKC) {
    const int kc = min(PKC, k1 - c0);
    // ---- stage the samples; state slot 0 = current state
    if (lane < kc) {
      const int k = c0 + lane;
      double* smp = sSt + lane * PST + 10;
      smp[0] = ts[k] - (lane == 0 ? sx[MSCKF_IMU_TIME] : ts[k - 1]);      // dt of this sample
      for (int i = 0; i < 3; ++i) {
        smp[1 + i] = gyro_m[3 * k + i] - sx[MSCKF_IMU_BG + i];
        smp[4 + i] = acc_m[3 * k + i] - sx[MSCKF_IMU_BA + i];
      }
      quat_step_coeffs(smp[0], smp + 1, smp + 7);
    }
    if (lane < 4) sSt[lane] = sx[MSCKF_IMU_Q + lane];
    if (lane < 3) { sSt[4 + lane] = sx[MSCKF_IMU_V + lane]; sSt[7 + lane] = sx[MSCKF_IMU_P + lane]; }
    __syncthreads();
    // ---- 1. RK4 chain
    if (lane == 0) {
      double q[4], v[3], p[3];
      for (int i = 0; i < 4; ++i) q[i] = sSt[i];
      for (int i = 0; i < 3; ++i) { v[i] = sSt[4 + i]; p[i] = sSt[7 + i]; }
      for (int c = 0; c < kc; ++c) {
        const double* smp = sSt + c * PST + 10;
        predict_new_state(smp[0], smp + 1, smp + 4, smp + 7, sx + MSCKF_IMU_GRAVITY, q, v, p);
        double* nx = sSt + (c + 1) * PST;
        for (int i = 0; i < 4; ++i) nx[i] = q[i];
        for (int i = 0; i < 3; ++i) { nx[4 + i] = v[i]; nx[7 + i] = p[i]; }
      }
      sx[MSCKF_IMU_TIME] = ts[c0 + kc - 1];
    }
    __syncthreads();
    // ---- 2. Phi of every sample of the chunk (null state of the first sample = the stored one, then the previous state)
    if (lane < kc) {
      const double* x0 = sSt + lane * PST;
      const double* x1 = sSt + (lane + 1) * PST;
      const bool first = (lane == 0);
      build_phi(x0, x1, first ? sx + MSCKF_IMU_Q_NULL : x0, first ? sx + MSCKF_IMU_V_NULL : x0 + 4,
                first ? sx + MSCKF_IMU_P_NULL : x0 + 7, x0 + 10, sx + MSCKF_IMU_GRAVITY, prm.Qc[2], sPc + lane * PHC);
    }
    __syncthreads();
    // state after the chunk; null states <- current (msckf.py:366-368)
    {
      const double* xe = sSt + kc * PST;
      if (lane < 4) { sx[MSCKF_IMU_Q + lane] = xe[lane]; sx[MSCKF_IMU_Q_NULL + lane] = xe[lane]; }
      if (lane < 3) {
        sx[MSCKF_IMU_V + lane] = xe[4 + lane]; sx[MSCKF_IMU_V_NULL + lane] = xe[4 + lane];
        sx[MSCKF_IMU_P + lane] = xe[7 + lane]; sx[MSCKF_IMU_P_NULL + lane] = xe[7 + lane];
      }
      if (lane == 0) sx[MSCKF_IMU_NULL_ALIASED] = 1.0;
    }
    // ---- 3. covariance, sample by sample, on DMMA tiles (3 x 3 tiles of 8 x 8, K = 24):
    //   P_II <- sym( Phi (P_II + dt G Qc G^T) Phi^T ),  Phi_acc <- Phi Phi_acc            (jit_utils.py:130-135, msckf.py:346-362)
    // G Qc G^T = diag(sg I, sbg I, sa R^T R, sba I, 0): R^T R is the identity to rounding, so the noise goes onto the diagonal.
    for (int c = 0; c < kc; ++c) {
      const double* pc = sPc + c * PHC;
      const double dt = pc[72];
      if (lane < 12) sPII[lane * PLD + lane] += dt * prm.Qc[lane / 3];
      for (int e = lane; e < 72; e += NT) {
        const int blk = e / 9, r = (e % 9) / 3, cc = e % 3;
        // block -> (row block, column block): 00 01 20 21 23 40 41 43
        const int rb = blk < 2 ? 0 : (blk < 5 ? 6 : 12);
        const int cb = (blk == 0 || blk == 2 || blk == 5) ? 0 : ((blk == 1 || blk == 3 || blk == 6) ? 3 : 9);
        sPhi[PLD * (rb + r) + cb + cc] = pc[e];
      }
      if (lane < 3) sPhi[PLD * (12 + lane) + 6 + lane] = 1.0 * dt;
      __syncthreads();
      // T = Phi M (tiles 0..8) and Phi_acc' = Phi Phi_acc (tiles 9..17, kept in registers until the old one has been read)
      constexpr int UT = (18 + PW - 1) / PW, U0 = (10 - PW + PW - 1) / PW, UX = (9 + PW - 1) / PW;   // job rounds; first round with a Phi_acc tile
      // all accumulators first, stores afterwards: the rounds are independent DMMA chains the scheduler can interleave
      double ac0[UT], ac1[UT];
#pragma unroll
      for (int u = 0; u < UT; ++u) {
        const int job = wrp + PW * u;
        ac0[u] = 0.0;
        ac1[u] = 0.0;
        if (job < 18) {
          const int t = job % 9, it = t / 3, jt = t - 3 * it;
          const double* B = (job < 9) ? sPII : sAcc;
          if (it < 2) {
            // rows 0..15 of Phi have nothing right of column 15 (phi_row_width): four 4-deep steps instead of six
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              pdmma(ac0[u], ac1[u], sPhi[(8 * it + g) * PLD + 4 * ks + tg], B[(4 * ks + tg) * PLD + 8 * jt + g]);
          } else {
            // rows 16..20 of Phi are identity rows (extrinsics), 21..23 padding: the product tile is a copy of B's tile
            ac0[u] = B[(16 + g) * PLD + 8 * jt + 2 * tg];
            ac1[u] = B[(16 + g) * PLD + 8 * jt + 2 * tg + 1];
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UT; ++u) {
        const int job = wrp + PW * u;
        if (job < 9) {
          const int it = job / 3, jt = job - 3 * it;
          sT[(8 * it + g) * PLD + 8 * jt + 2 * tg] = ac0[u];
          sT[(8 * it + g) * PLD + 8 * jt + 2 * tg + 1] = ac1[u];
        }
      }
      __syncthreads();
      // P_II <- (X + X^T) / 2 with X = T Phi^T: a tile of X^T = Phi T^T is the same products in the same order as the mirrored
      // tile of X, so both halves of the sum come out of this warp's own DMMAs and P_II stays exactly symmetric.
      // Phi_acc <- Phi_acc'
      double xc0[UX], xc1[UX], xe0[UX], xe1[UX];
#pragma unroll
      for (int u = 0; u < UX; ++u) {
        const int t = wrp + PW * u;
        xc0[u] = xc1[u] = xe0[u] = xe1[u] = 0.0;
        if (t < 9) {
          const int it = t / 3, jt = t - 3 * it;
          // X = T Phi^T and its mirror Phi T^T: a tile whose Phi operand is the identity tile row (16..23) is a copy of T
          if (jt < 2) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) pdmma(xc0[u], xc1[u], sT[(8 * it + g) * PLD + 4 * ks + tg], sPhi[(8 * jt + g) * PLD + 4 * ks + tg]);
          } else {
            xc0[u] = (16 + 2 * tg < 21) ? sT[(8 * it + g) * PLD + 16 + 2 * tg] : 0.0;
            xc1[u] = (17 + 2 * tg < 21) ? sT[(8 * it + g) * PLD + 17 + 2 * tg] : 0.0;
          }
          if (it < 2) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) pdmma(xe0[u], xe1[u], sPhi[(8 * it + g) * PLD + 4 * ks + tg], sT[(8 * jt + g) * PLD + 4 * ks + tg]);
          } else {
            xe0[u] = (16 + g < 21) ? sT[(8 * jt + 2 * tg) * PLD + 16 + g] : 0.0;
            xe1[u] = (16 + g < 21) ? sT[(8 * jt + 2 * tg + 1) * PLD + 16 + g] : 0.0;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UX; ++u) {
        const int t = wrp + PW * u;
        if (t < 9) {
          const int it = t / 3, jt = t - 3 * it;
          sPII[(8 * it + g) * PLD + 8 * jt + 2 * tg] = (xc0[u] + xe0[u]) / 2.;
          sPII[(8 * it + g) * PLD + 8 * jt + 2 * tg + 1] = (xc1[u] + xe1[u]) / 2.;
        }
      }
#pragma unroll
      for (int u = U0; u < UT; ++u) {
        const int job = wrp + PW * u;
        if (job >= 9 && job < 18) {
          const int t = job - 9, it = t / 3, jt = t - 3 * it;
          sAcc[(8 * it + g) * PLD + 8 * jt + 2 * tg] = ac0[u];
          sAcc[(8 * it + g) * PLD + 8 * jt + 2 * tg + 1] = ac1[u];
        }
      }
      __syncthreads();
    }
  }

  // ---- write back: IMU state, P_II, P_IC = Phi_acc P_IC and its mirror
  // (the tracking rate belongs to add_feature_observations, which runs beside this kernel inside a frame)
  for (int e = lane; e < IMU_STRIDE; e += NT)
    if (e != MSCKF_IMU_TRACKING_RATE) imu[e] = sx[e];
  if (lane == 0) d.counters[NCNT * f + MSCKF_CNT_IMU_STEPS] += k1 - k0;
  for (int e = lane; e < 441; e += NT) {
    const int i = e / 21, j = e - 21 * i;
    P[(size_t)i * ld + j] = sPII[i * PLD + j];
  }
  // P_IC <- Phi_acc P_IC and its mirror: a warp owns an 8-column tile (its B fragments are read before anything is written)
  const int nct = (n - NI + 7) >> 3;
  double bf[6], bn[6];
  auto load_b = [&](int jt, double* o) {
    const int j = NI + 8 * jt + g;
#pragma unroll
    for (int ks = 0; ks < 6; ++ks) {
      const int k = 4 * ks + tg;
      o[ks] = (jt < nct && k < NI && j < n) ? P[(size_t)k * ld + j] : 0.0;
    }
  };
  load_b(wrp, bf);
  for (int jt = wrp; jt < nct; jt += PW) {
    load_b(jt + PW, bn);                                   // next tile's fragments are in flight while this one is multiplied
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      double c0v = 0.0, c1v = 0.0;
#pragma unroll
      for (int ks = 0; ks < 6; ++ks) pdmma(c0v, c1v, sAcc[(8 * it + g) * PLD + 4 * ks + tg], bf[ks]);
      const int i = 8 * it + g, jj = NI + 8 * jt + 2 * tg;
      if (i < NI) {
        if (jj < n) { P[(size_t)i * ld + jj] = c0v; P[(size_t)jj * ld + i] = c0v; }
        if (jj + 1 < n) { P[(size_t)i * ld + jj + 1] = c1v; P[(size_t)(jj + 1) * ld + i] = c1v; }
      }
    }
#pragma unroll
    for (int ks = 0; ks < 6; ++ks) bf[ks] = bn[ks];
  }
}

// ------------------------------------------------------------------ state augmentation
// Where the camera state of this frame will sit: its index in the window and the physical observation slot (lowest free
// bit).  Neither depends on the propagation, so add_feature_observations (k_observe) can run beside k_propagate / k_augment
// inside a frame; both read the slot from here.
__global__ void __launch_bounds__(256) k_cam_slot(Dev d) {
  const int f = blockIdx.x * 256 + threadIdx.x;
  if (f >= d.F) return;
  const int N = d.ncam[f];
  if (N >= d.Nmax) {
    atomicOr(d.err, ERRBIT_CAM_CAP);
    d.obs_N[f] = N;
    d.obs_phys[f] = -1;
    return;
  }
  d.obs_N[f] = N + 1;
  d.obs_phys[f] = __ffsll((long long)~d.phys_used[f]) - 1;
}

// msckf.py:385-407 + jit_utils.py:137-167.  One CTA per filter.
__global__ void __launch_bounds__(PT) k_augment(Dev d, const long long* __restrict__ new_ids, const FrameIn* fin) {
  if (fin) new_ids = fin->cam_ids;
  const int f = blockIdx.x, tid = threadIdx.x;
  const int N = d.ncam[f];
  if (N >= d.Nmax || d.obs_phys[f] < 0) return;     // window full: reported by k_cam_slot
  __shared__ double sJ[6 * 21];
  __shared__ double sB[6 * 21];   // J P[:21,:21]
  const double* imu = d.imu + (size_t)f * IMU_STRIDE;
  double* P = cov_of(d, f);
  const int ld = d.ld, n = NI + 6 * N;
  if (tid == 0) {
    double Rwi[9], Rwc[9], tw[3], q[4];
    const double* Ric = imu + MSCKF_IMU_R_IC;
    const double* tci = imu + MSCKF_IMU_T_CI;
    to_rotation(imu + MSCKF_IMU_Q, Rwi);
    mat3_mul(Ric, Rwi, Rwc);
    mat3t_vec(Rwi, tci, tw);                      // R_w_i^T t_c_i
    to_quaternion(Rwc, q);
    double* c = d.cam + ((size_t)f * d.Nmax + N) * CAM_STRIDE;
    for (int i = 0; i < 4; ++i) { c[i] = q[i]; c[7 + i] = q[i]; }
    for (int i = 0; i < 3; ++i) { const double pi = imu[MSCKF_IMU_P + i] + tw[i]; c[4 + i] = pi; c[11 + i] = pi; }
    d.cam_id[(size_t)f * d.Nmax + N] = new_ids[f];
    // physical observation slot: lowest free bit (k_cam_slot)
    const int slot = d.obs_phys[f];
    d.cam_phys[(size_t)f * d.Nmax + N] = slot;
    d.phys_used[f] |= 1ull << slot;
    for (int e = 0; e < 126; ++e) sJ[e] = 0.0;
    double St[9];
    skew3(tw, St);
    for (int i = 0; i < 3; ++i) {
      for (int j = 0; j < 3; ++j) { sJ[21 * i + j] = Ric[3 * i + j]; sJ[21 * (3 + i) + j] = St[3 * i + j]; }
      sJ[21 * i + 15 + i] = 1.0;
      sJ[21 * (3 + i) + 12 + i] = 1.0;
      sJ[21 * (3 + i) + 18 + i] = 1.0;
    }
  }
  __syncthreads();
  // new rows: J P[:21, :n]; mirrored into the new columns.  J (jit_utils.py:141-156) is [R_i_c 0 .. I_3 at 15] over
  // [skew(t) 0 .. I_3 at 12 .. I_3 at 18]: only the rows 0..2 and one or two more of P enter a new row (the zero
  // terms of the dense product change nothing, the others are added in the same ascending order)
  for (int e = tid; e < 6 * n; e += PT) {
    const int i = e / n, j = e - i * n;
    const double* Pj = P + j;
    double s = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) s += sJ[21 * i + k] * Pj[(size_t)k * ld];
    if (i < 3) {
      s += Pj[(size_t)(15 + i) * ld];
    } else {
      s += Pj[(size_t)(12 + i - 3) * ld];
      s += Pj[(size_t)(18 + i - 3) * ld];
    }
    if (j < 21) sB[21 * i + j] = s;
    P[(size_t)(n + i) * ld + j] = s;
    P[(size_t)j * ld + n + i] = s;
  }
  __syncthreads();
  if (tid < 36) {
    const int i = tid / 6, j = tid - 6 * i;
    double a = 0, b = 0;
    for (int k = 0; k < 21; ++k) { a += sB[21 * i + k] * sJ[21 * j + k]; b += sB[21 * j + k] * sJ[21 * i + k]; }
    P[(size_t)(n + i) * ld + n + j] = (a + b) / 2.;
  }
  if (tid == 0) d.ncam[f] = N + 1;
}

// ------------------------------------------------------------------ reset of P (msckf.py:820-830)
__device__ void reset_cov(const Dev& d, int f, int tid, int nthreads) {
  double* P = cov_of(d, f);
  const Params& prm = *d.prm;
  for (int e = tid; e < 21 * d.ld; e += nthreads) {
    const int i = e / d.ld, j = e - i * d.ld;
    double v = 0;
    if (i == j) {
      if (i >= 3 && i < 6) v = prm.P0[0];
      else if (i >= 6 && i < 9) v = prm.P0[1];
      else if (i >= 9 && i < 12) v = prm.P0[2];
      else if (i >= 15 && i < 18) v = prm.P0[3];
      else if (i >= 18 && i < 21) v = prm.P0[4];
    }
    P[e] = v;
  }
}

__device__ void clear_window_and_map(const Dev& d, int f, int tid, int nthreads) {
  for (int s = tid; s < d.Fcap; s += nthreads) {
    d.f_flags[(size_t)f * d.Fcap + s] = 0;
    d.f_mask[(size_t)f * d.Fcap + s] = 0ull;
    d.free_stack[(size_t)f * d.Fcap + s] = d.Fcap - 1 - s;   // pop order: slot 0 first
  }
  if (tid == 0) {
    d.nlive[f] = 0;
    d.nfree[f] = d.Fcap;
    d.ncam[f] = 0;
    d.phys_used[f] = 0ull;
  }
}

// online_reset (msckf.py:859-886).  Also snapshots the IMU block for publish() BEFORE resetting.
__global__ void __launch_bounds__(PT) k_finish_frame(Dev d, int do_reset_check) {
  const int f = blockIdx.x, tid = threadIdx.x;
  const double* imu = d.imu + (size_t)f * IMU_STRIDE;
  for (int e = tid; e < IMU_STRIDE; e += PT) d.imu_pub[(size_t)f * IMU_STRIDE + e] = imu[e];
  if (tid == 0) d.counters[NCNT * f + MSCKF_CNT_FRAMES] += 1;
  if (!do_reset_check) return;
  __shared__ int s_reset;
  if (tid == 0) {
    const double thr = d.prm->pos_std_thr;
    int r = 0;
    if (thr > 0) {
      const double* P = cov_of(d, f);
      const double sx = sqrt(P[(size_t)12 * d.ld + 12]), sy = sqrt(P[(size_t)13 * d.ld + 13]), sz = sqrt(P[(size_t)14 * d.ld + 14]);
      const double m = fmax(fmax(sx, sy), sz);
      r = !(m < thr);
    }
    s_reset = r;
    d.did_reset[f] = (unsigned char)r;
  }
  __syncthreads();
  if (s_reset) {
    clear_window_and_map(d, f, tid, PT);
    reset_cov(d, f, tid, PT);
  }
}

// full reset of one filter (msckf.py:832-857) or of all filters (f < 0) at creation
__global__ void __launch_bounds__(PT) k_reset(Dev d, int only_filter, int reset_imu) {
  const int f = only_filter >= 0 ? only_filter : blockIdx.x;
  const int tid = threadIdx.x;
  if (only_filter < 0) {                 // creation: every filter starts in its first covariance buffer
    if (tid == 0) {
      d.Pcur[f] = d.P + (size_t)f * d.ld * d.ld;
      d.p_compacted[f] = 0;
    }
    __syncthreads();
  }
  clear_window_and_map(d, f, tid, PT);
  reset_cov(d, f, tid, PT);
  if (reset_imu && tid == 0) {
    double* imu = d.imu + (size_t)f * IMU_STRIDE;
    double keepR[9], keept[3], keepg[3];
    for (int i = 0; i < 9; ++i) keepR[i] = imu[MSCKF_IMU_R_IC + i];
    for (int i = 0; i < 3; ++i) { keept[i] = imu[MSCKF_IMU_T_CI + i]; keepg[i] = imu[MSCKF_IMU_GRAVITY + i]; }
    for (int i = 0; i < IMU_STRIDE; ++i) imu[i] = 0.0;
    imu[MSCKF_IMU_Q + 3] = 1.0;
    imu[MSCKF_IMU_Q_NULL + 3] = 1.0;
    for (int i = 0; i < 9; ++i) imu[MSCKF_IMU_R_IC + i] = keepR[i];
    for (int i = 0; i < 3; ++i) { imu[MSCKF_IMU_T_CI + i] = keept[i]; imu[MSCKF_IMU_GRAVITY + i] = keepg[i]; }
  }
}

// ------------------------------------------------------------------ camera-state removal (msckf.py:803-818)
// Compaction of P, the camera table and the physical-slot mask.  P is copied without the rows / columns of the two camera
// states into the filter's other covariance buffer, which becomes the current one (no ordering hazards: every thread
// copies on its own).
__global__ void __launch_bounds__(PT) k_remove_cams(Dev d) {
  const int f = blockIdx.x, tid = threadIdx.x;
  const int r0 = d.rm_idx[2 * f], r1 = d.rm_idx[2 * f + 1];
  if (r0 < 0) return;
  const int N = d.ncam[f], n = NI + 6 * N, ld = d.ld;
  // the pruning update already wrote P without the two camera states (k_upd_dmma, stage 1): only the camera table is left
  const bool compacted = d.p_compacted[f] != 0;
  __syncthreads();
  if (!compacted) {
    const double* P = cov_of(d, f);
    double* Pout = cov_other(d, f);
    const int n2 = n - 12;
    auto src_of = [&](int i) {   // new index -> old index
      if (i < NI + 6 * r0) return i;
      if (i < NI + 6 * (r1 - 1)) return i + 6;
      return i + 12;
    };
    // a warp takes four rows at a time, lanes run along them: 16 loads in flight per thread before the first store, no
    // index arithmetic beyond the two jumps over the removed blocks
    const int lane = tid & 31, warp = tid >> 5;
    for (int i0 = 4 * warp; i0 < n2; i0 += 4 * (PT / 32)) {
      const double* src[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) src[u] = P + (size_t)src_of(min(i0 + u, n2 - 1)) * ld;
      for (int j0 = lane; j0 < n2; j0 += 128) {
        double v[4][4];
        int sj[4];
#pragma unroll
        for (int w = 0; w < 4; ++w) sj[w] = src_of(min(j0 + 32 * w, n2 - 1));
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int w = 0; w < 4; ++w) v[u][w] = src[u][sj[w]];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int w = 0; w < 4; ++w)
            if (i0 + u < n2 && j0 + 32 * w < n2) Pout[(size_t)(i0 + u) * ld + j0 + 32 * w] = v[u][w];
      }
    }
    if (tid == 0) d.Pcur[f] = Pout;
  } else if (tid == 0) {
    d.p_compacted[f] = 0;
  }
  // camera table: stage every surviving entry, then write
  __shared__ double scam[64 * CAM_STRIDE];
  __shared__ long long sid[64];
  __shared__ int sphys[64];
  const int N2 = N - 2;
  for (int e = tid; e < N2 * CAM_STRIDE; e += PT) {
    const int c = e / CAM_STRIDE, k = e - c * CAM_STRIDE;
    const int oc = c < r0 ? c : (c < r1 - 1 ? c + 1 : c + 2);
    scam[e] = d.cam[((size_t)f * d.Nmax + oc) * CAM_STRIDE + k];
    if (k == 0) { sid[c] = d.cam_id[(size_t)f * d.Nmax + oc]; sphys[c] = d.cam_phys[(size_t)f * d.Nmax + oc]; }
  }
  __syncthreads();
  for (int e = tid; e < N2 * CAM_STRIDE; e += PT) {
    const int c = e / CAM_STRIDE, k = e - c * CAM_STRIDE;
    d.cam[((size_t)f * d.Nmax + c) * CAM_STRIDE + k] = scam[e];
    if (k == 0) { d.cam_id[(size_t)f * d.Nmax + c] = sid[c]; d.cam_phys[(size_t)f * d.Nmax + c] = sphys[c]; }
  }
  if (tid == 0) {
    d.ncam[f] = N2;
    d.phys_used[f] &= ~d.rm_mask[f];
  }
}

// ------------------------------------------------------------------ launchers
void launch_propagate(const Dev& d, const int* imu_off, const double* ts, const double* gyro, const double* acc, cudaStream_t st, const FrameIn* fin) {
  static DevOnce once;     // function attributes are per device
  const size_t smem = (size_t)PROP_DOUBLES * sizeof(double);
  once.run([&] {
    cudaFuncSetAttribute(k_propagate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  });
  KLAUNCH(d, KID_PROPAGATE, st, (k_propagate<<<d.F, PW * 32, smem, st>>>(d, imu_off, ts, gyro, acc, fin)));
}
void launch_cam_slot(const Dev& d, cudaStream_t st) { KLAUNCH(d, KID_MISC, st, (k_cam_slot<<<(d.F + 255) / 256, 256, 0, st>>>(d))); }
void launch_augment(const Dev& d, const long long* ids, cudaStream_t st, const FrameIn* fin, bool slot_done) {
  if (!slot_done) launch_cam_slot(d, st);
  KLAUNCH(d, KID_AUGMENT, st, (k_augment<<<d.F, PT, 0, st>>>(d, ids, fin)));
}
void launch_finish_frame(const Dev& d, int do_reset_check, cudaStream_t st) {
  KLAUNCH(d, KID_FINISH_FRAME, st, (k_finish_frame<<<d.F, PT, 0, st>>>(d, do_reset_check)));
}
void launch_reset(const Dev& d, int only_filter, int reset_imu, cudaStream_t st) {
  KLAUNCH(d, KID_RESET, st, (k_reset<<<only_filter >= 0 ? 1 : d.F, PT, 0, st>>>(d, only_filter, reset_imu)));
}
void launch_remove_cams(const Dev& d, cudaStream_t st) {
  KLAUNCH(d, KID_REMOVE_CAMS, st, (k_remove_cams<<<d.F, PT, 0, st>>>(d)));
}

}  // namespace msckf
