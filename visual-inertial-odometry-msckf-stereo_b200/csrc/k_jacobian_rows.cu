// measurement_jacobian / feature_jacobian / gating_test (msckf.py:429-541, 606-614) for short tracks, one LANE per
// Jacobian row: a group of G = 8 / 16 lanes holds the 4M rows of a landmark with M = 2 / 3..4 observations in
// registers (H_x row, H_f row, residual).  The three Householder reflectors on H_f, the compact-WY projection
// H' = E - V Z, the gate products T1 = H' Pcc, S = T1 H'^T + sigma^2 I and its Cholesky factorisation all run on
// registers with group-wide shuffles; shared memory only carries what every row needs from the others
// (Z, Pcc, the projected rows - which are also the record the QR kernel reads).
// Same algebra as k_jacobian.cu (the CTA / warp-per-landmark kernels that take the longer tracks).
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int JR_T = 128;                 // threads per CTA
constexpr unsigned JFULL = 0xffffffffu;

template <int G>
__device__ __forceinline__ double jr_gsum(double v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(JFULL, v, o);
  return v;
}

template <int M>
__host__ __device__ constexpr int jr_group_doubles() {
  // Pcc [6M][6M] | Z [3][6M] | H' rows + residual [4M-3][6M+1] (padded to even) | cl [M] ints
  // even, so that every group's Pcc block is 16-byte aligned and its rows are read with 128-bit loads
  return (36 * M * M + 18 * M + (((4 * M - 3) * (6 * M + 1) + 1) & ~1) + ((M + 1) / 2) + 2 + 1) & ~1;
}

// One row of msckf.py:429-498: q = 0,1 -> cam0 u,v ; q = 2,3 -> cam1 u,v.
__device__ __forceinline__ void jr_row(const double* __restrict__ pre, const double* p_w, const double* __restrict__ z, const double* g,
                                       const double* R01, int q, double* Hx /*[6]*/, double& res) {
  const int h = q >> 1, i = q & 1;
  double R0[9];
#pragma unroll
  for (int e = 0; e < 9; ++e) R0[e] = pre[e];
  const double d0[3] = {p_w[0] - pre[18], p_w[1] - pre[19], p_w[2] - pre[20]};
  double pc0[3], p[3];
  mat3_vec(R0, d0, pc0);
  if (h) {
    const double d1[3] = {p_w[0] - pre[21], p_w[1] - pre[22], p_w[2] - pre[23]};
    mat3_vec(pre + 9, d1, p);
  } else {
    p[0] = pc0[0]; p[1] = pc0[1]; p[2] = pc0[2];
  }
  const double iz = 1.0 / p[2];
  const double pn = (i ? p[1] : p[0]);
  const double b2 = -(pn * iz) * iz;                                // one reciprocal of the depth serves the row and the residual
  const double b[3] = {i ? 0.0 : iz, i ? iz : 0.0, b2};            // row of dz/dp
  double c[3];
  if (h) {
    mat3t_vec(R01, b, c);                                           // b^T R_01
  } else {
    c[0] = b[0]; c[1] = b[1]; c[2] = b[2];
  }
  double A[6];
  cross3(c, pc0, A);                                                // c^T [pc0]x = (c x pc0)^T
  double cr[3];
  mat3t_vec(R0, c, cr);                                             // c^T R0
  A[3] = -cr[0]; A[4] = -cr[1]; A[5] = -cr[2];
  // observability constraint (msckf.py:483-490)
  double u[6];
  u[0] = pre[24]; u[1] = pre[25]; u[2] = pre[26];
  const double dn[3] = {p_w[0] - pre[27], p_w[1] - pre[28], p_w[2] - pre[29]};
  cross3(dn, g, u + 3);
  double uu = 0, au = 0;
#pragma unroll
  for (int e = 0; e < 6; ++e) { uu += u[e] * u[e]; au += A[e] * u[e]; }
  const double sc = au / uu;
#pragma unroll
  for (int e = 0; e < 6; ++e) Hx[e] = A[e] - sc * u[e];
  res = z[q] - pn * iz;
}


// registers: the projected row (6M), its product with Pcc (6M) and the row of S (4M-3) live in registers per lane.
// Register budget (launch bounds): 128 / 128 / 168 registers, no spills.  CTAs per SM actually launched: fewer than that
// budget would hold, see launch_rows.
template <int M>
constexpr int jr_min_blocks() { return M == 4 ? 3 : 4; }
template <int M>
constexpr int jr_ctas_per_sm() { return M == 2 ? 3 : (M == 3 ? 2 : 1); }

template <int M, int G>
__global__ void __launch_bounds__(JR_T, jr_min_blocks<M>()) k_jacobian_rows(Dev d, int cls, int write_all) {
  constexpr int R4 = 4 * M, C6 = 6 * M, R = R4 - 3, LDR = C6 + 1, GPW = 32 / G, GD = jr_group_doubles<M>();
  static_assert(R4 <= G, "one lane per row");
  extern __shared__ __align__(16) double jsm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int grp = lane / G, r = lane - grp * G, gbase = grp * G;
  const int n = (int)d.stage_ctr[SC_NJAC_SMALL + cls];
  if (n == 0) return;
  const int2* wl = d.wl_jac_small + (size_t)cls * d.F * d.Fcap;
  double* sP = jsm + (size_t)(warp * GPW + grp) * GD;     // Pcc [C6][C6]
  double* sZ = sP + C6 * C6;                               // Z [3][C6]
  double* sH = sZ + 3 * C6;                                // projected rows | residual [R][LDR]
  int* cl = (int*)(sH + ((R * LDR + 1) & ~1));             // logical camera of observation a
  const Params& prm = *d.prm;
  const bool rowv = r < R4;                                // lanes beyond the 4M rows are padding
  const int a = rowv ? (r >> 2) : 0, q = r & 3;
  const int ii = r - 3;                                    // index of the projected row (rows 0..2 are eliminated)
  // Every group takes a contiguous piece of the work list: k_plan_items queues the landmarks of a filter next to each
  // other, and in the pruning stage they all sit on the same two cameras, so the gathered Pcc block is reused from one
  // landmark to the next (P does not change while this kernel runs).
  const int ngroups = gridDim.x * (JR_T / 32) * GPW;
  const int per = (n + ngroups - 1) / ngroups;
  const int wbeg = ((blockIdx.x * (JR_T / 32) + warp) * GPW + grp) * per;
  int have_f = -1;
  unsigned long long have_lmask = 0;
  for (int wi = 0; wi < per; ++wi) {
    const bool active = wbeg + wi < n;
    const int2 it = wl[active ? wbeg + wi : n - 1];
    const int f = it.x, k = it.y;
    const size_t ik = (size_t)f * d.Fcap + k;
    const int N = d.ncam[f], ld = d.ld;
    const int* phys = d.cam_phys + (size_t)f * d.Nmax;
    // ---- observation mask in logical camera order (k_plan_items), cameras of the M observations
    const unsigned long long lmask = d.item_lmask[ik];
    if (r < M) {
      unsigned long long mm = lmask;
      for (int x = 0; x < r; ++x) mm &= mm - 1;
      cl[r] = __ffsll((long long)mm) - 1;
    }
    __syncwarp();
    // ---- step 1: this lane's row of H_x, H_f, r
    const int slot = d.item_slot[ik];
    const size_t fs = (size_t)f * d.Fcap + slot;
    double Hx[6], v[3], res;
    {
      const int c = cl[a];
      const double p_w[3] = {d.f_pos[fs * 3], d.f_pos[fs * 3 + 1], d.f_pos[fs * 3 + 2]};
      jr_row(d.cam_pre + ((size_t)f * d.Nmax + c) * CAM_PRE, p_w, d.f_obs + (fs * d.Nmax + phys[c]) * 4,
             d.imu + (size_t)f * IMU_STRIDE + MSCKF_IMU_GRAVITY, prm.R01, q, Hx, res);
      if (!rowv) {
#pragma unroll
        for (int e = 0; e < 6; ++e) Hx[e] = 0.0;
        res = 0.0;
      }
      v[0] = -Hx[3]; v[1] = -Hx[4]; v[2] = -Hx[5];
    }
    // ---- Pcc into shared memory (group-cooperative gather; P is exactly symmetric)
    if (f != have_f || lmask != have_lmask) {               // group-uniform
      have_f = f;
      have_lmask = lmask;
      const double* P = cov_of(d, f);
      for (int e = r; e < C6 * C6; e += G) {
        const int i = e / C6, c = e - i * C6;
        sP[e] = P[(size_t)(NI + 6 * cl[i / 6] + i % 6) * ld + NI + 6 * cl[c / 6] + c % 6];
      }
    }
    // ---- step 2: three Householder reflectors on H_f (rows across the lanes), compact-WY T, Q^T r
    double tau[3];
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      const double xn2 = jr_gsum<G>(r > kk ? v[kk] * v[kk] : 0.0);
      const double alpha = __shfl_sync(JFULL, v[kk], kk, G);
      // 1 / |beta| from one reciprocal square root (the FP64 sqrt and divide sequences are most of this kernel's instructions)
      const double n2 = alpha * alpha + xn2;
      const bool nz = xn2 != 0.0;
      const double rn = nz ? rsqrt(n2) : 0.0;
      const double nrm = n2 * rn;
      const double beta = (alpha >= 0) ? -nrm : nrm;
      tau[kk] = nz ? (beta - alpha) * ((alpha >= 0) ? -rn : rn) : 0.0;
      const double sc = nz ? 1.0 / (alpha - beta) : 1.0;
      v[kk] = (r > kk) ? v[kk] * sc : (r == kk ? 1.0 : 0.0);
#pragma unroll
      for (int c = kk + 1; c < 3; ++c) {
        const double dot = jr_gsum<G>(r >= kk ? v[kk] * v[c] : 0.0) * tau[kk];
        if (r >= kk) v[c] -= dot * v[kk];
      }
    }
    double T00, T01, T02, T11, T12, T22;
    {
      const double d01 = jr_gsum<G>(v[0] * v[1]), d02 = jr_gsum<G>(v[0] * v[2]), d12 = jr_gsum<G>(v[1] * v[2]);
      const double y0 = jr_gsum<G>(v[0] * res), y1 = jr_gsum<G>(v[1] * res), y2 = jr_gsum<G>(v[2] * res);
      T00 = tau[0]; T11 = tau[1]; T22 = tau[2];
      T01 = -tau[1] * (T00 * d01);
      T02 = -tau[2] * (T00 * d02 + T01 * d12);
      T12 = -tau[2] * (T11 * d12);
      const double z0 = T00 * y0, z1 = T01 * y0 + T11 * y1, z2 = T02 * y0 + T12 * y1 + T22 * y2;
      res -= v[0] * z0 + v[1] * z1 + v[2] * z2;
    }
    // ---- step 3: Z = T^T V^T E; the four rows of an observation reduce among themselves
    {
      double Y[3][6];
#pragma unroll
      for (int kk = 0; kk < 3; ++kk)
#pragma unroll
        for (int e = 0; e < 6; ++e) {
          double y = v[kk] * Hx[e];
          y += __shfl_xor_sync(JFULL, y, 1);
          y += __shfl_xor_sync(JFULL, y, 2);
          Y[kk][e] = y;
        }
      if (q == 0 && rowv) {
#pragma unroll
        for (int e = 0; e < 6; ++e) {
          sZ[6 * a + e] = T00 * Y[0][e];
          sZ[C6 + 6 * a + e] = T01 * Y[0][e] + T11 * Y[1][e];
          sZ[2 * C6 + 6 * a + e] = T02 * Y[0][e] + T12 * Y[1][e] + T22 * Y[2][e];
        }
      }
    }
    __syncwarp();
    // ---- projected row of this lane: H'(r, :) = E(r, :) - V(r, :) Z
    double hrow[C6];
#pragma unroll
    for (int c = 0; c < C6; ++c) {
      const double e_rc = (c / 6 == a) ? Hx[c % 6] : 0.0;
      hrow[c] = e_rc - (v[0] * sZ[c] + v[1] * sZ[C6 + c] + v[2] * sZ[2 * C6 + c]);
    }
    if (ii >= 0 && rowv) {
#pragma unroll
      for (int c = 0; c < C6; ++c) sH[ii * LDR + c] = hrow[c];
      sH[ii * LDR + C6] = res;
    }
    __syncwarp();
    // ---- gate: row ii of S = H' Pcc H'^T + sigma^2 I, Cholesky across the lanes with y = L^-1 r' riding along
    double srow[R];
    {
      double t[C6];
#pragma unroll
      for (int c = 0; c < C6; ++c) t[c] = 0.0;
#pragma unroll
      for (int e = 0; e < C6; ++e) {
        const double he = hrow[e];
#pragma unroll
        for (int c = 0; c < C6; ++c) t[c] += he * sP[e * C6 + c];
      }
#pragma unroll
      for (int j = 0; j < R; ++j) {
        double s = (j == ii) ? prm.obs_noise : 0.0;
#pragma unroll
        for (int c = 0; c < C6; ++c) s += t[c] * sH[j * LDR + c];
        srow[j] = s;
      }
    }
    bool spd = true;
    double gamma = 0.0;
#pragma unroll
    for (int j = 0; j < R; ++j) {
      const double djj = __shfl_sync(JFULL, srow[j], j + 3, G);
      if (!(djj > 0.0)) spd = false;                       // group-uniform
      const double il = rsqrt(djj);                        // only 1 / L(j,j) is needed
      const double lij = srow[j] * il;                     // meaningful on rows ii > j
      const double yj = __shfl_sync(JFULL, res * il, j + 3, G);
      gamma += yj * yj;
      if (ii > j) res -= lij * yj;
#pragma unroll
      for (int c = j + 1; c < R; ++c) {
        const double lcj = __shfl_sync(JFULL, lij, c + 3, G);
        if (ii >= c) srow[c] -= lij * lcj;
      }
    }
    if (!spd) gamma = INFINITY;
    const int pass = (gamma < prm.chi2[d.item_dof[ik]]) ? 1 : 0;
    if (active && r == 0) {
      // a landmark whose S is not positive definite to rounding (gross outlier) is counted and its gate redone with an
      // LU solve, as the reference computes it (k_gate_lu, k_jacobian.cu)
      if (!spd) {
        atomicAdd((unsigned long long*)(d.counters + (size_t)NCNT * f + MSCKF_CNT_GATE_NOT_SPD), 1ull);
        queue_gate_lu(d.wl_lu, d.stage_ctr, f, k);
      }
      d.item_gamma[ik] = gamma;
      d.item_pass[ik] = pass;
    }
    // ---- record for the stacking / QR kernel: the projected rows and residual, row-major [R][6M+1]
    if (active && (pass || write_all || !spd) && d.item_off[ik] + jac_record_doubles(M) <= d.pool_cap) {
      double* out = d.pool + d.item_off[ik];
      for (int e = r; e < R * LDR; e += G) out[e] = sH[e];
    }
    __syncwarp();
  }
}

template <int M, int G>
static void launch_rows(const Dev& d, int cls, int write_all, cudaStream_t st) {
  static DevOnce once;     // function attributes are per device
  const size_t smem = (size_t)(JR_T / G) * jr_group_doubles<M>() * sizeof(double);
  once.run([&] {
    cudaFuncSetAttribute(k_jacobian_rows<M, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  });
  const int sms = sm_count();
  // One wave of resident CTAs, every group walking a contiguous piece of the list -- but well below what the register file
  // holds (3 / 2 / 1 CTAs per SM, it would hold 4 / 4 / 3): a full wave owns every register of every SM for the whole launch
  // (50-100 us) and whatever else is ready (the other classes on their lanes, the other half-batch) has to wait; leaving
  // room is worth 2-3 % of a frame although the kernel alone gets 30 % slower.
  int per_sm = (int)std::min<size_t>(jr_ctas_per_sm<M>(), (220 * 1024) / (smem + 1024));
  per_sm = std::max(per_sm, 1);
  KLAUNCH(d, KID_JACOBIAN_ROWS, st, (k_jacobian_rows<M, G><<<sms * per_sm, JR_T, smem, st>>>(d, cls, write_all)));
}

void launch_jacobian_rows(const Dev& d, int cls, int write_all, cudaStream_t st) {
  if (cls == 0) launch_rows<2, 8>(d, cls, write_all, st);
  else if (cls == 1) launch_rows<3, 16>(d, cls, write_all, st);
  else launch_rows<4, 16>(d, cls, write_all, st);
}

}  // namespace msckf
