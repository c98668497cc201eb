// measurement_update (msckf.py:555-604) for updates that do not fit one CTA's shared memory: wide camera windows
// (up to 64 camera states, n = 405) and caller-supplied dense H.  Every contraction runs on the FP64 tensor pipe
// (mma.sync.m8n8k4.f64 -> DMMA) and every operand row is staged into shared memory by the TMA engine (cp.async.bulk
// tracked by mbarriers -> UBLKCP), three stages deep; the filters of a batch and the tiles of a filter spread over the
// whole grid.  Algebra (k_update.cu header), with the upper Cholesky factor S = U^T U:
//     k_wupd_T         T = H_thin P[act,:] | r          (k x (n+1));  also T[:,act] transposed, the operand of S
//     k_wupd_S         S = T[:,act] H_thin^T + sigma^2 I (upper 64 x 64 tiles)
//     k_wupd_chol      S = U^T U in place, right looking with 32-row panels, plus the inverses of the diagonal blocks
//     k_wupd_trsm      W = U^-T [T | r]                  (32-column strips resident in shared memory)
//     k_upd_inject     dx = W^T y, state injection       (msckf.py:565-595)
//     k_wupd_downdate  P <- P - W^T W                    (upper 64 x 64 tiles, mirrored store: exactly symmetric)
// All GEMMs have the form C = A^T B with both operands "k-major" in memory (row kk of A holds A[kk][0..63] contiguously),
// which is how R (column-major), the rows of P, the rows of W and the rows of U lie in HBM: one bulk copy per operand row.
// Rows k .. k8-1 (k rounded up to 8) are padded with identity (S) / zeros (T) so that no kernel has ragged 8-row tiles.
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"
#include "tma.cuh"

namespace msckf {

constexpr int WT = 256;                  // threads per CTA: 8 warps as 2 (rows) x 4 (columns), warp tile 32 x 16
constexpr int WKC = 32;                  // operand rows (k) per pipeline stage
constexpr int WP = 68;                   // shared-memory pitch of a staged 64-double row: 4 (mod 16), conflict-free fragments
constexpr int WNS = 3;                   // pipeline stages
constexpr int WSTAGE = 2 * WKC * WP;     // doubles per stage: A rows, then B rows
constexpr size_t WGEMM_SMEM = (size_t)WNS * WSTAGE * sizeof(double);

struct Ring {
  double* buf;          // [WNS][2][WKC][WP]
  uint64_t* full;       // [WNS] "stage filled" barriers (one arrival: the producer lane that announces the byte count)
  unsigned it;          // chunks consumed so far (stage = it % WNS, parity = (it / WNS) & 1)
};

__device__ __forceinline__ void ring_init(Ring& rg, double* buf, uint64_t* full) {
  rg.buf = buf;
  rg.full = full;
  rg.it = 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < WNS; ++s) mbar_init(&full[s], 1);
    fence_mbar_init();
  }
  __syncthreads();
}

// acc (64 x 64 CTA tile, 32 x 16 per warp) += sum_{kk in [k0, k1)} A[kk][0..63]^T B[kk][0..63].
// rowA(kk) / rowB(kk): global address of the operand row; dblA / dblB: doubles to copy per row (even, <= 64).
template <class FA, class FB>
__device__ __forceinline__ void wide_gemm(Ring& rg, int k0, int k1, FA rowA, int dblA, FB rowB, int dblB, double (&acc)[4][2][2]) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3;
  const int wm = warp >> 2, wn = warp & 3;
  const int nch = (k1 - k0 + WKC - 1) / WKC;
  auto issue = [&](int c) {                                 // warp 0: lane l moves operand row l of chunk c
    const unsigned s = (rg.it + c) % WNS;
    double* a = rg.buf + (size_t)s * WSTAGE + lane * WP;
    double* b = a + WKC * WP;
    const int valid = min(WKC, k1 - k0 - c * WKC);
    if (lane >= valid && lane < ((valid + 3) & ~3)) {       // ragged tail: the rest of the last 4-deep step reads zeros
      for (int e = 0; e < 64; ++e) { a[e] = 0.0; b[e] = 0.0; }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive_expect_tx(&rg.full[s], (unsigned)(valid * (dblA + dblB) * sizeof(double)));
    __syncwarp();
    if (lane < valid) {
      const int kk = k0 + c * WKC + lane;
      bulk_g2s(a, rowA(kk), dblA * sizeof(double), &rg.full[s]);
      bulk_g2s(b, rowB(kk), dblB * sizeof(double), &rg.full[s]);
    }
  };
  __syncthreads();                                          // earlier readers of the ring are done
  if (warp == 0)
    for (int c = 0; c < min(nch, WNS - 1); ++c) issue(c);
  for (int c = 0; c < nch; ++c) {
    if (c > 0) __syncthreads();                             // chunk c - 1 is consumed: its stage may be refilled
    if (warp == 0 && c + WNS - 1 < nch) issue(c + WNS - 1);
    const unsigned s = (rg.it + c) % WNS, par = ((rg.it + c) / WNS) & 1u;
    mbar_wait(&rg.full[s], par);
    const double* a = rg.buf + (size_t)s * WSTAGE + tg * WP + wm * 32 + g;
    const double* b = rg.buf + (size_t)s * WSTAGE + WKC * WP + tg * WP + wn * 16 + g;
    const int nks = (min(WKC, k1 - k0 - c * WKC) + 3) >> 2;
#pragma unroll 2
    for (int ks = 0; ks < nks; ++ks) {
      const double a0 = a[0], a1 = a[8], a2 = a[16], a3 = a[24], b0 = b[0], b1 = b[8];
      dmma(acc[0][0][0], acc[0][0][1], a0, b0);
      dmma(acc[0][1][0], acc[0][1][1], a0, b1);
      dmma(acc[1][0][0], acc[1][0][1], a1, b0);
      dmma(acc[1][1][0], acc[1][1][1], a1, b1);
      dmma(acc[2][0][0], acc[2][0][1], a2, b0);
      dmma(acc[2][1][0], acc[2][1][1], a2, b1);
      dmma(acc[3][0][0], acc[3][0][1], a3, b0);
      dmma(acc[3][1][0], acc[3][1][1], a3, b1);
      a += 4 * WP;
      b += 4 * WP;
    }
  }
  rg.it += nch;
}

__device__ __forceinline__ void zero_acc(double (&acc)[4][2][2]) {
#pragma unroll
  for (int x = 0; x < 4; ++x)
#pragma unroll
    for (int y = 0; y < 2; ++y) acc[x][y][0] = acc[x][y][1] = 0.0;
}

// ------------------------------------------------------------------ T = H_thin P[act,:] | r,  and T[:,act]^T
// grid (column tiles of n + 1, row tiles of k, filters)
__global__ void __launch_bounds__(WT, 2) k_wupd_T(Dev d) {
  extern __shared__ __align__(128) double wsm[];
  __shared__ uint64_t s_full[WNS];
  __shared__ int s_inv[64];
  __shared__ int s_cm[NI + 6 * 64 + 3];
  Ring rg;
  ring_init(rg, wsm, s_full);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3, wm = warp >> 2, wn = warp & 3;
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  for (int w = blockIdx.z; w < nbig; w += gridDim.z) {
    const int f = d.wl_upd_big[w];
    const int k = d.k_rows[f], k8 = (k + 7) & ~7;
    const int n = NI + 6 * d.ncam[f];
    if (k == 0 || i0 >= k8 || j0 > n) continue;
    const int na = d.n_act[f], mp = d.m_pad[f], ld = d.ld, tld = Tld(d), sld = d.Sld;
    const bool tri = d.m_rows[f] > na;
    const double* A = d.A + (size_t)f * d.A_elems;
    const double* P = cov_of(d, f);
    const int* cm = d.colmap + (size_t)f * d.nmax;
    __syncthreads();                                        // the previous tile's epilogue has read s_inv / s_cm
    if (tid < 64) s_inv[tid] = -1;
    for (int c = tid; c < na; c += WT) s_cm[c] = cm[c];
    __syncthreads();
    for (int c = tid; c < na; c += WT) {
      const int j = s_cm[c] - j0;
      if (j >= 0 && j < 64) s_inv[j] = c;
    }
    double acc[4][2][2];
    zero_acc(acc);
    wide_gemm(rg, tri ? i0 : 0, na,
              [&](int c) { return A + (size_t)c * mp + i0; }, min(64, k8 - i0),
              [&](int c) { return P + (size_t)s_cm[c] * ld + j0; }, min(64, ld - j0), acc);
    double* T = d.T + (size_t)f * d.Tk * tld;
    double* Tt = d.TactT + (size_t)f * d.Tk * sld;
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      const int i = i0 + wm * 32 + mt * 8 + g;
      if (i >= k8) continue;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = j0 + wn * 16 + nt * 8 + 2 * tg + e;
          const double v = (i < k) ? acc[mt][nt][e] : 0.0;
          if (j < n) {
            T[(size_t)i * tld + j] = v;
            const int c = s_inv[j - j0];
            if (c >= 0) Tt[(size_t)c * sld + i] = v;
          } else if (j == n) {
            T[(size_t)i * tld + n] = (i < k) ? A[(size_t)na * mp + i] : 0.0;
          }
        }
    }
  }
}

// ------------------------------------------------------------------ S = T[:,act] H_thin^T + sigma^2 I  (upper tiles)
// grid (column tiles of k, row tiles of k, filters)
__global__ void __launch_bounds__(WT, 2) k_wupd_S(Dev d) {
  extern __shared__ __align__(128) double wsm[];
  __shared__ uint64_t s_full[WNS];
  Ring rg;
  ring_init(rg, wsm, s_full);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3, wm = warp >> 2, wn = warp & 3;
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  if (j0 < i0) return;
  const double sigma2 = d.prm->obs_noise;
  for (int w = blockIdx.z; w < nbig; w += gridDim.z) {
    const int f = d.wl_upd_big[w];
    const int k = d.k_rows[f], k8 = (k + 7) & ~7;
    if (k == 0 || j0 >= k8) continue;
    const int na = d.n_act[f], mp = d.m_pad[f], sld = d.Sld;
    const bool tri = d.m_rows[f] > na;
    const double* A = d.A + (size_t)f * d.A_elems;
    const double* Tt = d.TactT + (size_t)f * d.Tk * sld;
    double acc[4][2][2];
    zero_acc(acc);
    wide_gemm(rg, tri ? j0 : 0, na,
              [&](int c) { return Tt + (size_t)c * sld + i0; }, min(64, k8 - i0),
              [&](int c) { return A + (size_t)c * mp + j0; }, min(64, k8 - j0), acc);
    double* S = d.S + (size_t)f * d.Tk * sld;
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      const int i = i0 + wm * 32 + mt * 8 + g;
      if (i >= k8) continue;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int j = j0 + wn * 16 + nt * 8 + 2 * tg + e;
          if (j >= k8) continue;
          S[(size_t)i * sld + j] = (i < k && j < k) ? acc[mt][nt][e] + (i == j ? sigma2 : 0.0) : (i == j ? 1.0 : 0.0);
        }
    }
  }
}

// ------------------------------------------------------------------ S = U^T U  (upper factor, in place) + diagonal-block inverses
// One CTA per filter, right looking with 32-row panels.  Per panel: warp 0 factors the 32 x 32 diagonal block in
// registers (lane = column, entries travel by shuffle) and inverts the factor; the panel row S(p, p+1:) arrives by bulk
// copy meanwhile; U(p, p+1:) = U_pp^-T S(p, p+1:) and the trailing update S(I, J) -= U(p, I)^T U(p, J) are DMMA tiles fed
// from the shared-memory panel, the trailing matrix is read and written in L2.
constexpr int CB = 32;                   // panel height
constexpr int CDP = 36;                  // pitch of the 32 x 32 blocks: 4 (mod 16)

__host__ __device__ inline int chol_pan_pitch(int Tk) { return ((Tk + 15) & ~15) + 4; }
static size_t chol_smem_bytes(int Tk) { return ((size_t)CB * chol_pan_pitch(Tk) + 2 * CB * CDP + CB) * sizeof(double); }

// warp-level Cholesky D = U^T U of a 32 x 32 block (lane c holds column c), then the inverse of U; results to sD / sUi.
__device__ __forceinline__ bool chol32_inverse(double* sD, double* sUi, double* sRs, int lane) {
  double col[CB];
#pragma unroll
  for (int r = 0; r < CB; ++r) col[r] = sD[r * CDP + lane];
  bool bad = false;
#pragma unroll
  for (int j = 0; j < CB; ++j) {
    const double djj = __shfl_sync(0xffffffffu, col[j], j);
    if (!(djj > 0.0)) bad = true;                           // warp-uniform
    const double rs = rsqrt(djj);
    if (lane == j) sRs[j] = rs;                             // 1 / U(j,j)
    const double ujc = (lane == j) ? djj * rs : col[j] * rs;
    col[j] = ujc;
#pragma unroll
    for (int r = j + 1; r < CB; ++r) {
      const double ujr = __shfl_sync(0xffffffffu, ujc, r);
      if (lane >= r) col[r] -= ujr * ujc;
    }
  }
  __syncwarp();
#pragma unroll
  for (int r = 0; r < CB; ++r) sD[r * CDP + lane] = (r <= lane) ? col[r] : 0.0;
  __syncwarp();
  // column `lane` of U^-1 by back substitution (x_r = 0 below the diagonal)
#pragma unroll
  for (int r = CB - 1; r >= 0; --r) {
    double s = (r == lane) ? 1.0 : 0.0;
#pragma unroll
    for (int q = r + 1; q < CB; ++q) s -= sD[r * CDP + q] * col[q];
    col[r] = (r <= lane) ? s * sRs[r] : 0.0;
  }
#pragma unroll
  for (int r = 0; r < CB; ++r) sUi[r * CDP + lane] = col[r];
  return bad;
}

__global__ void __launch_bounds__(WT, 1) k_wupd_chol(Dev d) {
  extern __shared__ __align__(128) double wsm[];
  __shared__ uint64_t s_bar;
  __shared__ int s_bad;
  const int PP = chol_pan_pitch(d.Tk);
  double* sPan = wsm;                      // [32][PP]   S(p, p0+32 ..) -> U(p, p0+32 ..)
  double* sD = sPan + CB * PP;             // [32][CDP]  diagonal block -> U_pp
  double* sUi = sD + CB * CDP;             // [32][CDP]  U_pp^-1
  double* sRs = sUi + CB * CDP;            // [32]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3;
  if (tid == 0) { mbar_init(&s_bar, 1); fence_mbar_init(); }
  __syncthreads();
  unsigned par = 0;
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int sld = d.Sld;
  for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
    const int f = d.wl_upd_big[w];
    const int k = d.k_rows[f], k8 = (k + 7) & ~7;
    if (k == 0) continue;
    double* S = d.S + (size_t)f * d.Tk * sld;
    double* Ui = d.Uinv + (size_t)f * d.Uinv_stride;
    if (tid == 0) s_bad = 0;
    const int nb = (k8 + CB - 1) / CB;
    for (int p = 0; p < nb; ++p) {
      const int p0 = CB * p, bs = min(CB, k8 - p0), c1 = p0 + CB, rest = max(0, k8 - c1);
      __syncthreads();                                      // the previous trailing update is complete (and fenced)
      // panel row by bulk copy (warp 1), diagonal block by plain loads (everyone)
      if (rest > 0 && warp == 1) {
        if (lane == 0) mbar_arrive_expect_tx(&s_bar, (unsigned)(CB * rest * sizeof(double)));
        __syncwarp();
        bulk_g2s(sPan + lane * PP, S + (size_t)(p0 + lane) * sld + c1, rest * sizeof(double), &s_bar);
      }
      for (int e = tid; e < CB * CB; e += WT) {
        const int r = e >> 5, c = e & 31;
        sD[r * CDP + c] = (r < bs && c < bs) ? ((c >= r) ? S[(size_t)(p0 + r) * sld + p0 + c] : 0.0) : (r == c ? 1.0 : 0.0);
      }
      __syncthreads();
      if (warp == 0) {
        if (chol32_inverse(sD, sUi, sRs, lane) && lane == 0) s_bad = 1;
      }
      __syncthreads();
      if (s_bad) {
        if (rest > 0) { mbar_wait(&s_bar, par); par ^= 1u; }   // nothing stays in flight
        break;
      }
      for (int e = tid; e < CB * CB; e += WT) {             // U_pp and its inverse to HBM
        const int r = e >> 5, c = e & 31;
        if (r < bs && c < bs && c >= r) S[(size_t)(p0 + r) * sld + p0 + c] = sD[r * CDP + c];
        Ui[(size_t)p * CB * CB + e] = sUi[r * CDP + c];
      }
      if (rest == 0) continue;
      mbar_wait(&s_bar, par);
      par ^= 1u;
      // U(p, J) = U_pp^-T S(p, J): a warp owns whole 8-column tiles (reads its B fragments before it writes)
      for (int nt = warp; nt < rest / 8; nt += WT / 32) {
        double bf[8];
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) bf[ks] = sPan[(4 * ks + tg) * PP + 8 * nt + g];
        double c[4][2];
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          c[mt][0] = c[mt][1] = 0.0;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            if (ks <= 2 * mt + 1) dmma(c[mt][0], c[mt][1], sUi[(4 * ks + tg) * CDP + 8 * mt + g], bf[ks]);
        }
        __syncwarp();
#pragma unroll
        for (int mt = 0; mt < 4; ++mt) {
          *(double2*)(sPan + (8 * mt + g) * PP + 8 * nt + 2 * tg) = make_double2(c[mt][0], c[mt][1]);
          *(double2*)(S + (size_t)(p0 + 8 * mt + g) * sld + c1 + 8 * nt + 2 * tg) = make_double2(c[mt][0], c[mt][1]);
        }
      }
      __syncthreads();
      // trailing update, 32 x 32 tiles (I <= J) of the remaining rows / columns, one per warp
      const int nbr = (rest + CB - 1) / CB, ntile = nbr * (nbr + 1) / 2;
      for (int t = warp; t < ntile; t += WT / 32) {
        int J = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
        while ((J + 1) * (J + 2) / 2 <= t) ++J;
        while (J * (J + 1) / 2 > t) --J;
        const int I = t - J * (J + 1) / 2;                  // I <= J
        const int a0 = CB * I, b0 = CB * J;                 // column offsets inside the panel
        const int mtn = min(4, (rest - a0) / 8), ntn = min(4, (rest - b0) / 8);
        double c[4][4][2];
#pragma unroll
        for (int mt = 0; mt < 4; ++mt)
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) {
            c[mt][nt][0] = c[mt][nt][1] = 0.0;
            if (mt < mtn && nt < ntn && !(I == J && mt > nt)) {
              const double2 v = *(const double2*)(S + (size_t)(c1 + a0 + 8 * mt + g) * sld + c1 + b0 + 8 * nt + 2 * tg);
              c[mt][nt][0] = v.x;
              c[mt][nt][1] = v.y;
            }
          }
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          double af[4], bf[4];
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            af[x] = (x < mtn) ? -sPan[(4 * ks + tg) * PP + a0 + 8 * x + g] : 0.0;
            bf[x] = (x < ntn) ? sPan[(4 * ks + tg) * PP + b0 + 8 * x + g] : 0.0;
          }
#pragma unroll
          for (int mt = 0; mt < 4; ++mt)
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
              if (mt < mtn && nt < ntn && !(I == J && mt > nt)) dmma(c[mt][nt][0], c[mt][nt][1], af[mt], bf[nt]);
        }
#pragma unroll
        for (int mt = 0; mt < 4; ++mt)
#pragma unroll
          for (int nt = 0; nt < 4; ++nt)
            if (mt < mtn && nt < ntn && !(I == J && mt > nt))
              *(double2*)(S + (size_t)(c1 + a0 + 8 * mt + g) * sld + c1 + b0 + 8 * nt + 2 * tg) = make_double2(c[mt][nt][0], c[mt][nt][1]);
      }
      fence_proxy_async();                                  // the next panel row is read back by the TMA engine
    }
    __syncthreads();
    if (s_bad) {
      if (tid == 0) {
        raise_not_spd(d, 3, f, k, d.n_act[f], d.m_rows[f]);
        d.k_rows[f] = 0;                                    // later stages skip this filter
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ W = U^-T [T | r]
// One CTA per (filter, 32-column strip).  The strip X (k x 32) stays in shared memory; the rows of U stream through the
// ring in [32][<=64] chunks: for row block q first the inverse of its diagonal block (X_q <- U_qq^-T X_q), then the
// column chunks of U(q, q+1:) (X_I -= U(q, I)^T X_q).
constexpr int XP = 36;                   // pitch of the strip: 4 (mod 16)
static size_t trsm_smem_bytes(int Tk) { return ((size_t)((Tk + 31) & ~31) * XP + (size_t)WNS * WKC * WP) * sizeof(double); }

__global__ void __launch_bounds__(WT, 1) k_wupd_trsm(Dev d) {
  extern __shared__ __align__(128) double wsm[];
  __shared__ uint64_t s_full[WNS], s_ld;
  const int krows = (d.Tk + 31) & ~31;
  double* sX = wsm;                        // [krows][XP]
  double* ring = sX + (size_t)krows * XP;  // [WNS][WKC][WP]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3;
  if (tid == 0) {
    for (int s = 0; s < WNS; ++s) mbar_init(&s_full[s], 1);
    mbar_init(&s_ld, 1);
    fence_mbar_init();
  }
  __syncthreads();
  unsigned it = 0, ldpar = 0;
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int sld = d.Sld, tld = Tld(d);
  const int j0 = blockIdx.x * 32;
  for (int w = blockIdx.y; w < nbig; w += gridDim.y) {
    const int f = d.wl_upd_big[w];
    const int k = d.k_rows[f], k8 = (k + 7) & ~7;
    const int n = NI + 6 * d.ncam[f];
    if (k == 0 || j0 > n) continue;
    const double* S = d.S + (size_t)f * d.Tk * sld;
    const double* Ui = d.Uinv + (size_t)f * d.Uinv_stride;
    double* T = d.T + (size_t)f * d.Tk * tld;
    const int nb = (k8 + CB - 1) / CB, kp = nb * CB;
    const int wx = min(32, tld - j0);                       // strip columns that exist in T
    __syncthreads();                                        // the previous strip is stored
    // ---- strip in: rows < k8 by bulk copy, zero rows up to the block boundary
    if (tid == 0) mbar_arrive_expect_tx(&s_ld, (unsigned)(k8 * wx * sizeof(double)));
    for (int e = tid; e < (kp - k8) * XP; e += WT) sX[(size_t)k8 * XP + e] = 0.0;
    __syncthreads();
    for (int i = tid; i < k8; i += WT) bulk_g2s(sX + (size_t)i * XP, T + (size_t)i * tld + j0, wx * sizeof(double), &s_ld);
    // ---- chunk schedule: for q: [inverse block q], [U(q, c1 + 64 c ..) for c = 0 ..]
    // chunk index -> (q, kind) is recomputed by walking; producer (warp 0) and consumers walk the same sequence
    int total = 0;
    for (int q = 0; q < nb; ++q) total += 1 + (max(0, k8 - CB * (q + 1)) + 63) / 64;
    auto chunk_of = [&](int ci, int& q, int& cc) {          // cc = -1: inverse block, else column chunk
      int q_ = 0, c_ = ci;
      while (true) {
        const int cnt = 1 + (max(0, k8 - CB * (q_ + 1)) + 63) / 64;
        if (c_ < cnt) break;
        c_ -= cnt;
        ++q_;
      }
      q = q_;
      cc = c_ - 1;
    };
    auto issue = [&](int ci) {                              // warp 0
      int q, cc;
      chunk_of(ci, q, cc);
      const unsigned s = (it + ci) % WNS;
      double* dst = ring + (size_t)s * WKC * WP + lane * WP;
      if (cc < 0) {
        if (lane == 0) mbar_arrive_expect_tx(&s_full[s], (unsigned)(CB * CB * sizeof(double)));
        __syncwarp();
        bulk_g2s(dst, Ui + (size_t)q * CB * CB + lane * CB, CB * sizeof(double), &s_full[s]);
      } else {
        const int col0 = CB * (q + 1) + 64 * cc, len = min(64, k8 - col0);
        const int rows = min(CB, k8 - CB * q);              // == 32: a block with columns to its right is complete
        if (lane == 0) mbar_arrive_expect_tx(&s_full[s], (unsigned)(rows * len * sizeof(double)));
        __syncwarp();
        if (lane < rows) bulk_g2s(dst, S + (size_t)(CB * q + lane) * sld + col0, len * sizeof(double), &s_full[s]);
      }
    };
    if (warp == 0)
      for (int ci = 0; ci < min(total, WNS - 1); ++ci) issue(ci);
    mbar_wait(&s_ld, ldpar);
    ldpar ^= 1u;
    for (int ci = 0; ci < total; ++ci) {
      __syncthreads();                                      // chunk ci - 1 consumed (stage free), its writes to X visible
      if (warp == 0 && ci + WNS - 1 < total) issue(ci + WNS - 1);
      const unsigned s = (it + ci) % WNS, par = ((it + ci) / WNS) & 1u;
      mbar_wait(&s_full[s], par);
      const double* ch = ring + (size_t)s * WKC * WP;
      int q, cc;
      chunk_of(ci, q, cc);
      const int q0 = CB * q;
      if (cc < 0) {
        // X_q <- U_qq^-T X_q : 4 x 4 tiles of 8 x 8, two per warp; results held until every read is done
        const int mt = warp >> 1, ntb = (warp & 1) * 2;
        double c[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          if (ks > 2 * mt + 1) break;                       // the inverse is upper triangular
          const double af = ch[(4 * ks + tg) * WP + 8 * mt + g];
#pragma unroll
          for (int x = 0; x < 2; ++x) dmma(c[x][0], c[x][1], af, sX[(size_t)(q0 + 4 * ks + tg) * XP + 8 * (ntb + x) + g]);
        }
        __syncthreads();
#pragma unroll
        for (int x = 0; x < 2; ++x) *(double2*)(sX + (size_t)(q0 + 8 * mt + g) * XP + 8 * (ntb + x) + 2 * tg) = make_double2(c[x][0], c[x][1]);
      } else {
        // X(col0 .. col0+len, :) -= U(q, col0 ..)^T X_q : warp = 8-row tile of the chunk, 4 column tiles
        const int col0 = q0 + CB + 64 * cc, len = min(64, k8 - col0);
        if (8 * warp < len) {
          double2* cp[4];
          double2 c[4];
#pragma unroll
          for (int x = 0; x < 4; ++x) {
            cp[x] = (double2*)(sX + (size_t)(col0 + 8 * warp + g) * XP + 8 * x + 2 * tg);
            c[x] = *cp[x];
          }
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const double af = -ch[(4 * ks + tg) * WP + 8 * warp + g];
#pragma unroll
            for (int x = 0; x < 4; ++x) dmma(c[x].x, c[x].y, af, sX[(size_t)(q0 + 4 * ks + tg) * XP + 8 * x + g]);
          }
#pragma unroll
          for (int x = 0; x < 4; ++x) *cp[x] = c[x];
        }
      }
    }
    it += total;
    __syncthreads();
    // ---- strip out: W rows < k8, columns <= n
    for (int e = tid; e < k8 * 16; e += WT) {
      const int i = e >> 4, c2 = (e & 15) * 2, j = j0 + c2;
      if (j + 1 <= n) *(double2*)(T + (size_t)i * tld + j) = *(const double2*)(sX + (size_t)i * XP + c2);
      else if (j <= n) T[(size_t)i * tld + j] = sX[(size_t)i * XP + c2];
    }
  }
}

// ------------------------------------------------------------------ dx = W^T y, state injection (msckf.py:565-595)
__global__ void __launch_bounds__(WT) k_upd_inject(Dev d) {
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int tid = threadIdx.x;
  for (int w = blockIdx.x; w < nbig; w += gridDim.x) {
    const int f = d.wl_upd_big[w], k = d.k_rows[f];
    if (k == 0) continue;
    const int n = NI + 6 * d.ncam[f], tld = Tld(d);
    const double* T = d.T + (size_t)f * d.Tk * tld;
    double* dx = d.dx + (size_t)f * d.nmax;
    __shared__ double s_y[NI + 6 * 64 + 8];                 // y = L^-1 r_thin, read by every column
    __syncthreads();
    for (int i = tid; i < k; i += WT) s_y[i] = T[(size_t)i * tld + n];
    __syncthreads();
    for (int j = tid; j < n; j += WT) {                      // four independent chains: the loads of a column overlap
      double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
      int i = 0;
      for (; i + 4 <= k; i += 4) {
        s0 += T[(size_t)i * tld + j] * s_y[i];
        s1 += T[(size_t)(i + 1) * tld + j] * s_y[i + 1];
        s2 += T[(size_t)(i + 2) * tld + j] * s_y[i + 2];
        s3 += T[(size_t)(i + 3) * tld + j] * s_y[i + 3];
      }
      for (; i < k; ++i) s0 += T[(size_t)i * tld + j] * s_y[i];
      dx[j] = (s0 + s1) + (s2 + s3);
    }
    __syncthreads();
    inject_state(d, f, dx, tid, WT);
    __syncthreads();
  }
}

// ------------------------------------------------------------------ P <- P - W^T W  (upper 64 x 64 tiles, mirrored store)
// grid (column tiles of n, row tiles of n, filters)
__global__ void __launch_bounds__(WT, 2) k_wupd_downdate(Dev d) {
  extern __shared__ __align__(128) double wsm[];
  __shared__ uint64_t s_full[WNS];
  Ring rg;
  ring_init(rg, wsm, s_full);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3, wm = warp >> 2, wn = warp & 3;
  const int nbig = (int)d.stage_ctr[SC_NUPD_BIG];
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  if (j0 < i0) return;
  for (int w = blockIdx.z; w < nbig; w += gridDim.z) {
    const int f = d.wl_upd_big[w];
    const int k = d.k_rows[f], k8 = (k + 7) & ~7;
    const int n = NI + 6 * d.ncam[f];
    if (k == 0 || j0 >= n) continue;
    const int ld = d.ld, tld = Tld(d);
    const double* W = d.T + (size_t)f * d.Tk * tld;
    double acc[4][2][2];
    zero_acc(acc);
    wide_gemm(rg, 0, k8,
              [&](int r) { return W + (size_t)r * tld + i0; }, min(64, tld - i0),
              [&](int r) { return W + (size_t)r * tld + j0; }, min(64, tld - j0), acc);
    double* P = cov_of(d, f);
#pragma unroll
    for (int mt = 0; mt < 4; ++mt) {
      const int a = i0 + wm * 32 + mt * 8 + g;
      if (a >= n) continue;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int b = j0 + wn * 16 + nt * 8 + 2 * tg + e;
          if (b < n && b >= a) {
            const double v = P[(size_t)a * ld + b] - acc[mt][nt][e];
            P[(size_t)a * ld + b] = v;
            P[(size_t)b * ld + a] = v;
          }
        }
    }
  }
}

void launch_update_tiled(const Dev& d, cudaStream_t st) {
  // work-list driven (wl_upd_big): no wide update queued -> every grid exits at once
  static DevOnce once;     // function attributes are per device
  const size_t sm_chol = chol_smem_bytes(d.Tk), sm_trsm = trsm_smem_bytes(d.Tk);
  once.run([&] {
    cudaFuncSetAttribute(k_wupd_T, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WGEMM_SMEM);
    cudaFuncSetAttribute(k_wupd_S, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WGEMM_SMEM);
    cudaFuncSetAttribute(k_wupd_downdate, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WGEMM_SMEM);
    cudaFuncSetAttribute(k_wupd_chol, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(k_wupd_trsm, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  });
  const int sms = sm_count();
  const int tn = (d.nmax + 1 + 63) / 64, tk = (d.Tk + 63) / 64, tp = (d.nmax + 63) / 64, ts = (d.nmax + 1 + 31) / 32;
  // enough filter lanes (grid z / y) to fill the GPU a few times over; each CTA walks the work list with that stride
  auto lanes = [&](int tiles) { return std::max(1, std::min(d.F, (4 * sms + tiles - 1) / tiles)); };
  KLAUNCH(d, KID_UPD_T, st, (k_wupd_T<<<dim3(tn, tk, lanes(tn * tk)), WT, WGEMM_SMEM, st>>>(d)));
  KLAUNCH(d, KID_UPD_S, st, (k_wupd_S<<<dim3(tk, tk, lanes(tk * (tk + 1) / 2)), WT, WGEMM_SMEM, st>>>(d)));
  KLAUNCH(d, KID_UPD_CHOL, st, (k_wupd_chol<<<std::min(d.F, sms), WT, sm_chol, st>>>(d)));
  KLAUNCH(d, KID_UPD_TRSM, st, (k_wupd_trsm<<<dim3(ts, lanes(ts)), WT, sm_trsm, st>>>(d)));
  KLAUNCH(d, KID_UPD_INJECT, st, (k_upd_inject<<<std::min(d.F, 2 * sms), WT, 0, st>>>(d)));
  KLAUNCH(d, KID_UPD_DOWNDATE, st, (k_wupd_downdate<<<dim3(tp, tp, lanes(tp * (tp + 1) / 2)), WT, WGEMM_SMEM, st>>>(d)));
}

}  // namespace msckf
