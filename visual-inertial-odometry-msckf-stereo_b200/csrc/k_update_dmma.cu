// measurement_update (msckf.py:555-604) on the FP64 tensor pipe: one CTA per filter, every dense contraction of the
// Kalman gain / covariance update issued as mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4) from shared-memory fragments.
//
// Applies when the stacked Jacobian was compressed (m > n_a, so H_thin = R is upper triangular with k = n_a rows,
// msckf.py:549-553) and k fits the class capacity KC.  Algebra as in k_update.cu:
//     T = R P[act,:]                       (k x n, DMMA; P rows gathered through the active-column map)
//     S = T[:,act] R^T + sigma^2 I         (k x k lower tiles, DMMA)
//     blocked Cholesky of S fused with the forward substitution W = L^-1 [T | r]: per 8-row panel the diagonal
//     block is factorised and inverted by one warp, panel solves and trailing updates are DMMA tiles
//     dx = W^T y, state injection (msckf.py:565-595)
//     P <- P - W^T W                       (n x n upper 16x16 warp tiles, DMMA, mirrored store)
// Shared memory: X [KC][KC+4] (R in the upper triangle, later S / L in the lower), T [KC][ldT] (T -> W, column n = r -> y).
// Leading dimensions are 4 (mod 16) doubles so that every fragment load (8 rows x 4 k or 4 k x 8 columns) is
// bank-conflict free.
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"
#include "dmma_tiles.cuh"

namespace msckf {

__host__ __device__ inline int dmma_ldT(int nmax) {       // smallest ld >= nmax + 1 with ld = 4 (mod 16)
  int l = (nmax + 15) & ~15;                              // 16-wide downdate tiles stay inside a row
  if (l < nmax + 1) l = nmax + 1;
  while ((l & 15) != 4) ++l;
  return l;
}
size_t upd_dmma_smem_bytes(int nmax, int KC) {
  const size_t dbl = (size_t)KC * (KC + 4) + (size_t)KC * dmma_ldT(nmax) + 8 * LDL + ((nmax + 7) & ~7) + KC / 2 + 8;
  return dbl * sizeof(double);
}

template <int KC, int DT>
__global__ void __launch_bounds__(DT, KC > 64 ? 1 : 512 / DT) k_upd_dmma(Dev d, int cls, int stage) {
  constexpr int LDX = KC + 4;
  constexpr int NKT = KC / 8;                  // row tiles at capacity
  constexpr int NW = DT / 32;                  // warps per CTA
  constexpr int S_PER_WARP = (NKT * (NKT + 1) / 2 + NW - 1) / NW;
  extern __shared__ double us[];
  const int ldT = dmma_ldT(d.nmax);
  double* sX = us;                              // [KC][LDX]
  double* sT = sX + KC * LDX;                   // [KC][ldT]
  double* sLinv = sT + KC * ldT;                // [8][LDL]
  double* sdx = sLinv + 8 * LDL;                // [nmax rounded to 8]
  int* scm = (int*)(sdx + ((d.nmax + 7) & ~7)); // [KC]
  __shared__ int s_bad;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, tg = lane & 3;       // fragment coordinates: group id, thread in group
  const int nlist = (int)d.stage_ctr[SC_NUPD_DMMA + cls];
  const int* wl = d.wl_upd_dmma + (size_t)cls * d.F;
  const double sigma2 = d.prm->obs_noise;

  for (int w = blockIdx.x; w < nlist; w += gridDim.x) {
    const int f = wl[w];
    const int k = d.k_rows[f], na = d.n_act[f], mp = d.m_pad[f], n = NI + 6 * d.ncam[f], ld = d.ld;
    const int kp = (k + 7) & ~7, nkt = kp >> 3, nkk = kp >> 2;     // rows of H_thin: 8-row tiles, 4-deep steps over the rows
    const int nks = ((na + 7) & ~7) >> 2;                          // 4-deep steps over the active columns
    const bool tri = d.m_rows[f] > na;                             // compressed: H_thin = R is upper triangular (k = n_a)
    const int ntn = (n + 7) >> 3;                // column tiles of the state part
    const int nct = (n + 8) >> 3;                // column tiles including the residual column n
    const double* A = d.A + (size_t)f * d.A_elems;
    double* P = cov_of(d, f);
    const int* cm = d.colmap + (size_t)f * d.nmax;

    // ---- stage H_thin (zero padded), r and the column map
    for (int e = tid; e < KC * LDX; e += DT) sX[e] = 0.0;
    for (int e = tid; e < kp * (nct * 8 - n); e += DT) {
      const int i = e / (nct * 8 - n), c = n + e - i * (nct * 8 - n);
      sT[i * ldT + c] = (c == n && i < k) ? A[(size_t)na * mp + i] : 0.0;
    }
    for (int c = tid; c < KC; c += DT) scm[c] = (c < na) ? cm[c] : 0;
    if (tid == 0) s_bad = 0;
    __syncthreads();
    for (int e = tid; e < k * na; e += DT) {
      const int c = e / k, i = e - c * k;
      sX[i * LDX + c] = A[(size_t)c * mp + i];
    }
    __syncthreads();

    // ---- T = R P[act, :] : a warp owns a column tile (B fragments of P in registers), loops over the row tiles
    for (int jt = warp; jt < ntn; jt += DT / 32) {
      const int j = 8 * jt + g;
      double bf[KC / 4];
#pragma unroll
      for (int ks = 0; ks < KC / 4; ++ks) {
        const int c = 4 * ks + tg;
        bf[ks] = (ks < nks && c < na && j < n) ? P[(size_t)scm[c] * ld + j] : 0.0;
      }
#pragma unroll
      for (int it = 0; it < NKT; ++it) {
        if (it < nkt) {
          double c0 = 0.0, c1 = 0.0;
          const double* arow = sX + (8 * it + g) * LDX + tg;
#pragma unroll
          for (int ks = 0; ks < KC / 4; ++ks)
            if (ks < nks && (!tri || ks >= 2 * it)) dmma884(c0, c1, arow[4 * ks], bf[ks]);
          const int col = 8 * jt + 2 * tg;
          double* trow = sT + (8 * it + g) * ldT;
          if (col < n) trow[col] = c0;
          if (col + 1 < n) trow[col + 1] = c1;
        }
      }
    }
    __syncthreads();

    // ---- S = T[:, act] R^T + sigma^2 I, lower tiles, kept in registers until every read of R is done
    const int nst = nkt * (nkt + 1) / 2;
    double sv0[S_PER_WARP], sv1[S_PER_WARP];
    int sit, sjt;                                  // tile of this warp: decoded once, then stepped (tri_advance)
    tri_decode(warp, sit, sjt);
#pragma unroll
    for (int q = 0; q < S_PER_WARP; ++q) {
      sv0[q] = sv1[q] = 0.0;
      const int t = warp + NW * q;
      if (t < nst) {
        const int it = sit, jt = sjt;
        tri_advance(sit, sjt, NW);
        const double* trow = sT + (8 * it + g) * ldT;
        const double* brow = sX + (8 * jt + g) * LDX + tg;
        for (int ks = tri ? 2 * jt : 0; ks < nks; ++ks) dmma884(sv0[q], sv1[q], trow[scm[4 * ks + tg]], brow[4 * ks]);
      }
    }
    __syncthreads();
    tri_decode(warp, sit, sjt);
#pragma unroll
    for (int q = 0; q < S_PER_WARP; ++q) {
      const int t = warp + NW * q;
      if (t < nst) {
        const int it = sit, jt = sjt;
        tri_advance(sit, sjt, NW);
        const int row = 8 * it + g, col = 8 * jt + 2 * tg;
        sX[row * LDX + col] = sv0[q] + (row == col ? sigma2 : 0.0);
        sX[row * LDX + col + 1] = sv1[q] + (row == col + 1 ? sigma2 : 0.0);
      }
    }
    __syncthreads();

    // ---- blocked Cholesky of S fused with W = L^-1 [T | r]  (right looking, 8-row panels)
    bool bad = false;
    for (int p = 0; p < nkt; ++p) {
      const int p0 = 8 * p;
      if (warp == 0) {
        if (chol8_inverse(sX, LDX, p0, sLinv, lane) && lane == 0) s_bad = 1;
      }
      __syncthreads();
      if (s_bad) { bad = true; break; }
      const double li0 = sLinv[g * LDL + tg], li1 = sLinv[g * LDL + 4 + tg];   // A fragment of L^-1 == B fragment of L^-T
      const int nbelow = nkt - p - 1;
      // panel: W_p = L_pp^-1 T_p (column tiles), L_ip = S_ip L_pp^-T (row tiles below)
      for (int item = warp; item < nct + nbelow; item += DT / 32) {
        double c0 = 0.0, c1 = 0.0;
        if (item < nct) {
          double* base = sT + p0 * ldT + 8 * item;
          const double b0 = base[tg * ldT + g], b1 = base[(4 + tg) * ldT + g];
          dmma884(c0, c1, li0, b0);
          dmma884(c0, c1, li1, b1);
          base[g * ldT + 2 * tg] = c0;
          base[g * ldT + 2 * tg + 1] = c1;
        } else {
          double* base = sX + (8 * (p + 1 + item - nct)) * LDX + p0;
          const double a0 = base[g * LDX + tg], a1 = base[g * LDX + 4 + tg];
          dmma884(c0, c1, a0, li0);
          dmma884(c0, c1, a1, li1);
          base[g * LDX + 2 * tg] = c0;
          base[g * LDX + 2 * tg + 1] = c1;
        }
      }
      __syncthreads();
      // trailing update: T_i -= L_ip W_p ;  S_ij -= L_ip L_jp^T  (p < j <= i)
      const int nT = nbelow * nct, nS = nbelow * (nbelow + 1) / 2;
      int tib = 0, tjb = 0;
      bool tdec = false;
      for (int item = warp; item < nT + nS; item += DT / 32) {
        if (item < nT) {
          const int ib = item / nct, jt = item - ib * nct, it = p + 1 + ib;
          const double* lrow = sX + (8 * it + g) * LDX + p0 + tg;
          const double a0 = -lrow[0], a1 = -lrow[4];
          const double* wb = sT + p0 * ldT + 8 * jt;
          const double b0 = wb[tg * ldT + g], b1 = wb[(4 + tg) * ldT + g];
          double* cp = sT + (8 * it + g) * ldT + 8 * jt + 2 * tg;
          double c0 = cp[0], c1 = cp[1];
          dmma884(c0, c1, a0, b0);
          dmma884(c0, c1, a1, b1);
          cp[0] = c0;
          cp[1] = c1;
        } else {
          if (!tdec) { tri_decode(item - nT, tib, tjb); tdec = true; }
          const int ib = tib, jb = tjb;
          tri_advance(tib, tjb, DT / 32);
          const int it = p + 1 + ib, jt = p + 1 + jb;
          const double* lrow = sX + (8 * it + g) * LDX + p0 + tg;
          const double a0 = -lrow[0], a1 = -lrow[4];
          const double* lcol = sX + (8 * jt + g) * LDX + p0 + tg;
          const double b0 = lcol[0], b1 = lcol[4];
          double* cp = sX + (8 * it + g) * LDX + 8 * jt + 2 * tg;
          double c0 = cp[0], c1 = cp[1];
          dmma884(c0, c1, a0, b0);
          dmma884(c0, c1, a1, b1);
          cp[0] = c0;
          cp[1] = c1;
        }
      }
      __syncthreads();
    }
    if (bad) {
      if (tid == 0) raise_not_spd(d, 4, f, k, na, d.m_rows[f]);
      __syncthreads();
      continue;
    }

    // ---- dx = W^T y, state injection (msckf.py:565-595)
    for (int j = tid; j < n; j += DT) {
      double s = 0.0;
      for (int i = 0; i < kp; ++i) s += sT[i * ldT + j] * sT[i * ldT + n];
      sdx[j] = s;
    }
    __syncthreads();
    inject_state(d, f, sdx, tid, DT);
    __syncwarp();

    // ---- P <- P - W^T W : 16x16 warp tiles over the upper triangle, mirrored store
    // In the pruning stage the two camera states being removed (msckf.py:803-815) are dropped on the way out: the result is
    // written compacted into the filter's other covariance buffer, which becomes the current one (k_remove_cams then only
    // compacts the camera table and P is not streamed a second time).
    const bool compact = stage == 1 && d.rm_idx[2 * f] >= 0;
    const int cs0 = compact ? NI + 6 * d.rm_idx[2 * f] : 1 << 30, cs1 = compact ? NI + 6 * d.rm_idx[2 * f + 1] : 1 << 30;
    double* Pout = compact ? cov_other(d, f) : P;
    auto cmap = [&](int i) { return i < cs0 ? i : (i < cs0 + 6 ? -1 : (i < cs1 ? i - 6 : (i < cs1 + 6 ? -1 : i - 12))); };
    const int nb = (n + 15) >> 4, nblk = nb * (nb + 1) / 2;
    // the P values of the warp's next tile are requested before this tile's DMMAs and stores (tiles are disjoint, and the
    // mirrored stores only touch the strict lower triangle, which nobody reads)
    double pv[2][2][2], pn[2][2][2];
    auto load_tile = [&](int bj, int bi, double (&o)[2][2][2]) {
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int y = 0; y < 2; ++y) {
          const int a = 16 * bi + 8 * x + g, b = 16 * bj + 8 * y + 2 * tg;
          o[x][y][0] = (a < n && b < n) ? P[(size_t)a * ld + b] : 0.0;
          o[x][y][1] = (a < n && b + 1 < n) ? P[(size_t)a * ld + b + 1] : 0.0;
        }
    };
    int dbj, dbi;                                  // this warp's tile (bi <= bj), stepped from tile to tile
    tri_decode(warp, dbj, dbi);
    if (warp < nblk) load_tile(dbj, dbi, pv);
    for (int item = warp; item < nblk; item += DT / 32) {
      const int bj = dbj, bi = dbi;
      tri_advance(dbj, dbi, DT / 32);
      const int a0 = 16 * bi, b0 = 16 * bj;
      if (item + DT / 32 < nblk) load_tile(dbj, dbi, pn);
      double acc[2][2][2];
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int y = 0; y < 2; ++y) acc[x][y][0] = acc[x][y][1] = 0.0;
      const double* wa = sT + tg * ldT + a0 + g;
      const double* wb = sT + tg * ldT + b0 + g;
      for (int ks = 0; ks < nkk; ++ks) {
        const double fa0 = wa[0], fa1 = wa[8], fb0 = wb[0], fb1 = wb[8];
        dmma884(acc[0][0][0], acc[0][0][1], fa0, fb0);
        dmma884(acc[0][1][0], acc[0][1][1], fa0, fb1);
        if (bi != bj) dmma884(acc[1][0][0], acc[1][0][1], fa1, fb0);
        dmma884(acc[1][1][0], acc[1][1][1], fa1, fb1);
        wa += 4 * ldT;
        wb += 4 * ldT;
      }
      // destination indices of this thread's two rows and four columns of the tile, once per tile (-1: outside / removed)
      int ma[2], mb[2][2];
#pragma unroll
      for (int x = 0; x < 2; ++x) {
        const int a = a0 + 8 * x + g;
        ma[x] = (a < n) ? (compact ? cmap(a) : a) : -1;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int b = b0 + 8 * x + 2 * tg + e;
          mb[x][e] = (b < n) ? (compact ? cmap(b) : b) : -1;
        }
      }
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int y = 0; y < 2; ++y) {
          if (bi == bj && x == 1 && y == 0) continue;
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int a = a0 + 8 * x + g, b = b0 + 8 * y + 2 * tg + e;
            if (b >= a && ma[x] >= 0 && mb[y][e] >= 0) {
              const double v = pv[x][y][e] - acc[x][y][e];
              Pout[(size_t)ma[x] * ld + mb[y][e]] = v;
              Pout[(size_t)mb[y][e] * ld + ma[x]] = v;
            }
          }
        }
#pragma unroll
      for (int x = 0; x < 2; ++x)
#pragma unroll
        for (int y = 0; y < 2; ++y) { pv[x][y][0] = pn[x][y][0]; pv[x][y][1] = pn[x][y][1]; }
    }
    __syncthreads();
    if (compact && tid == 0) {
      d.Pcur[f] = Pout;
      d.p_compacted[f] = 1;
    }
  }
}

template <int KC, int DT>
static void launch_upd_class(const Dev& d, int cls, int kid, cudaStream_t st, int stage) {
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_upd_dmma<KC, DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  });
  const int sms = sm_count();
  const size_t sm = upd_dmma_smem_bytes(d.nmax, KC);
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(512 / DT, (224 * 1024) / (sm + 1024)));    // 128 registers per thread
  KLAUNCH(d, kid, st, (k_upd_dmma<KC, DT><<<std::min(d.F, sms * per_sm), DT, sm, st>>>(d, cls, stage)));
}

// update class c = compression class c: k, n_a + 1 <= 16 / 32 / 48 / 64 / 80
void launch_upd_dmma_class(const Dev& d, int cls, cudaStream_t st, int stage) {
  if (!d.upd_dmma_ok[cls]) return;
  if (cls == 0) launch_upd_class<16, 128>(d, 0, KID_UPD_DMMA16, st, stage);
  else if (cls == 1) launch_upd_class<32, 128>(d, 1, KID_UPD_DMMA32, st, stage);
  else if (cls == 2) launch_upd_class<48, 256>(d, 2, KID_UPD_DMMA48, st, stage);
  else if (cls == 3) launch_upd_class<64, 256>(d, 3, KID_UPD_DMMA64, st, stage);
  else launch_upd_class<80, 256>(d, 4, KID_UPD_DMMA80, st, stage);       // n_a = 66 .. 78: one filter in a few thousand updates, but the tiled path costs it 0.3 ms
}
void launch_update_dmma(const Dev& d, Fork& fk) {
  for (int c = 0; c < UPD_DMMA_CLASSES; ++c) launch_upd_dmma_class(d, c, fk.lane(c == 3 ? 0 : c + 1), 0);
}

}  // namespace msckf
