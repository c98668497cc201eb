// Stacked-Jacobian compression of measurement_update (msckf.py:549-553): blocked Householder QR of the m x (n_a + 1)
// matrix [H | r] on the FP64 tensor pipe, one CTA per filter, for compressed updates (m > n_a) with n_a + 1 <= 64.
//
// The stacked matrix is never materialised in HBM.  Rows are expanded from the compact per-landmark records of
// k_jacobian_gate into a shared-memory chunk C (128 rows) and folded into the running factor R (shared memory,
// Q^T r riding along as column n_a) with the triangular-on-top-of-pentagonal scheme of LAPACK's tpqrt, 8 columns at
// a time:
//   1. panel: warp 0 holds the 128 x 8 panel of C in registers (4 rows per lane) and applies up to 8 Householder
//      reflectors v = [e_j ; C(:,j) / (alpha - beta)] with one batched warp reduction per reflector, then forms the
//      8 x 8 triangular factor T of the compact-WY form I - V T V^T from V^T V;
//   2. trailing columns, one 8-column tile per warp:  W = R(panel rows, tile) + V^T C(:, tile)   (DMMA, K = rows)
//                                                   Y = T^T W                                   (DMMA)
//                                                   R(panel rows, tile) -= Y,  C(:, tile) -= V Y (DMMA, K = 8).
// Rows of C and R have a leading dimension of 4 (mod 16) doubles: every fragment load is bank-conflict free.
// Look-ahead: the panel warp first applies panel p to the columns of panel p+1 and factors it while the other warps
// finish the trailing tiles of panel p (T is double-buffered) -- one CTA barrier per panel, nobody waits for the panel.
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"
#include "qr_panel.cuh"

namespace msckf {

constexpr int QD_T = 256;          // threads per CTA
constexpr int QD_CR = 128;         // rows per chunk (4 per lane of the panel warp)
constexpr int QD_LDL = 12;         // leading dimension of T and of the per-warp W / Y staging tiles
constexpr unsigned QFULL = 0xffffffffu;

__device__ __forceinline__ void qd_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double qd_allsum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(QFULL, v, o);
  return v;
}

size_t qr_dmma_smem_bytes(int cap) {
  const size_t ldr = cap + 4;
  const size_t dbl = (size_t)(cap + QD_CR) * ldr + 2 * 8 * QD_LDL + 8 * 8 * QD_LDL + 8;
  return dbl * sizeof(double) + 64 * sizeof(int) + 16;
}

// Panel factorisation by one warp (qr_panel.cuh): the diagonal block of R is read and written in place.
__device__ __forceinline__ void qd_panel(double* sR, double* sC, double* sTm, int ldr, int j0, int nrefl, bool need_T, int lane) {
  double* Rd = sR + j0 * ldr + j0;
  qr_panel<4>(Rd, ldr, Rd, ldr, sC + j0, ldr, sTm, QD_LDL, nrefl, need_T, lane);
}

// Applies the block reflector of panel j0 (V in the chunk, T in sTm) to the 8-column tile at c0 (R rows and chunk columns).
__device__ __forceinline__ void qd_apply(double* sR, double* sC, const double* sTm, double* sYw, int ldr, int j0, int c0, int crk, int crt,
                                         int g, int tg) {
  double* rp = sR + (j0 + g) * ldr + c0 + 2 * tg;
  const double r_old0 = rp[0], r_old1 = rp[1];
  double w0[4] = {r_old0, 0.0, 0.0, 0.0}, w1[4] = {r_old1, 0.0, 0.0, 0.0};
  const double* va = sC + tg * ldr + j0 + g;
  const double* cb = sC + tg * ldr + c0 + g;
  int ks = 0;
  for (; ks + 4 <= crk; ks += 4) {
#pragma unroll
    for (int u = 0; u < 4; ++u) qd_dmma(w0[u], w1[u], va[(4 * (ks + u)) * ldr], cb[(4 * (ks + u)) * ldr]);
  }
  for (; ks < crk; ++ks) qd_dmma(w0[0], w1[0], va[4 * ks * ldr], cb[4 * ks * ldr]);
  const double ww0 = (w0[0] + w0[1]) + (w0[2] + w0[3]), ww1 = (w1[0] + w1[1]) + (w1[2] + w1[3]);
  // Y = T^T W
  sYw[g * QD_LDL + 2 * tg] = ww0;
  sYw[g * QD_LDL + 2 * tg + 1] = ww1;
  __syncwarp();
  const double wb0 = sYw[tg * QD_LDL + g], wb1 = sYw[(4 + tg) * QD_LDL + g];
  const double ta0 = sTm[tg * QD_LDL + g], ta1 = sTm[(4 + tg) * QD_LDL + g];
  double y0 = 0.0, y1 = 0.0;
  qd_dmma(y0, y1, ta0, wb0);
  qd_dmma(y0, y1, ta1, wb1);
  rp[0] = r_old0 - y0;
  rp[1] = r_old1 - y1;
  __syncwarp();
  sYw[g * QD_LDL + 2 * tg] = y0;
  sYw[g * QD_LDL + 2 * tg + 1] = y1;
  __syncwarp();
  const double yb0 = sYw[tg * QD_LDL + g], yb1 = sYw[(4 + tg) * QD_LDL + g];
  // C(:, tile) -= V Y
  for (int rt = 0; rt < crt; ++rt) {
    const double* vr = sC + (8 * rt + g) * ldr + j0 + tg;
    double2* cp = (double2*)(sC + (8 * rt + g) * ldr + c0 + 2 * tg);
    double2 cv = *cp;
    qd_dmma(cv.x, cv.y, -vr[0], yb0);
    qd_dmma(cv.x, cv.y, -vr[4], yb1);
    *cp = cv;
  }
  __syncwarp();
}

__global__ void __launch_bounds__(QD_T, 2) k_qr_dmma(Dev d, int cls, int cap) {
  extern __shared__ __align__(16) double qs[];
  const int ldr = cap + 4;
  double* sR = qs;                                   // [cap][ldr]
  double* sC = sR + cap * ldr;                       // [QD_CR][ldr]
  double* sTm = sC + QD_CR * ldr;                    // [2][8][QD_LDL]  T of the panel being applied / being factored
  double* sY = sTm + 2 * 8 * QD_LDL;                 // [8 warps][8][QD_LDL]
  int* camcol = (int*)(sY + 8 * 8 * QD_LDL);         // [64] logical camera -> active column block
  __shared__ int s_c0;
  const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = nthr >> 5;
  const int g = lane >> 2, tg = lane & 3;
  const int nlist = (int)d.stage_ctr[SC_NQR + cls];
  const int* wl = d.wl_qr + (size_t)cls * d.F;
  double* sYw = sY + warp * 8 * QD_LDL;

  for (int w = blockIdx.x; w < nlist; w += gridDim.x) {
    const int f = wl[w];
    const int m = d.m_rows[f], na = d.n_act[f], nc = na + 1, mp = d.m_pad[f];
    const int nct = (nc + 7) >> 3, npan = (na + 7) >> 3;
    const size_t fb = (size_t)f * d.Fcap;
    const int* cm = d.colmap + (size_t)f * d.nmax;
    const int* rmap = d.rowmap + (size_t)f * d.rowmap_cap;
    for (int e = tid; e < cap * ldr; e += nthr) sR[e] = 0.0;
    for (int c = tid; c < 64; c += nthr) camcol[c] = -1;
    __syncthreads();
    for (int b = tid; b < na / 6; b += nthr) camcol[(cm[6 * b] - NI) / 6] = b;

    for (int r0 = 0; r0 < m; r0 += QD_CR) {
      const int cr = min(QD_CR, m - r0);
      const int crk = (cr + 3) >> 2, crt = (cr + 7) >> 3;
      // ---- expand the rows of this chunk (zero fill, then the 6-column blocks of every observation and the residual)
      for (int e = tid; e < QD_CR * ldr / 2; e += nthr) ((double2*)sC)[e] = make_double2(0.0, 0.0);
      if (tid == 0) s_c0 = na;
      __syncthreads();
      for (int i = tid; i < cr; i += nthr) {
        const int rm = rmap[r0 + i];
        const int kk = rm >> 8, li = rm & 255;
        const int rows = d.item_rows[fb + kk], M = (rows + 3) / 4;
        unsigned long long lmask = d.item_lmask[fb + kk];
        const double* rec = d.pool + d.item_off[fb + kk] + (size_t)li * (6 * M + 1);
        double* row = sC + i * ldr;
        const int first = camcol[__ffsll((long long)lmask) - 1];
        for (int a = 0; a < M; ++a) {
          const int cam = __ffsll((long long)lmask) - 1;
          lmask &= lmask - 1;
          double* dst = row + 6 * camcol[cam];
#pragma unroll
          for (int e = 0; e < 6; ++e) dst[e] = rec[6 * a + e];
        }
        row[na] = rec[6 * M];
        atomicMin(&s_c0, 6 * first);
      }
      __syncthreads();
      const int pb0 = s_c0 >> 3;                      // columns before the chunk's first non-zero column need no elimination
      if (warp == 0) qd_panel(sR, sC, sTm + (pb0 & 1) * 8 * QD_LDL, ldr, 8 * pb0, min(8, na - 8 * pb0), pb0 + 1 < nct, lane);
      __syncthreads();
      for (int pb = pb0; pb < npan; ++pb) {
        const int j0 = 8 * pb;
        const double* Tm = sTm + (pb & 1) * 8 * QD_LDL;
        if (pb + 1 < npan) {                          // another panel follows: warp 0 looks ahead
          if (warp == 0) {
            qd_apply(sR, sC, Tm, sYw, ldr, j0, j0 + 8, crk, crt, g, tg);
            qd_panel(sR, sC, sTm + ((pb + 1) & 1) * 8 * QD_LDL, ldr, j0 + 8, min(8, na - j0 - 8), pb + 2 < nct, lane);
          } else {
            for (int ct = pb + 2 + (warp - 1); ct < nct; ct += nwarp - 1) qd_apply(sR, sC, Tm, sYw, ldr, j0, 8 * ct, crk, crt, g, tg);
          }
        } else {
          for (int ct = pb + 1 + warp; ct < nct; ct += nwarp) qd_apply(sR, sC, Tm, sYw, ldr, j0, 8 * ct, crk, crt, g, tg);
        }
        __syncthreads();
      }
    }
    // ---- H_thin = R (k = n_a rows), r_thin = first n_a entries of Q^T r
    double* A = d.A + (size_t)f * d.A_elems;
    for (int e = tid; e < na * nc; e += nthr) {
      const int i = e % na, c = e / na;
      A[(size_t)c * mp + i] = (c >= i) ? sR[i * ldr + c] : 0.0;
    }
    __syncthreads();
  }
}

void launch_qr_dmma_class(const Dev& d, int c, cudaStream_t st) {
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_qr_dmma, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
  });
  const int sms = sm_count();
  const int caps[QR_DMMA_CLASSES] = {16, 32, 48, 64, 80};
  const int thr[QR_DMMA_CLASSES] = {64, 128, 128, 256, 256};
  const int kid[QR_DMMA_CLASSES] = {KID_QR_D16, KID_QR_D32, KID_QR_D48, KID_QR_D64, KID_QR_D80};
  const size_t sm = qr_dmma_smem_bytes(caps[c]);
  int per_sm = (int)std::min<size_t>((224 * 1024) / (sm + 1024), 512 / thr[c]);   // 128 registers per thread
  per_sm = std::max(1, per_sm);
  KLAUNCH(d, kid[c], st, (k_qr_dmma<<<std::min(d.F, sms * per_sm), thr[c], sm, st>>>(d, c, caps[c])));
}
void launch_qr_dmma(const Dev& d, Fork& fk) {
  for (int c = 0; c < QR_DMMA_CLASSES; ++c) launch_qr_dmma_class(d, c, fk.lane(c));
}

}  // namespace msckf
