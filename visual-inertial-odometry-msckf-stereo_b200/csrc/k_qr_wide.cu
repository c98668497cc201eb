// Stacked-Jacobian compression of measurement_update (msckf.py:549-553, numba _fastQR jit_utils.py:173-175) for stacks
// that are too wide for the one-CTA classes of k_qr_dmma (more than 79 active columns: windows of 14 .. 64 camera states)
// or too tall for one CTA to be reasonable (the 1e5-row stress frame): a tall-skinny QR (TSQR) over the whole grid.
//
//   leaves   k_qr_wide<RPL>(level = -1): the rows of a filter's stack are cut into nleaf contiguous pieces; one CTA per
//            piece expands its rows chunk by chunk from the per-landmark records into shared memory and folds them into
//            the piece's own triangular factor (the tpqrt scheme of k_qr_dmma.cu: panel of 8 reflectors by one warp,
//            compact-WY trailing update W = R + V^T C, Y = T^T W, R -= Y, C -= V Y on DMMA tiles);
//   tree     k_qr_wide<RPL>(level = 0, 1, ...): node a folds the factor of node a + 2^level into its own (the partner's
//            rows arrive as chunks by bulk copy), log2(nleaf) launches;
//   finish   k_qr_wide_finish: H_thin = R of node 0, r_thin = its last column, into the filter's compressed slot.
//
// The running factor R does not fit shared memory next to a chunk (185 x 185 doubles at a 30-state window, 393 x 393 at
// 64), so it lives in L2 / HBM and only the 8-row slab a panel step touches is resident: slabs p+1 and p+2 are fetched by the TMA
// engine (cp.async.bulk + mbarrier, two buffers) while panel p is applied, updated R tiles go straight back to global.
// Look-ahead: the warp that factors panels first applies panel p to the columns of panel p+1 and factors it while the
// other warps finish the trailing tiles of panel p -- one CTA barrier per panel and no warp waits for the factorisation.
// Row signs of R differ from a sequential Householder QR; the update is invariant to them (as in k_qr_tall.cu).
#include <algorithm>
#include <cstdlib>
#include "msckf_dev.cuh"
#include "kernels.h"
#include "tma.cuh"
#include "qr_panel.cuh"

namespace msckf {

constexpr int QW_T = 256;
constexpr int QW_LDL = 12;         // leading dimension of T and of the per-warp W / Y staging tiles
constexpr unsigned QWFULL = 0xffffffffu;
constexpr int QW_NSLAB = 3;       // slabs of R in flight at most: the one being applied, the one the look-ahead panel needs, one more on its way

__host__ __device__ inline int qw_cap(int na) { return (na + 1 + 7) & ~7; }          // columns of [H | r], padded to 8
// leaves of a stack of m rows: pieces of at least 512 rows, at most leaf_cap of them
__host__ __device__ inline int qw_rows_per_leaf(int m, int leaf_cap, int CR) {
  int nl = (m + 511) / 512;
  nl = nl < 1 ? 1 : (nl > leaf_cap ? leaf_cap : nl);
  const int per = (m + nl - 1) / nl;
  return (per + CR - 1) / CR * CR;
}

__host__ __device__ inline size_t qw_smem_bytes(int cap, int CR) {
  const size_t LP = cap + 4;
  return ((size_t)CR * LP + QW_NSLAB * 8 * LP + 2 * 8 * QW_LDL + 8 * 8 * QW_LDL) * sizeof(double) + 64 * sizeof(int) + 64;
}
// rows per chunk of a stack with `cap` columns: as many as the shared memory of the launch holds next to the slabs
// (128, 64 or 32: the panel warp keeps 4, 2 or 1 rows per lane)
__host__ __device__ inline int qw_chunk_rows(int cap, size_t smem_bytes) {
  for (int cr = 128; cr > 32; cr >>= 1)
    if (qw_smem_bytes(cap, cr) <= smem_bytes) return cr;
  return 32;
}

// Panel factorisation by one warp (qr_panel.cuh).  The 8 x 8 diagonal block of R comes from the shared-memory slab (row jj
// of the slab = row j0 + jj of R), its new values go to global memory.  V replaces the panel columns of the chunk.
template <int RPL, bool FOLD>
__device__ __forceinline__ void qw_panel(const double* slab, double* Rg, int ldg, double* sC, double* sTm, int LP, int j0, int nrefl,
                                         bool need_T, int lane) {
  qr_panel<RPL, FOLD>(slab + j0, LP, Rg + (size_t)j0 * ldg + j0, ldg, sC + j0, LP, sTm, QW_LDL, nrefl, need_T, lane);
}

// Applies the block reflector of panel j0 (V in the chunk, T in sTm) to the 8-column tile at c0: R rows from the slab,
// new R rows to global memory, chunk columns updated in place.
__device__ __forceinline__ void qw_apply(const double* slab, double* Rg, int ldg, double* sC, const double* sTm, double* sYw, int LP, int j0,
                                         int c0, int crk, int crt, int g, int tg) {
  const double r_old0 = slab[g * LP + c0 + 2 * tg], r_old1 = slab[g * LP + c0 + 2 * tg + 1];
  double w0[4] = {r_old0, 0.0, 0.0, 0.0}, w1[4] = {r_old1, 0.0, 0.0, 0.0};
  const double* va = sC + (size_t)tg * LP + j0 + g;
  const double* cb = sC + (size_t)tg * LP + c0 + g;
  int ks = 0;
  for (; ks + 4 <= crk; ks += 4) {
#pragma unroll
    for (int u = 0; u < 4; ++u) dmma(w0[u], w1[u], va[(size_t)(4 * (ks + u)) * LP], cb[(size_t)(4 * (ks + u)) * LP]);
  }
  for (; ks < crk; ++ks) dmma(w0[0], w1[0], va[(size_t)4 * ks * LP], cb[(size_t)4 * ks * LP]);
  const double ww0 = (w0[0] + w0[1]) + (w0[2] + w0[3]), ww1 = (w1[0] + w1[1]) + (w1[2] + w1[3]);
  // Y = T^T W
  sYw[g * QW_LDL + 2 * tg] = ww0;
  sYw[g * QW_LDL + 2 * tg + 1] = ww1;
  __syncwarp();
  const double wb0 = sYw[tg * QW_LDL + g], wb1 = sYw[(4 + tg) * QW_LDL + g];
  const double ta0 = sTm[tg * QW_LDL + g], ta1 = sTm[(4 + tg) * QW_LDL + g];
  double y0 = 0.0, y1 = 0.0;
  dmma(y0, y1, ta0, wb0);
  dmma(y0, y1, ta1, wb1);
  *(double2*)(Rg + (size_t)(j0 + g) * ldg + c0 + 2 * tg) = make_double2(r_old0 - y0, r_old1 - y1);
  __syncwarp();
  sYw[g * QW_LDL + 2 * tg] = y0;
  sYw[g * QW_LDL + 2 * tg + 1] = y1;
  __syncwarp();
  const double yb0 = sYw[tg * QW_LDL + g], yb1 = sYw[(4 + tg) * QW_LDL + g];
  // C(:, tile) -= V Y
  for (int rt = 0; rt < crt; ++rt) {
    const double* vr = sC + (size_t)(8 * rt + g) * LP + j0 + tg;
    double2* cp = (double2*)(sC + (size_t)(8 * rt + g) * LP + c0 + 2 * tg);
    double2 cv = *cp;
    dmma(cv.x, cv.y, -vr[0], yb0);
    dmma(cv.x, cv.y, -vr[4], yb1);
    *cp = cv;
  }
  __syncwarp();
}

template <bool FOLD>
__device__ __forceinline__ void qw_panel_rows(int rpl, const double* slab, double* Rg, int ldg, double* sC, double* sTm, int LP, int j0,
                                              int nrefl, bool need_T, int lane) {
  if (rpl == 4) qw_panel<4, FOLD>(slab, Rg, ldg, sC, sTm, LP, j0, nrefl, need_T, lane);
  else if (rpl == 2) qw_panel<2, FOLD>(slab, Rg, ldg, sC, sTm, LP, j0, nrefl, need_T, lane);
  else qw_panel<1, FOLD>(slab, Rg, ldg, sC, sTm, LP, j0, nrefl, need_T, lane);
}

template <bool FOLD>
__global__ void __launch_bounds__(QW_T, 1) k_qr_wide(Dev d, int level, unsigned smem_bytes, int nslab) {
  extern __shared__ __align__(128) double qs[];
  __shared__ uint64_t s_slab[QW_NSLAB], s_rows;
  __shared__ int s_c0;
  const int ldg = d.qr_capcols;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3;
  if (tid == 0) {
    for (int b = 0; b < QW_NSLAB; ++b) mbar_init(&s_slab[b], 1);
    mbar_init(&s_rows, 1);
    fence_mbar_init();
  }
  __syncthreads();
  unsigned su[QW_NSLAB] = {0u, 0u, 0u}, srows = 0u;   // completed uses of each barrier (phase parity)
  const int nlist = (int)d.stage_ctr[SC_NQR_BUILD];
  const int leaf_cap = d.qr_leaf_cap;

  for (int item = blockIdx.x; item < nlist * leaf_cap; item += gridDim.x) {
    const int f = d.wl_qrb[item / leaf_cap], node = item % leaf_cap;
    const int m = d.m_rows[f], na = d.n_act[f];
    if (m <= na) continue;                            // not compressed (msckf.py:549): k_qr_build copies the rows out
    const int cap = qw_cap(na), nct = cap >> 3, npan = (na + 7) >> 3;
    // shared-memory layout of this stack: the chunk is as tall as its own width allows
    const int LP = cap + 4, CR = qw_chunk_rows(cap, smem_bytes), RPL = CR >> 5;
    double* sC = qs;                                    // [CR][LP]   chunk of stacked rows
    double* sSlab = sC + (size_t)CR * LP;               // [QW_NSLAB][8][LP] the 8 rows of R a panel step reads
    double* sTm = sSlab + QW_NSLAB * 8 * LP;            // [2][8][QW_LDL]
    double* sY = sTm + 2 * 8 * QW_LDL;                  // [8 warps][8][QW_LDL]
    int* camcol = (int*)(sY + 8 * 8 * QW_LDL);          // [64] logical camera -> active column block
    double* sYw = sY + warp * 8 * QW_LDL;
    const int rpl = qw_rows_per_leaf(m, leaf_cap, CR), nleaf = (m + rpl - 1) / rpl;
    int rb = 0, re = 0;
    const double* Rb = nullptr;
    if (level < 0) {
      if (node >= nleaf) continue;
      rb = node * rpl;
      re = min(m, rb + rpl);
    } else {
      if ((node & ((2 << level) - 1)) != 0 || node + (1 << level) >= nleaf) continue;
      Rb = d.Rnode + ((size_t)f * leaf_cap + node + (1 << level)) * d.Rnode_stride;
      re = na;                                        // the partner's factor has n_a rows
    }
    double* Rg = d.Rnode + ((size_t)f * leaf_cap + node) * d.Rnode_stride;
    const size_t fb = (size_t)f * d.Fcap;
    const int* cm = d.colmap + (size_t)f * d.nmax;
    const int* rmap = d.rowmap + (size_t)f * d.rowmap_cap;
    __syncthreads();
    if (level < 0) {
      for (int e = tid; e < cap * (ldg / 2); e += QW_T) ((double2*)Rg)[e] = make_double2(0.0, 0.0);
      for (int c = tid; c < 64; c += QW_T) camcol[c] = -1;
      fence_proxy_async();
      __syncthreads();
      for (int b = tid; b < na / 6; b += QW_T) camcol[(cm[6 * b] - NI) / 6] = b;
    }

    for (int r0 = rb; r0 < re; r0 += CR) {
      const int cr = min(CR, re - r0);
      int pb0;
      __syncthreads();                                // the previous chunk is folded (and its R stores fenced)
      if (level < 0) {
        // ---- expand the rows of this chunk from the landmark records (zero fill, then 6-column blocks and the residual)
        for (int e = tid; e < CR * LP / 2; e += QW_T) ((double2*)sC)[e] = make_double2(0.0, 0.0);
        if (tid == 0) s_c0 = na;
        __syncthreads();
        for (int i = tid; i < cr; i += QW_T) {
          const int rm = rmap[r0 + i];
          const int kk = rm >> 8, li = rm & 255;
          const int rows = d.item_rows[fb + kk], M = (rows + 3) / 4;
          unsigned long long lmask = d.item_lmask[fb + kk];
          const double* rec = d.pool + d.item_off[fb + kk] + (size_t)li * (6 * M + 1);
          double* row = sC + (size_t)i * LP;
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
        pb0 = s_c0 >> 3;                              // columns before the chunk's first non-zero column need no elimination
      } else {
        // ---- rows r0 .. r0 + cr of the partner's factor, columns r0 .. cap (zero before the diagonal), by bulk copy
        pb0 = r0 >> 3;
        for (int e = tid; e < (CR - cr) * LP; e += QW_T) sC[(size_t)cr * LP + e] = 0.0;
        if (tid == 0) mbar_arrive_expect_tx(&s_rows, (unsigned)((size_t)cr * (cap - r0) * sizeof(double)));
        __syncthreads();
        for (int i = tid; i < cr; i += QW_T) bulk_g2s(sC + (size_t)i * LP + r0, Rb + (size_t)(r0 + i) * ldg + r0, (cap - r0) * sizeof(double), &s_rows);
        mbar_wait(&s_rows, srows & 1u);
        srows += 1;
      }
      auto issue_slab = [&](int pb) {                 // warp 1, lanes 0..7: row i of slab pb, columns 8 pb .. cap
        const int b = pb % nslab;
        if (lane == 0) mbar_arrive_expect_tx(&s_slab[b], (unsigned)(8 * (cap - 8 * pb) * sizeof(double)));
        __syncwarp();
        if (lane < 8) bulk_g2s(sSlab + (size_t)(b * 8 + lane) * LP + 8 * pb, Rg + (size_t)(8 * pb + lane) * ldg + 8 * pb, (cap - 8 * pb) * sizeof(double), &s_slab[b]);
      };
      if (warp == 1) {
        issue_slab(pb0);
        if (nslab > 2 && pb0 + 1 < npan) issue_slab(pb0 + 1);
      }
      if (warp == 0) {
        const int sb = pb0 % nslab;
        mbar_wait(&s_slab[sb], su[sb] & 1u);
        qw_panel_rows<FOLD>(RPL, sSlab + (size_t)sb * 8 * LP, Rg, ldg, sC, sTm + (pb0 & 1) * 8 * QW_LDL, LP, 8 * pb0, min(8, na - 8 * pb0), pb0 + 1 < nct, lane);
      }
      __syncthreads();
      for (int pb = pb0; pb < npan; ++pb) {
        const int cur = pb & 1, j0 = 8 * pb, sb = pb % nslab, sbn = (pb + 1) % nslab;
        // rows of the chunk that can be non-zero in this panel: all of them for stacked landmark rows; a triangular factor
        // being folded (tree levels) has nothing below its diagonal, so row r0 + i only matters once 8 pb + 7 >= r0 + i
        const int rows_nz = level < 0 ? cr : min(cr, j0 + 8 - r0);
        const int crk = (rows_nz + 3) >> 2, crt = (rows_nz + 7) >> 3;
        const double* slab = sSlab + (size_t)sb * 8 * LP;
        const double* Tm = sTm + cur * 8 * QW_LDL;
        const bool ahead = pb + 1 < npan;             // another panel follows: warp 0 looks ahead
        if (warp == 1 && pb + nslab - 1 < npan) issue_slab(pb + nslab - 1);   // nslab - 1 slabs ahead
        mbar_wait(&s_slab[sb], su[sb] & 1u);          // (already complete for warp 0)
        if (ahead) {
          if (warp == 0) {
            qw_apply(slab, Rg, ldg, sC, Tm, sYw, LP, j0, j0 + 8, crk, crt, g, tg);
            mbar_wait(&s_slab[sbn], su[sbn] & 1u);
            qw_panel_rows<FOLD>(RPL, sSlab + (size_t)sbn * 8 * LP, Rg, ldg, sC, sTm + (cur ^ 1) * 8 * QW_LDL, LP, j0 + 8, min(8, na - j0 - 8), pb + 2 < nct, lane);
          } else {
            for (int ct = pb + 2 + (warp - 1); ct < nct; ct += QW_T / 32 - 1) qw_apply(slab, Rg, ldg, sC, Tm, sYw, LP, j0, 8 * ct, crk, crt, g, tg);
          }
        } else {
          for (int ct = pb + 1 + warp; ct < nct; ct += QW_T / 32) qw_apply(slab, Rg, ldg, sC, Tm, sYw, LP, j0, 8 * ct, crk, crt, g, tg);
        }
        su[sb] += 1;
        if (pb + 1 == npan) fence_proxy_async();      // R rows written with ordinary stores are read back by the TMA engine
        __syncthreads();
      }
    }
  }
}

// H_thin = R (k = n_a rows), r_thin = first n_a entries of Q^T r, from the root of the filter's tree
__global__ void __launch_bounds__(QW_T) k_qr_wide_finish(Dev d) {
  const int nlist = (int)d.stage_ctr[SC_NQR_BUILD];
  for (int w = blockIdx.x; w < nlist; w += gridDim.x) {
    const int f = d.wl_qrb[w];
    const int m = d.m_rows[f], na = d.n_act[f], nc = na + 1, mp = d.m_pad[f];
    if (m <= na) continue;
    const double* Rg = d.Rnode + (size_t)f * d.qr_leaf_cap * d.Rnode_stride;
    const int ldg = d.qr_capcols;
    double* A = d.A + (size_t)f * d.A_elems;
    for (int e = threadIdx.x; e < na * nc; e += QW_T) {
      const int i = e % na, c = e / na;
      A[(size_t)c * mp + i] = (c >= i) ? Rg[(size_t)i * ldg + c] : 0.0;
    }
  }
}

int qr_wide_rows_per_chunk(int Nmax) {                // of the widest possible stack; narrower stacks get taller chunks
  return qw_chunk_rows(qw_cap(6 * Nmax), 226 * 1024);
}
int qr_wide_leaf_rows(int m, int leaf_cap, int CR) { return qw_rows_per_leaf(m, leaf_cap, CR); }

void launch_qr_wide(const Dev& d, cudaStream_t st) {
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_qr_wide<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    cudaFuncSetAttribute(k_qr_wide<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
  });
  const bool fold = getenv("MSCKF_QW_NOFOLD") == nullptr;                  // development knobs (A/B on one box)
  const int nslab = getenv("MSCKF_QW_SLABS") ? std::max(2, std::min(QW_NSLAB, atoi(getenv("MSCKF_QW_SLABS")))) : QW_NSLAB;
  const int sms = sm_count();
  // a 128-row chunk of the widest stack if that fits, else everything the SM has: every stack then sizes its own chunk
  const size_t sm = std::min<size_t>(qw_smem_bytes(d.qr_capcols, 128), 226 * 1024);
  const int grid = (int)std::min<long long>((long long)d.F * d.qr_leaf_cap, 2LL * sms);
  for (int level = -1; level < d.qr_levels; ++level)
    if (fold) KLAUNCH(d, KID_QR_WIDE, st, (k_qr_wide<true><<<grid, QW_T, sm, st>>>(d, level, (unsigned)sm, nslab)));
    else KLAUNCH(d, KID_QR_WIDE, st, (k_qr_wide<false><<<grid, QW_T, sm, st>>>(d, level, (unsigned)sm, nslab)));
  KLAUNCH(d, KID_QR_WIDE, st, (k_qr_wide_finish<<<std::min(d.F, 2 * sms), QW_T, 0, st>>>(d)));
}

}  // namespace msckf
