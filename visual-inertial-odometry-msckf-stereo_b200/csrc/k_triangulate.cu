// Batched Levenberg-Marquardt landmark triangulation: Feature.check_motion,
// generate_initial_guess, initialize_position, cost and jacobian (feature.py:33-295).
// One warp per landmark; the landmark's 2M views (R, t relative to the first cam0, and the
// measurement) are staged in shared memory as a structure of arrays, lanes stride over views,
// and the 3x3 normal equations are reduced with warp shuffles.
#include <algorithm>
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int TRI_WPB = 4;          // warps (= landmarks) per CTA
constexpr int VIEW_DOUBLES = 14;    // R[9] t[3] z[2]

struct TriCfg {
  double thr, eps, precision, lam0;
  int outer_max, inner_max;
  double R01[9], t01[3];
};

__device__ __forceinline__ TriCfg load_tricfg(const Params& p) {
  TriCfg c;
  c.thr = p.tri_translation_thr; c.eps = p.huber_eps; c.precision = p.est_precision; c.lam0 = p.init_damping;
  c.outer_max = p.outer_max; c.inner_max = p.inner_max;
  for (int i = 0; i < 9; ++i) c.R01[i] = p.R01[i];
  for (int i = 0; i < 3; ++i) c.t01[i] = p.t01[i];
  return c;
}

// view v of the warp's shared block: component k at sv[k*stride + v]
struct Views {
  double* sv;
  int stride;
  __device__ __forceinline__ double& at(int k, int v) const { return sv[k * stride + v]; }
};

// np.linalg.solve for 3x3 (LAPACK dgesv: LU with partial pivoting)
__device__ __forceinline__ void solve3(const double* A_in, const double* b_in, double* x) {
  double A[9], b[3];
  for (int i = 0; i < 9; ++i) A[i] = A_in[i];
  for (int i = 0; i < 3; ++i) b[i] = b_in[i];
  for (int k = 0; k < 3; ++k) {
    int piv = k;
    double best = fabs(A[3 * k + k]);
    for (int i = k + 1; i < 3; ++i)
      if (fabs(A[3 * i + k]) > best) { best = fabs(A[3 * i + k]); piv = i; }
    if (piv != k) {
      for (int j = 0; j < 3; ++j) { const double t = A[3 * k + j]; A[3 * k + j] = A[3 * piv + j]; A[3 * piv + j] = t; }
      const double t = b[k]; b[k] = b[piv]; b[piv] = t;
    }
    for (int i = k + 1; i < 3; ++i) {
      const double l = A[3 * i + k] / A[3 * k + k];
      for (int j = k + 1; j < 3; ++j) A[3 * i + j] -= l * A[3 * k + j];
      b[i] -= l * b[k];
    }
  }
  x[2] = b[2] / A[8];
  x[1] = (b[1] - A[5] * x[2]) / A[4];
  x[0] = (b[0] - A[1] * x[1] - A[2] * x[2]) / A[0];
}

__device__ __forceinline__ double view_cost(const Views& V, int v, const double* x) {   // feature.py:33-55
  const double h0 = V.at(0, v) * x[0] + V.at(1, v) * x[1] + V.at(2, v) + x[2] * V.at(9, v);
  const double h1 = V.at(3, v) * x[0] + V.at(4, v) * x[1] + V.at(5, v) + x[2] * V.at(10, v);
  const double h2 = V.at(6, v) * x[0] + V.at(7, v) * x[1] + V.at(8, v) + x[2] * V.at(11, v);
  const double ih = 1.0 / h2;                          // one reciprocal per view (an FP64 divide is a ~35 instruction sequence)
  const double e0 = h0 * ih - V.at(12, v), e1 = h1 * ih - V.at(13, v);
  return e0 * e0 + e1 * e1;
}

// generate_initial_guess (feature.py:99-122) between view 0 and view vl (both cam0 views):
// the landmark in the frame of view 0.
__device__ __forceinline__ void initial_guess(const Views& V, int vl, double* p) {
  const double z1x = V.at(12, 0), z1y = V.at(13, 0), z2x = V.at(12, vl), z2y = V.at(13, vl);
  const double m0 = V.at(0, vl) * z1x + V.at(1, vl) * z1y + V.at(2, vl);
  const double m1 = V.at(3, vl) * z1x + V.at(4, vl) * z1y + V.at(5, vl);
  const double m2 = V.at(6, vl) * z1x + V.at(7, vl) * z1y + V.at(8, vl);
  const double a0 = m0 - z2x * m2, a1 = m1 - z2y * m2;
  const double b0 = z2x * V.at(11, vl) - V.at(9, vl), b1 = z2y * V.at(11, vl) - V.at(10, vl);
  const double depth = (a0 * b0 + a1 * b1) / (a0 * a0 + a1 * a1);
  p[0] = z1x * depth; p[1] = z1y * depth; p[2] = 1.0 * depth;
}

// The solver proper.  Views 0..nv-1 are staged in V by the caller (all lanes, then __syncwarp).
// Rf/tf = pose of the first cam0 (camera -> world).  Returns validity; writes world position.
template <int G>
__device__ __forceinline__ double grp_sum(double v, unsigned m) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(m, v, o);
  return v;
}

// G lanes (a whole warp, or 8 lanes so that four landmarks share a warp) cooperate; gmask = their lane mask,
// lane = index inside the group.
template <int G>
__device__ bool lm_solve(const Views& V, int nv, const TriCfg& cfg, const double* Rf, const double* tf, int lane, unsigned gmask,
                         double* pos_w, int* n_solves) {
  // ---- initial guess from the first and the last cam0 view (feature.py:99-122, 216-218)
  const int vl = nv - 2;
  double x[3];
  {
    double pg[3];
    initial_guess(V, vl, pg);
    x[0] = pg[0] / pg[2]; x[1] = pg[1] / pg[2]; x[2] = 1.0 / pg[2];
  }
  // ---- initial cost
  double total = 0;
  for (int v = lane; v < nv; v += G) total += view_cost(V, v, x);
  total = grp_sum<G>(total, gmask);
  // ---- normal equations at x (the reference rebuilds them every outer iteration, but x only
  //      changes on the single accepted step, after which nothing else is observable)
  double A[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
  for (int v = lane; v < nv; v += G) {   // feature.py:57-97, 241-248
    const double h1 = V.at(0, v) * x[0] + V.at(1, v) * x[1] + V.at(2, v) + x[2] * V.at(9, v);
    const double h2 = V.at(3, v) * x[0] + V.at(4, v) * x[1] + V.at(5, v) + x[2] * V.at(10, v);
    const double h3 = V.at(6, v) * x[0] + V.at(7, v) * x[1] + V.at(8, v) + x[2] * V.at(11, v);
    const double W0[3] = {V.at(0, v), V.at(1, v), V.at(9, v)};
    const double W1[3] = {V.at(3, v), V.at(4, v), V.at(10, v)};
    const double W2[3] = {V.at(6, v), V.at(7, v), V.at(11, v)};
    double J0[3], J1[3];
    const double ih3 = 1.0 / h3, u1 = h1 * ih3, u2 = h2 * ih3;     // one reciprocal of the depth per view instead of fourteen divides
    for (int i = 0; i < 3; ++i) {
      J0[i] = (W0[i] - W2[i] * u1) * ih3;
      J1[i] = (W1[i] - W2[i] * u2) * ih3;
    }
    const double r0 = u1 - V.at(12, v), r1 = u2 - V.at(13, v);
    const double e2 = r0 * r0 + r1 * r1;
    // Huber weight eps / (2 e): e only enters through 1 / e.  The reference branches on e <= eps (feature.py:91-95), e = sqrt(e2):
    // comparing the squares decides the same way except within rounding of eps^2, where the square root is taken as written.
    const double t2 = cfg.eps * cfg.eps;
    const bool inside = (e2 < t2 * (1.0 - 1e-15)) || (e2 <= t2 * (1.0 + 1e-15) && sqrt(e2) <= cfg.eps);
    const double w = inside ? 1.0 : 0.5 * cfg.eps * rsqrt(e2);
    const double ww = w * w;
    for (int i = 0; i < 3; ++i) {
      const double s0 = ww * J0[i], s1 = ww * J1[i];
      for (int j = 0; j < 3; ++j) A[3 * i + j] += s0 * J0[j] + s1 * J1[j];
      b[i] += s0 * r0 + s1 * r1;
    }
  }
  for (int i = 0; i < 9; ++i) A[i] = grp_sum<G>(A[i], gmask);
  for (int i = 0; i < 3; ++i) b[i] = grp_sum<G>(b[i], gmask);

  // ---- damped steps (feature.py:220-276).  is_cost_reduced is never cleared after a success,
  //      so at most one step is accepted; failed attempts walk lambda up in groups of inner_max.
  double lam = cfg.lam0, delta_norm = INFINITY;
  bool reduced = false;
  int outer = 0, solves = 0;
  while (outer < cfg.outer_max && delta_norm > cfg.precision && !reduced) {
    int inner = 0;
    while (inner < cfg.inner_max && !reduced) {
      double Ad[9];
      for (int i = 0; i < 9; ++i) Ad[i] = A[i];
      Ad[0] += lam; Ad[4] += lam; Ad[8] += lam;
      double delta[3];
      solve3(Ad, b, delta);
      ++solves;
      const double xn[3] = {x[0] - delta[0], x[1] - delta[1], x[2] - delta[2]};
      delta_norm = sqrt(delta[0] * delta[0] + delta[1] * delta[1] + delta[2] * delta[2]);
      double nc = 0;
      for (int v = lane; v < nv; v += G) nc += view_cost(V, v, xn);
      nc = grp_sum<G>(nc, gmask);
      if (nc < total) {
        reduced = true;
        x[0] = xn[0]; x[1] = xn[1]; x[2] = xn[2];
        total = nc;
        lam = fmax(lam / 10., 1e-10);
      } else {
        lam = fmin(lam * 10., 1e12);
      }
      ++inner;
    }
    ++outer;
  }
  if (n_solves) *n_solves = solves;

  // ---- back to Euclidean, cheirality in every view (feature.py:278-295)
  const double fp[3] = {x[0] / x[2], x[1] / x[2], 1.0 / x[2]};
  bool ok = true;
  for (int v = lane; v < nv; v += G) {
    const double z = V.at(6, v) * fp[0] + V.at(7, v) * fp[1] + V.at(8, v) * fp[2] + V.at(11, v);
    if (z <= 0) ok = false;
  }
  ok = __all_sync(gmask, ok);
  mat3_vec(Rf, fp, pos_w);
  for (int i = 0; i < 3; ++i) pos_w[i] += tf[i];
  return ok;
}

// Stage the two views of one stereo observation (feature.py:185-213).
// R0 = to_rotation(q) (world -> cam0), p = cam0 position; Rf/tf = first cam0 pose (cam -> world).
__device__ __forceinline__ void stage_views(const Views& V, int a, const double* q, const double* p, const double* z,
                                            const double* Rf, const double* tf, const TriCfg& cfg) {
  double R0[9];
  to_rotation(q, R0);
  // cam0: pose^-1 * T_c0_w = (R0 Rf, R0 tf + (-(R0 p)))
  double Rr[9], a1[3], a2[3];
  mat3_mul(R0, Rf, Rr);
  mat3_vec(R0, tf, a1);
  mat3_vec(R0, p, a2);
  for (int k = 0; k < 9; ++k) V.at(k, 2 * a) = Rr[k];
  for (int k = 0; k < 3; ++k) V.at(9 + k, 2 * a) = a1[k] + (-a2[k]);
  V.at(12, 2 * a) = z[0]; V.at(13, 2 * a) = z[1];
  // cam1 pose = cam0 pose * (R01^T, -R01^T t01):  R1 = R0^T R01^T, t1 = R0^T (-R01^T t01) + p
  double c10[3], tmp[3], t1[3], R1t[9];
  mat3t_vec(cfg.R01, cfg.t01, tmp);
  for (int k = 0; k < 3; ++k) c10[k] = -tmp[k];
  mat3t_vec(R0, c10, tmp);
  for (int k = 0; k < 3; ++k) t1[k] = tmp[k] + p[k];
  mat3_mul(cfg.R01, R0, R1t);                 // R1^T = R01 R0
  mat3_mul(R1t, Rf, Rr);
  mat3_vec(R1t, tf, a1);
  mat3_vec(R1t, t1, a2);
  for (int k = 0; k < 9; ++k) V.at(k, 2 * a + 1) = Rr[k];
  for (int k = 0; k < 3; ++k) V.at(9 + k, 2 * a + 1) = a1[k] + (-a2[k]);
  V.at(12, 2 * a + 1) = z[2]; V.at(13, 2 * a + 1) = z[3];
}

// Same with R0 = R(q) and R1^T' = R_01 R0 taken from the per-stage camera table (bitwise the products above).
__device__ __forceinline__ void stage_views_pre(const Views& V, int a, const double* pre, const double* p, const double* z,
                                                const double* Rf, const double* tf, const TriCfg& cfg) {
  double R0[9], R1t[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) { R0[k] = pre[k]; R1t[k] = pre[9 + k]; }
  double Rr[9], a1[3], a2[3];
  mat3_mul(R0, Rf, Rr);
  mat3_vec(R0, tf, a1);
  mat3_vec(R0, p, a2);
  for (int k = 0; k < 9; ++k) V.at(k, 2 * a) = Rr[k];
  for (int k = 0; k < 3; ++k) V.at(9 + k, 2 * a) = a1[k] + (-a2[k]);
  V.at(12, 2 * a) = z[0]; V.at(13, 2 * a) = z[1];
  double c10[3], tmp[3], t1[3];
  mat3t_vec(cfg.R01, cfg.t01, tmp);
  for (int k = 0; k < 3; ++k) c10[k] = -tmp[k];
  mat3t_vec(R0, c10, tmp);
  for (int k = 0; k < 3; ++k) t1[k] = tmp[k] + p[k];
  mat3_mul(R1t, Rf, Rr);
  mat3_vec(R1t, tf, a1);
  mat3_vec(R1t, t1, a2);
  for (int k = 0; k < 9; ++k) V.at(k, 2 * a + 1) = Rr[k];
  for (int k = 0; k < 3; ++k) V.at(9 + k, 2 * a + 1) = a1[k] + (-a2[k]);
  V.at(12, 2 * a + 1) = z[2]; V.at(13, 2 * a + 1) = z[3];
}

// check_motion (feature.py:124-165)
__device__ __forceinline__ bool check_motion(const TriCfg& cfg, const double* Rf, const double* tf, const double* z_first,
                                             const double* p_last) {
  if (cfg.thr < 0) return true;
  double dcam[3] = {z_first[0], z_first[1], 1.0};
  const double nn = norm3(dcam);
  for (int i = 0; i < 3; ++i) dcam[i] /= nn;
  double dw[3];
  mat3_vec(Rf, dcam, dw);
  const double tr[3] = {p_last[0] - tf[0], p_last[1] - tf[1], p_last[2] - tf[2]};
  const double par = tr[0] * dw[0] + tr[1] * dw[1] + tr[2] * dw[2];
  const double o[3] = {tr[0] - par * dw[0], tr[1] - par * dw[1], tr[2] - par * dw[2]};
  return norm3(o) > cfg.thr;
}

// ------------------------------------------------------------------ store-driven kernel
// Persistent grid over the work list wl_tri[0 .. n) of (filter, candidate) pairs that still need a position
// (built by k_classify_lost / k_prune_select).  Tracks are short (a handful of stereo observations), so eight lanes
// work on one landmark and four landmarks share a warp.
constexpr int TRI_G = 8;
__global__ void __launch_bounds__(TRI_WPB * 32) k_triangulate_store(Dev d) {
  extern __shared__ double smem[];
  constexpr int GPW = 32 / TRI_G;
  const int warp = threadIdx.x >> 5, wl = threadIdx.x & 31, wpb = blockDim.x >> 5;     // warps per CTA chosen by the launcher
  const int grp = wl / TRI_G, lane = wl - grp * TRI_G, gbase = grp * TRI_G;
  const unsigned gmask = ((1u << TRI_G) - 1u) << gbase;
  const int n = (int)d.stage_ctr[SC_NTRI];
  const TriCfg cfg = load_tricfg(*d.prm);
  const int stride = 2 * d.Nmax;
  Views V{smem + (size_t)(warp * GPW + grp) * (VIEW_DOUBLES * stride + 8), stride};
  for (int w = (blockIdx.x * wpb + warp) * GPW + grp; w < n; w += gridDim.x * wpb * GPW) {
    const int2 it = d.wl_tri[w];
    const int f = it.x, ci = it.y;
    const int slot = d.cand[(size_t)f * d.Fcap + ci];
    const size_t fs = (size_t)f * d.Fcap + slot;
    const unsigned long long mask = d.f_mask[fs];
    const int N = d.ncam[f];
    // logical-order observation mask
    unsigned long long lmask = 0;
    for (int c0 = 0; c0 < N; c0 += TRI_G) {
      const int c = c0 + lane;
      const int bit = (c < N) ? (int)((mask >> d.cam_phys[(size_t)f * d.Nmax + c]) & 1ull) : 0;
      const unsigned b = (__ballot_sync(gmask, bit) >> gbase) & ((1u << TRI_G) - 1u);
      lmask |= (unsigned long long)b << c0;
    }
    const int M = __popcll(lmask);
    const int c_first = __ffsll((long long)lmask) - 1;
    const int c_last = 63 - __clzll((long long)lmask);
    const double* pref = d.cam_pre + ((size_t)f * d.Nmax + c_first) * CAM_PRE;
    const double* caml = d.cam + ((size_t)f * d.Nmax + c_last) * CAM_STRIDE;
    double Rf[9], tf[3];
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) Rf[3 * i + j] = pref[3 * j + i];
    for (int i = 0; i < 3; ++i) tf[i] = pref[18 + i];
    const double* obs_base = d.f_obs + fs * d.Nmax * 4;
    const double* zf = obs_base + (size_t)d.cam_phys[(size_t)f * d.Nmax + c_first] * 4;
    bool ok = check_motion(cfg, Rf, tf, zf, caml + 4);
    double pos[3] = {0, 0, 0};
    __syncwarp(gmask);                  // the previous landmark's views are no longer read
    if (ok) {
      for (int c = lane; c < N; c += TRI_G) {
        if ((lmask >> c) & 1ull) {
          const int a = __popcll(lmask & ((1ull << c) - 1ull));
          const double* pre = d.cam_pre + ((size_t)f * d.Nmax + c) * CAM_PRE;
          stage_views_pre(V, a, pre, pre + 18, obs_base + (size_t)d.cam_phys[(size_t)f * d.Nmax + c] * 4, Rf, tf, cfg);
        }
      }
      __syncwarp(gmask);
      int ns = 0;
      ok = lm_solve<TRI_G>(V, 2 * M, cfg, Rf, tf, lane, gmask, pos, &ns);
      if (lane == 0) {
        d.f_pos[fs * 3 + 0] = pos[0]; d.f_pos[fs * 3 + 1] = pos[1]; d.f_pos[fs * 3 + 2] = pos[2];
        unsigned long long* cnt = (unsigned long long*)(d.counters + (size_t)NCNT * f);
        atomicAdd(cnt + MSCKF_CNT_TRIANGULATIONS, 1ull);
        atomicAdd(cnt + MSCKF_CNT_TRI_OBS, (unsigned long long)M);
        atomicAdd(cnt + MSCKF_CNT_TRI_SOLVES, (unsigned long long)ns);
      }
    }
    if (lane == 0) {
      int fl = d.f_flags[fs];
      fl = ok ? (fl | FF_INIT) : (fl & ~FF_INIT);
      d.f_flags[fs] = fl;
      d.cand_status[(size_t)f * d.Fcap + ci] = ok ? CS_READY : MSCKF_FEAT_TRI_FAILED;
    }
  }
}

// ------------------------------------------------------------------ stateless kernel (explicit poses)
__global__ void __launch_bounds__(TRI_WPB * 32) k_triangulate_raw(Dev d, int n, const int* __restrict__ obs_off,
                                                                   const double* __restrict__ uv, const double* __restrict__ cam_q,
                                                                   const double* __restrict__ cam_p, double* __restrict__ pos_out,
                                                                   unsigned char* __restrict__ valid, int* __restrict__ solves,
                                                                   int max_obs, int mode) {
  extern __shared__ double smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x * TRI_WPB + warp;
  if (i >= n) return;
  const int o0 = obs_off[i], M = obs_off[i + 1] - o0;
  const TriCfg cfg = load_tricfg(*d.prm);
  const int stride = 2 * max_obs;
  Views V{smem + (size_t)warp * VIEW_DOUBLES * stride, stride};
  double R0f[9], Rf[9], tf[3];
  to_rotation(cam_q + 4 * (size_t)o0, R0f);
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) Rf[3 * r + c] = R0f[3 * c + r];
  for (int r = 0; r < 3; ++r) tf[r] = cam_p[3 * (size_t)o0 + r];
  bool ok = check_motion(cfg, Rf, tf, uv + 4 * (size_t)o0, cam_p + 3 * (size_t)(o0 + M - 1));
  double pos[3] = {0, 0, 0};
  int ns = 0;
  if (mode == 1) {                       // Feature.check_motion only
    if (lane == 0) valid[i] = ok ? 1 : 0;
    return;
  }
  if (mode >= 2) ok = true;              // 2: generate_initial_guess only, 3: initialize_position without check_motion
  if (ok) {
    for (int a = lane; a < M; a += 32)
      stage_views(V, a, cam_q + 4 * (size_t)(o0 + a), cam_p + 3 * (size_t)(o0 + a), uv + 4 * (size_t)(o0 + a), Rf, tf, cfg);
    __syncwarp();
    if (mode == 2) {
      initial_guess(V, 2 * M - 2, pos);
      if (lane == 0) { pos_out[3 * (size_t)i + 0] = pos[0]; pos_out[3 * (size_t)i + 1] = pos[1]; pos_out[3 * (size_t)i + 2] = pos[2]; }
      return;
    }
    ok = lm_solve<32>(V, 2 * M, cfg, Rf, tf, lane, 0xffffffffu, pos, &ns);
  }
  if (lane == 0) {
    pos_out[3 * (size_t)i + 0] = pos[0]; pos_out[3 * (size_t)i + 1] = pos[1]; pos_out[3 * (size_t)i + 2] = pos[2];
    valid[i] = ok ? 1 : 0;
    if (solves) solves[i] = ns;
  }
}

void launch_triangulate_store(const Dev& d, cudaStream_t st) {
  int wpb = TRI_WPB;                     // fewer warps per CTA for wide windows so that the views fit shared memory
  while (wpb > 1 && (size_t)wpb * (32 / TRI_G) * (VIEW_DOUBLES * 2 * d.Nmax + 8) * sizeof(double) > 96 * 1024) wpb >>= 1;
  const size_t smem = (size_t)wpb * (32 / TRI_G) * (VIEW_DOUBLES * 2 * d.Nmax + 8) * sizeof(double);
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_triangulate_store, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_triangulate_raw, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(12, (224 * 1024) / (smem + 1024)));     // 3 CTAs / SM at a 20-camera window
  KLAUNCH(d, KID_TRIANGULATE, st, (k_triangulate_store<<<sm_count() * per_sm, wpb * 32, smem, st>>>(d)));
}

void launch_triangulate_raw(const Dev& d, int n, const int* obs_off, const double* uv, const double* cam_q, const double* cam_p,
                            double* pos, unsigned char* valid, int* solves, int max_obs, int mode, cudaStream_t st) {
  const size_t smem = (size_t)TRI_WPB * VIEW_DOUBLES * 2 * max_obs * sizeof(double);
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_triangulate_raw, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  KLAUNCH(d, KID_TRIANGULATE, st,
          (k_triangulate_raw<<<(n + TRI_WPB - 1) / TRI_WPB, TRI_WPB * 32, smem, st>>>(d, n, obs_off, uv, cam_q, cam_p, pos, valid, solves,
                                                                                   max_obs, mode)));
}

}  // namespace msckf
