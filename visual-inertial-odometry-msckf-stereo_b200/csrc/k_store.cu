// Device-resident landmark store: the reference's map_server / observations dictionaries
// (msckf.py:409-427, 616-689, 730-801) as flat per-filter tables.  One CTA per filter.
//
//   * a landmark owns a slot; `ord` lists live slots in insertion order (Python dict order),
//   * observations are a 64-bit mask over physical camera slots plus uv[slot][4],
//   * id -> slot lookup uses an open-addressing table rebuilt every frame,
//   * order-preserving compaction (ballot prefix sums) keeps every list in map order, which is
//     what makes the processed / invalid id sequences and the row-cap break bit-exact.
#include <limits.h>
#include "msckf_dev.cuh"
#include "kernels.h"

namespace msckf {

constexpr int ST = 256;
constexpr long long EMPTY_KEY = LLONG_MIN;

// exclusive prefix of a 0/1 flag over the CTA (one tile of ST threads); total returned to all
__device__ __forceinline__ int block_rank(int flag, int* s_warp, int& total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const unsigned m = __ballot_sync(0xffffffffu, flag);
  const int r = __popc(m & ((1u << lane) - 1u));
  __syncthreads();                     // protects s_warp against the previous tile's readers
  if (lane == 0) s_warp[w] = __popc(m);
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < ST / 32; ++i) {
    const int c = s_warp[i];
    if (i < w) base += c;
    tot += c;
  }
  total = tot;
  return base + r;
}

// exclusive prefix of an int value over the CTA tile
__device__ __forceinline__ long long block_scan_ll(long long v, long long* s_warp, long long& total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  long long inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();
  if (lane == 31) s_warp[w] = inc;
  __syncthreads();
  long long base = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < ST / 32; ++i) {
    const long long c = s_warp[i];
    if (i < w) base += c;
    tot += c;
  }
  total = tot;
  return base + inc - v;
}

__device__ __forceinline__ unsigned hash64(long long k) {
  unsigned long long x = (unsigned long long)k;
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return (unsigned)x;
}

// ------------------------------------------------------------------ add_feature_observations (msckf.py:409-427)
__global__ void __launch_bounds__(ST) k_observe(Dev d, const int* __restrict__ feat_off, const long long* __restrict__ ids,
                                                 const double* __restrict__ uv, const FrameIn* fin, int use_smem, int pending_augment) {
  if (fin) { feat_off = fin->feat_off; ids = fin->ids; uv = fin->uv; }
  const int f = blockIdx.x, tid = threadIdx.x;
  __shared__ int s_warp[ST / 32];
  __shared__ int s_tracked;
  const int n0 = d.nlive[f];
  const int off = feat_off[f], cnt = feat_off[f + 1] - off;
  // inside a frame this kernel runs beside propagation / augmentation: the camera state being added is not in the window
  // yet, its index and observation slot come from k_cam_slot
  const int N = pending_augment ? d.obs_N[f] : d.ncam[f];
  // the per-frame id -> slot table lives in shared memory when it fits (launch_observe), else in its global scratch
  extern __shared__ __align__(16) unsigned char obs_sm[];
  long long* hk = use_smem ? (long long*)obs_sm : d.h_key + (size_t)f * d.Hs;
  int* hv = use_smem ? (int*)(obs_sm + (size_t)d.Hs * sizeof(long long)) : d.h_val + (size_t)f * d.Hs;
  const unsigned hmask = (unsigned)d.Hs - 1u;
  if (tid == 0) s_tracked = 0;
  for (int e = tid; e < d.Hs; e += ST) hk[e] = EMPTY_KEY;
  __syncthreads();
  if (N == 0) return;
  const int phys = pending_augment ? d.obs_phys[f] : d.cam_phys[(size_t)f * d.Nmax + N - 1];
  if (phys < 0) return;                              // window full (reported)
  const unsigned long long bit = 1ull << phys;
  int* ord = d.ord + (size_t)f * d.Fcap;
  for (int j = tid; j < n0; j += ST) {
    const int slot = ord[j];
    const long long key = d.f_id[(size_t)f * d.Fcap + slot];
    unsigned h = hash64(key) & hmask;
    while (true) {
      const unsigned long long old = atomicCAS((unsigned long long*)&hk[h], (unsigned long long)EMPTY_KEY, (unsigned long long)key);
      if (old == (unsigned long long)EMPTY_KEY) { hv[h] = slot; break; }
      h = (h + 1) & hmask;
    }
  }
  __syncthreads();
  int nfree = d.nfree[f];
  const int* fstack = d.free_stack + (size_t)f * d.Fcap;
  int base_new = 0;
  for (int t0 = 0; t0 < cnt; t0 += ST) {
    const int i = t0 + tid;
    const bool valid = i < cnt;
    long long key = 0;
    int slot = -1;
    if (valid) {
      key = ids[off + i];
      unsigned h = hash64(key) & hmask;
      while (true) {
        const long long k = hk[h];
        if (k == EMPTY_KEY) break;
        if (k == key) { slot = hv[h]; break; }     // (-2: the id was new earlier in this message -> reported below)
        h = (h + 1) & hmask;
      }
    }
    if (slot == -2) { atomicOr(d.err, ERRBIT_DUP_ID); slot = -1; }
    const int isnew = valid && slot < 0;
    int total;
    const int rank = block_rank(isnew, s_warp, total);
    if (isnew) {
      const int idx = base_new + rank;
      if (idx >= nfree) {
        atomicOr(d.err, ERRBIT_FEATURE_CAP);
      } else {
        slot = fstack[nfree - 1 - idx];
        const size_t fs = (size_t)f * d.Fcap + slot;
        d.f_id[fs] = key;
        d.f_mask[fs] = bit;
        d.f_flags[fs] = FF_LIVE;
        d.f_pos[fs * 3 + 0] = 0; d.f_pos[fs * 3 + 1] = 0; d.f_pos[fs * 3 + 2] = 0;
        ord[n0 + idx] = slot;
      }
      // the new id joins the frame's table (value -2: "new in this message"): a second occurrence of the same id, which
      // also found no live landmark, meets it here and is reported (the message contract requires unique ids)
      unsigned h = hash64(key) & hmask;
      while (true) {
        const unsigned long long old = atomicCAS((unsigned long long*)&hk[h], (unsigned long long)EMPTY_KEY, (unsigned long long)key);
        if (old == (unsigned long long)EMPTY_KEY) { hv[h] = -2; break; }
        if (old == (unsigned long long)key) { atomicOr(d.err, ERRBIT_DUP_ID); break; }
        h = (h + 1) & hmask;
      }
    } else if (valid) {
      d.f_mask[(size_t)f * d.Fcap + slot] |= bit;
      atomicAdd(&s_tracked, 1);
    }
    if (valid && slot >= 0) {
      double* o = d.f_obs + (((size_t)f * d.Fcap + slot) * d.Nmax + phys) * 4;
      const double* z = uv + 4 * (size_t)(off + i);
      o[0] = z[0]; o[1] = z[1]; o[2] = z[2]; o[3] = z[3];
    }
    base_new += total;
    __syncthreads();                                  // this tile's insertions are complete before the next tile looks ids up
  }
  __syncthreads();
  if (tid == 0) {
    const int added = min(base_new, nfree);
    d.nlive[f] = n0 + added;
    d.nfree[f] = nfree - added;
    d.imu[(size_t)f * IMU_STRIDE + MSCKF_IMU_TRACKING_RATE] = (double)s_tracked / ((double)n0 + 1e-5);
  }
}

// Appends (filter, candidate) to the triangulation work list: one atomic per warp, not per landmark.
__device__ __forceinline__ void wl_push_tri(const Dev& d, bool want, int f, int idx) {
  const unsigned bal = __ballot_sync(0xffffffffu, want);
  if (!bal) return;
  const unsigned lane = threadIdx.x & 31;
  const int leader = __ffs(bal) - 1;
  unsigned long long base = 0;
  if ((int)lane == leader) base = atomicAdd(d.stage_ctr + SC_NTRI, (unsigned long long)__popc(bal));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (want) d.wl_tri[base + __popc(bal & ((1u << lane) - 1u))] = make_int2(f, idx);
}

// ------------------------------------------------------------------ remove_lost_features, pass 1 (msckf.py:623-650)
// Splits the map into tracked (kept) and lost landmarks, both in map order.  Lost landmarks go to
// the candidate list with status FEW_OBS / CS_NEED_TRI / CS_READY; their slots are released
// (the data stays valid until the next frame's observations arrive).
__global__ void __launch_bounds__(ST) k_classify_lost(Dev d) {
  const int f = blockIdx.x, tid = threadIdx.x;
  __shared__ int s_warp[ST / 32];
  const int n = d.nlive[f], N = d.ncam[f];
  int* ord = d.ord + (size_t)f * d.Fcap;
  int* tmp = d.ord_tmp + (size_t)f * d.Fcap;
  int* cand = d.cand + (size_t)f * d.Fcap;
  int* cst = d.cand_status + (size_t)f * d.Fcap;
  if (N == 0) {
    if (tid == 0) { d.ncand[f] = 0; d.nitem[f] = 0; }
    return;
  }
  const unsigned long long bit = 1ull << d.cam_phys[(size_t)f * d.Nmax + N - 1];
  int nkeep = 0, nlost = 0;
  const int nfree0 = d.nfree[f];
  for (int t0 = 0; t0 < n; t0 += ST) {
    const int j = t0 + tid;
    const bool valid = j < n;
    int slot = 0;
    unsigned long long mask = 0;
    if (valid) { slot = ord[j]; mask = d.f_mask[(size_t)f * d.Fcap + slot]; }
    const int lost = valid && !(mask & bit);
    const int keep = valid && !lost;
    int tl, tk;
    const int rl = block_rank(lost, s_warp, tl);
    const int rk = block_rank(keep, s_warp, tk);
    if (keep) tmp[nkeep + rk] = slot;
    bool want_tri = false;
    int tri_idx = 0;
    if (lost) {
      const int ci = nlost + rl;
      cand[ci] = slot;
      int st;
      if (__popcll(mask) < 3) st = MSCKF_FEAT_FEW_OBS;
      else st = (d.f_flags[(size_t)f * d.Fcap + slot] & FF_INIT) ? CS_READY : CS_NEED_TRI;
      cst[ci] = st;
      want_tri = st == CS_NEED_TRI;
      tri_idx = ci;
      d.free_stack[(size_t)f * d.Fcap + nfree0 + ci] = slot;
    }
    wl_push_tri(d, want_tri, f, tri_idx);
    nkeep += tk;
    nlost += tl;
  }
  __syncthreads();
  for (int j = tid; j < nkeep; j += ST) ord[j] = tmp[j];
  if (tid == 0) {
    d.nlive[f] = nkeep;
    d.nfree[f] = nfree0 + nlost;
    d.ncand[f] = nlost;
    d.rm_mask[f] = 0ull;
  }
}

// ------------------------------------------------------------------ find_redundant_cam_states (msckf.py:691-727)
__device__ void find_redundant(const Dev& d, int f, int* rm /*[2] logical, sorted*/) {
  const Params& prm = *d.prm;
  const int N = d.ncam[f];
  const double* cam = d.cam + (size_t)f * d.Nmax * CAM_STRIDE;
  const double track = d.imu[(size_t)f * IMU_STRIDE + MSCKF_IMU_TRACKING_RATE];
  const int key = N - 4;
  int idx = key + 1, first = 0;
  double Rk[9];
  to_rotation(cam + (size_t)key * CAM_STRIDE, Rk);
  const double* pk = cam + (size_t)key * CAM_STRIDE + 4;
  for (int i = 0; i < 2; ++i) {
    const double* c = cam + (size_t)idx * CAM_STRIDE;
    double R[9], B[9], q[4];
    to_rotation(c, R);
    const double dp[3] = {c[4] - pk[0], c[5] - pk[1], c[6] - pk[2]};
    const double distance = norm3(dp);
    mat3_mul_nt(R, Rk, B);
    to_quaternion(B, q);
    const double angle = 2 * acos(q[3]);
    if (angle < prm.prune_rot_thr && distance < prm.prune_trans_thr && track > prm.prune_track_thr) {
      rm[i] = idx;
      idx += 1;
    } else {
      rm[i] = first;
      first += 1;
      idx += 1;
    }
  }
  if (rm[0] > rm[1]) { const int t = rm[0]; rm[0] = rm[1]; rm[1] = t; }
}

__global__ void k_find_redundant(Dev d, int f, long long* out) {
  if (threadIdx.x == 0) {
    int rm[2] = {-1, -1};
    if (d.ncam[f] >= 4) find_redundant(d, f, rm);
    out[0] = rm[0] >= 0 ? d.cam_id[(size_t)f * d.Nmax + rm[0]] : -1;
    out[1] = rm[1] >= 0 ? d.cam_id[(size_t)f * d.Nmax + rm[1]] : -1;
  }
}

// ------------------------------------------------------------------ prune_cam_state_buffer, pass 1 (msckf.py:730-769)
__global__ void __launch_bounds__(ST) k_prune_select(Dev d) {
  const int f = blockIdx.x, tid = threadIdx.x;
  __shared__ int s_warp[ST / 32];
  __shared__ unsigned long long s_rm;
  const int N = d.ncam[f];
  if (N < d.prm->max_cam) {
    if (tid == 0) {
      d.ncand[f] = 0; d.nitem[f] = 0; d.rm_idx[2 * f] = -1; d.rm_idx[2 * f + 1] = -1; d.rm_mask[f] = 0ull;
      d.log_rm[2 * f] = -1; d.log_rm[2 * f + 1] = -1;
    }
    return;
  }
  if (tid == 0) {
    int rm[2];
    find_redundant(d, f, rm);
    d.rm_idx[2 * f] = rm[0]; d.rm_idx[2 * f + 1] = rm[1];
    const unsigned long long m = (1ull << d.cam_phys[(size_t)f * d.Nmax + rm[0]]) | (1ull << d.cam_phys[(size_t)f * d.Nmax + rm[1]]);
    d.rm_mask[f] = m;
    s_rm = m;
    d.log_rm[2 * f] = d.cam_id[(size_t)f * d.Nmax + rm[0]];
    d.log_rm[2 * f + 1] = d.cam_id[(size_t)f * d.Nmax + rm[1]];
  }
  __syncthreads();
  const unsigned long long rmm = s_rm;
  const int n = d.nlive[f];
  const int* ord = d.ord + (size_t)f * d.Fcap;
  int* cand = d.cand + (size_t)f * d.Fcap;
  int* cst = d.cand_status + (size_t)f * d.Fcap;
  int nc = 0;
  for (int t0 = 0; t0 < n; t0 += ST) {
    const int j = t0 + tid;
    const bool valid = j < n;
    int slot = 0, c = 0;
    unsigned long long mask = 0;
    if (valid) {
      slot = ord[j];
      mask = d.f_mask[(size_t)f * d.Fcap + slot];
      c = __popcll(mask & rmm);
      if (c == 1) d.f_mask[(size_t)f * d.Fcap + slot] = mask & ~rmm;
    }
    int tot;
    const int r = block_rank(c >= 2, s_warp, tot);
    bool want_tri = false;
    if (c >= 2) {
      cand[nc + r] = slot;
      const int st = (d.f_flags[(size_t)f * d.Fcap + slot] & FF_INIT) ? CS_READY : CS_NEED_TRI;
      cst[nc + r] = st;
      want_tri = st == CS_NEED_TRI;
    }
    wl_push_tri(d, want_tri, f, nc + r);
    nc += tot;
  }
  if (tid == 0) d.ncand[f] = nc;
}

// ------------------------------------------------------------------ items: landmarks that reach feature_jacobian
// stage 0: all observations, dof = M-1 (msckf.py:645,669-671); stage 1: the pruned camera
// states only, dof = M (msckf.py:769,787-789); their observations are then dropped (:794-795).
// Scratch for each item is bump-allocated from the pool.
__global__ void __launch_bounds__(ST) k_plan_items(Dev d, int stage) {
  const int f = blockIdx.x, tid = threadIdx.x;
  __shared__ int s_warp[ST / 32];
  __shared__ long long s_warp_ll[ST / 32];
  __shared__ unsigned long long s_base;
  const int nc = d.ncand[f];
  const int* cand = d.cand + (size_t)f * d.Fcap;
  const int* cst = d.cand_status + (size_t)f * d.Fcap;
  const unsigned long long rmm = d.rm_mask[f];
  __shared__ int s_phys[64];                 // physical observation slot of logical camera c
  const int N = d.ncam[f];
  for (int c = tid; c < N; c += ST) s_phys[c] = d.cam_phys[(size_t)f * d.Nmax + c];
  __syncthreads();
  int ni = 0;
  long long elems_total = 0;
  long long w_obs = 0, w_flops = 0, w_bytes = 0;   // algorithmic work of the Jacobian / gate kernel
  // pass A: ordered item list + relative offsets
  for (int t0 = 0; t0 < nc; t0 += ST) {
    const int ci = t0 + tid;
    const bool valid = ci < nc;
    int isitem = 0, M = 0;
    unsigned long long im = 0;
    if (valid) {
      const int slot = cand[ci];
      const unsigned long long mask = d.f_mask[(size_t)f * d.Fcap + slot];
      isitem = cst[ci] == CS_READY;
      im = stage == 0 ? mask : (mask & rmm);
      M = __popcll(im);
      if (stage == 1) d.f_mask[(size_t)f * d.Fcap + slot] = mask & ~rmm;
    }
    const int rows = 4 * M - 3;
    const long long need = isitem ? jac_record_doubles(M) : 0;  // projected rows, see k_jacobian.cu
    int tot;
    const int r = block_rank(isitem, s_warp, tot);
    long long tot_e;
    const long long eoff = block_scan_ll(need, s_warp_ll, tot_e);
    if (isitem) {
      const size_t k = (size_t)f * d.Fcap + ni + r;
      d.item_cand[k] = ci;
      d.item_mask[k] = im;
      unsigned long long lm = 0;             // the mask in logical camera order: what the Jacobian and QR kernels index by
      for (int c = 0; c < N; ++c) lm |= ((im >> s_phys[c]) & 1ull) << c;
      d.item_lmask[k] = lm;
      d.item_slot[k] = cand[ci];
      d.item_rows[k] = rows;
      d.item_dof[k] = stage == 0 ? M - 1 : M;
      d.item_off[k] = elems_total + eoff;   // relative; rebased below
      d.item_pass[k] = 0;
      d.item_gamma[k] = 0.0;
      d.item_rowoff[k] = -1;
      const long long Ml = M, R = rows, C6 = 6 * Ml;
      w_obs += Ml;
      w_flops += 300 * Ml + 12 * (4 * Ml) * (C6 + 1) + 2 * R * C6 * C6 + 2 * R * R * C6 + (R * R * R) / 3 + 2 * R * R;
      w_bytes += 36 * Ml + 8 * R * (C6 + 1);
    }
    // queue the item for its Jacobian kernel class: one atomic per warp and class instead of one per item
    {
      const int cls = !isitem ? -1 : (M <= JAC_SMALL_M ? jac_class(M) : (M <= JAC_BIG_SPLIT_M ? JAC_CLASSES : JAC_CLASSES + 1));
      const unsigned lane = tid & 31;
#pragma unroll
      for (int c = 0; c <= JAC_CLASSES + 1; ++c) {
        const unsigned bal = __ballot_sync(0xffffffffu, cls == c);
        if (bal == 0) continue;
        unsigned long long base = 0;
        if (lane == (unsigned)(__ffs(bal) - 1))
          base = atomicAdd(d.stage_ctr + (c < JAC_CLASSES ? SC_NJAC_SMALL + c : (c == JAC_CLASSES ? SC_NJAC_BIG : SC_NJAC_BIG2)),
                           (unsigned long long)__popc(bal));
        base = __shfl_sync(0xffffffffu, base, __ffs(bal) - 1);
        if (cls == c) {
          const unsigned long long pos = base + __popc(bal & ((1u << lane) - 1u));
          if (c < JAC_CLASSES) d.wl_jac_small[(size_t)c * d.F * d.Fcap + pos] = make_int2(f, ni + r);
          else if (c == JAC_CLASSES) d.wl_jac_big[pos] = make_int2(f, ni + r);
          else d.wl_jac_big2[pos] = make_int2(f, ni + r);
        }
      }
    }
    ni += tot;
    elems_total += tot_e;
  }
  __syncthreads();
  if (tid == 0) {
    unsigned long long base = 0;
    if (elems_total > 0) {
      base = atomicAdd(d.stage_ctr + SC_POOL, (unsigned long long)elems_total);
      if ((long long)(base + elems_total) > d.pool_cap) {
        atomicOr(d.err, ERRBIT_POOL);
        ni = 0;
      }
    }
    s_base = base;
    d.nitem[f] = ni;
  }
  __syncthreads();
  const long long base = (long long)s_base;
  for (int k = tid; k < ni; k += ST) d.item_off[(size_t)f * d.Fcap + k] += base;
  if (w_obs && ni > 0) {
    unsigned long long* cnt = (unsigned long long*)(d.counters + (size_t)NCNT * f);
    atomicAdd(cnt + MSCKF_CNT_JAC_OBS, (unsigned long long)w_obs);
    atomicAdd(cnt + MSCKF_CNT_K2_FLOPS, (unsigned long long)w_flops);
    atomicAdd(cnt + MSCKF_CNT_K2_BYTES, (unsigned long long)w_bytes);
  }
}

// ------------------------------------------------------------------ stacking plan + decision log
// Applies the sequential semantics of msckf.py:662-679 / 776-795 to the gate results: a landmark
// is taken through the gate only while the rows stacked so far are <= row_cap (stage 0).
__global__ void __launch_bounds__(ST) k_plan_stack(Dev d, int stage) {
  const int f = blockIdx.x, tid = threadIdx.x;
  __shared__ long long s_warp_ll[ST / 32];
  __shared__ unsigned long long s_act;
  __shared__ int s_proc;
  const int ni = d.nitem[f], nc = d.ncand[f];
  const int cap = (stage == 0) ? d.prm->row_cap : 0;
  if (tid == 0) { s_act = 0ull; s_proc = 0; }
  __syncthreads();
  long long stacked = 0;   // rows stacked by earlier tiles (inclusive prefix)
  long long used_rows = 0;
  int* cst = d.cand_status + (size_t)f * d.Fcap;
  for (int t0 = 0; t0 < ni; t0 += ST) {
    const int k = t0 + tid;
    const bool valid = k < ni;
    const size_t ik = (size_t)f * d.Fcap + k;
    const int pass = valid ? d.item_pass[ik] : 0;
    const int rows = valid ? d.item_rows[ik] : 0;
    long long tot;
    const long long excl = stacked + block_scan_ll(pass ? rows : 0, s_warp_ll, tot);
    const bool processed = valid && (cap <= 0 || excl <= cap);
    const bool used = processed && pass;
    long long tot_u;
    const long long ro = used_rows + block_scan_ll(used ? rows : 0, s_warp_ll, tot_u);
    if (valid) {
      d.item_rowoff[ik] = used ? (int)ro : -1;
      if (used && ro + rows <= d.rowmap_cap) {
        int* rm = d.rowmap + (size_t)f * d.rowmap_cap + ro;
        for (int li = 0; li < rows; ++li) rm[li] = (k << 8) | li;
      }
      cst[d.item_cand[ik]] = processed ? (pass ? MSCKF_FEAT_GATE_PASS : MSCKF_FEAT_GATE_FAIL) : MSCKF_FEAT_AFTER_CAP;
      if (used) atomicOr(&s_act, d.item_mask[ik]);
      if (processed) atomicAdd(&s_proc, 1);
    }
    stacked += tot;
    used_rows += tot_u;
  }
  __syncthreads();
  // decision log in map order
  long long* lid = d.log_id + ((size_t)stage * d.F + f) * d.Fcap;
  int* lst = d.log_status + ((size_t)stage * d.F + f) * d.Fcap;
  double* lga = d.log_gamma + ((size_t)stage * d.F + f) * d.Fcap;
  int* ldf = d.log_dof + ((size_t)stage * d.F + f) * d.Fcap;
  for (int ci = tid; ci < nc; ci += ST) {
    lid[ci] = d.f_id[(size_t)f * d.Fcap + d.cand[(size_t)f * d.Fcap + ci]];
    lst[ci] = cst[ci];
    lga[ci] = 0.0;
    ldf[ci] = 0;
  }
  __syncthreads();
  for (int k = tid; k < ni; k += ST) {
    const size_t ik = (size_t)f * d.Fcap + k;
    lga[d.item_cand[ik]] = d.item_gamma[ik];
    ldf[d.item_cand[ik]] = d.item_dof[ik];
  }
  // active columns in logical order: warp 0, one camera per lane (a serial walk is a chain of dependent global loads)
  __shared__ int s_na;
  if (tid < 32) {
    const int N = d.ncam[f];
    const unsigned long long act = s_act;
    int* cm = d.colmap + (size_t)f * d.nmax;
    int nblk = 0;
    for (int c0 = 0; c0 < N; c0 += 32) {
      const int c = c0 + tid;
      const int bit = (c < N) ? (int)((act >> d.cam_phys[(size_t)f * d.Nmax + c]) & 1ull) : 0;
      const unsigned b = __ballot_sync(0xffffffffu, bit);
      if (bit) {
        const int o = 6 * (nblk + __popc(b & ((1u << tid) - 1u)));
        for (int e = 0; e < 6; ++e) cm[o + e] = NI + 6 * c + e;
      }
      nblk += __popc(b);
    }
    if (tid == 0) s_na = 6 * nblk;
    __syncwarp();
  }
  if (tid == 0) {
    int m = (int)used_rows;
    const int N = d.ncam[f];
    const int na = s_na;
    if (m > d.rowmap_cap) { atomicOr(d.err, ERRBIT_ROWS); m = 0; }
    const int kr = (m > na) ? na : m;
    int qcls = -1;
    if (m > 0) {
      // not compressed, wider than the one-CTA classes, or (when the batch leaves SMs idle) tall enough to be worth a tree: k_qr_build / k_qr_wide
      qcls = (m <= na || na + 1 > 80 || (d.qr_leaf_cap > 1 && m >= 2048)) ? QR_DMMA_CLASSES : (na + 1 <= 16 ? 0 : (na + 1 <= 32 ? 1 : (na + 1 <= 48 ? 2 : (na + 1 <= 64 ? 3 : 4))));
      // pruning stage: every landmark sits on the two cameras being removed (5 rows x 12 columns each): tall-skinny TSQR
      if (stage == 1 && qcls == 0 && na == 12 && d.qr_tall_ok) {
        qcls = QR_DMMA_CLASSES + 1;
        d.wl_qrt[atomicAdd(d.stage_ctr + SC_NQR_TALL, 1ull)] = f;
      } else if (qcls < QR_DMMA_CLASSES) d.wl_qr[(size_t)qcls * d.F + atomicAdd(d.stage_ctr + SC_NQR + qcls, 1ull)] = f;
      else d.wl_qrb[atomicAdd(d.stage_ctr + SC_NQR_BUILD, 1ull)] = f;
    }
    d.qr_cls[f] = qcls;
    d.m_rows[f] = m;
    d.m_pad[f] = max(4, ((kr + 3) / 4) * 4);
    d.n_act[f] = na;
    d.k_rows[f] = kr;
    int path = 0;
    if (kr > 0) {
      const int wid = max(kr, na) + 1;                  // same thresholds as the compression classes: QR and update of a class share a stream
      const int uc = wid <= 16 ? 0 : (wid <= 32 ? 1 : (wid <= 48 ? 2 : (wid <= 64 ? 3 : (wid <= 80 ? 4 : -1))));
      path = (uc >= 0 && d.upd_dmma_ok[uc]) ? 5 + uc : 2;           // one-CTA tensor-pipe kernel, else tiled
      if (path == 2) d.wl_upd_big[atomicAdd(d.stage_ctr + SC_NUPD_BIG, 1ull)] = f;
      if (path >= 5) d.wl_upd_dmma[(size_t)(path - 5) * d.F + atomicAdd(d.stage_ctr + SC_NUPD_DMMA + (path - 5), 1ull)] = f;
    }
    d.upd_path[f] = path;
    d.log_n[stage * d.F + f] = nc;
    d.log_upd[((size_t)stage * d.F + f) * 2 + 0] = m;
    d.log_upd[((size_t)stage * d.F + f) * 2 + 1] = NI + 6 * N;
    long long* cnt = d.counters + (size_t)NCNT * f;
    cnt[MSCKF_CNT_FEATURE_UPDATES] += s_proc;
    if (m > 0) {
      const long long mm = m, a = na, k = d.k_rows[f], n = NI + 6 * N;
      cnt[MSCKF_CNT_UPDATES] += 1;
      if (mm > a) {            // 2 m nc^2 - (2/3) nc^3 + 4 m nc ; bytes 8 m (nc+1) + 8 nc^2
        cnt[MSCKF_CNT_QR_FLOPS] += 2 * mm * a * a - (2 * a * a * a) / 3 + 4 * mm * a;
        cnt[MSCKF_CNT_QR_BYTES] += 8 * mm * (a + 1) + 8 * a * a;
      }
      cnt[MSCKF_CNT_K4_FLOPS] += 4 * k * n * n + 3 * k * k * n + (k * k * k) / 3;   // K4: 4kn^2 + 3k^2n + k^3/3
      cnt[MSCKF_CNT_K4_BYTES] += 16 * n * n + 8 * k * n;
      // what the kernels execute: T = H_thin P[act,:] over the active columns only (half of it when H_thin = R is
      // triangular), the upper triangle of S, Cholesky, W = U^-T [T | r], dx, the upper triangle of the downdate
      const bool tri = mm > a;
      cnt[MSCKF_CNT_K4_EXEC_FLOPS] += (tri ? 1 : 2) * k * a * (n + 1) + (tri ? (k * k * a) / 3 : k * k * a) + (k * k * k) / 3 +
                                      k * k * (n + 1) + k * n * n + 2 * k * n;
    }
  }
}

void launch_observe(const Dev& d, const int* feat_off, const long long* ids, const double* uv, cudaStream_t st, const FrameIn* fin,
                    int pending_augment) {
  const size_t table = (size_t)d.Hs * (sizeof(long long) + sizeof(int));
  const int use_smem = table <= 96 * 1024;
  static DevOnce once;     // function attributes are per device
  once.run([&] {
    cudaFuncSetAttribute(k_observe, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
  });
  KLAUNCH(d, KID_OBSERVE, st, (k_observe<<<d.F, ST, use_smem ? table : 0, st>>>(d, feat_off, ids, uv, fin, use_smem, pending_augment)));
}
void launch_classify_lost(const Dev& d, cudaStream_t st) {
  KLAUNCH(d, KID_CLASSIFY_LOST, st, (k_classify_lost<<<d.F, ST, 0, st>>>(d)));
}
void launch_plan_items(const Dev& d, int stage, cudaStream_t st) {
  KLAUNCH(d, KID_PLAN_ITEMS, st, (k_plan_items<<<d.F, ST, 0, st>>>(d, stage)));
}
void launch_plan_stack(const Dev& d, int stage, cudaStream_t st) {
  KLAUNCH(d, KID_PLAN_STACK, st, (k_plan_stack<<<d.F, ST, 0, st>>>(d, stage)));
}
void launch_prune_select(const Dev& d, cudaStream_t st) {
  KLAUNCH(d, KID_PRUNE_SELECT, st, (k_prune_select<<<d.F, ST, 0, st>>>(d)));
}
void launch_find_redundant(const Dev& d, int filter, long long* out_ids, cudaStream_t st) {
  KLAUNCH(d, KID_MISC, st, (k_find_redundant<<<1, 32, 0, st>>>(d, filter, out_ids)));
}

}  // namespace msckf
