// C ABI of libmsckf_b200.so (see include/msckf_b200.h): context life cycle, pinned staging,
// stage sequencing on one CUDA stream, state import / export.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "kernels.h"

using namespace msckf;

namespace msckf {
const char* const kKernelNames[KID_COUNT] = {
    "k_propagate", "k_augment", "k_cam_prepare", "k_observe", "k_classify_lost", "k_prune_select", "k_triangulate", "k_plan_items",
    "k_jacobian_gate", "k_jacobian_rows", "k_jacobian_big", "k_plan_stack", "k_qr", "k_qr_wide", "k_qr_tall", "k_qr_d16", "k_qr_d32", "k_qr_d48", "k_qr_d64", "k_qr_d80", "k_upd_dmma16", "k_upd_dmma32", "k_upd_dmma48", "k_upd_dmma64", "k_upd_dmma80", "k_upd_T", "k_upd_S", "k_upd_chol", "k_upd_trsm", "k_upd_inject",
    "k_upd_downdate", "k_remove_cams", "k_finish_frame", "k_reset", "misc"};

void KTimer::begin(int id, cudaStream_t st) {
  launches[id] += 1;
  if (!timing) return;
  if (pool.size() < 2 * (pending.size() + 1)) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    pool.push_back(a);
    pool.push_back(b);
  }
  cudaEventRecord(pool[2 * pending.size()], st);
  pending.push_back(id);
}
void KTimer::end(cudaStream_t st) {
  if (!timing) return;
  cudaEventRecord(pool[2 * pending.size() - 1], st);
  if (pending.size() >= 8192) flush(st);
}
void KTimer::flush(cudaStream_t st) {
  if (pending.empty()) return;
  cudaStreamSynchronize(st);
  for (size_t i = 0; i < pending.size(); ++i) {
    float t = 0;
    if (cudaEventElapsedTime(&t, pool[2 * i], pool[2 * i + 1]) == cudaSuccess) {
      ms[pending[i]] += t;
      timed[pending[i]] += 1;
    }
  }
  pending.clear();
}
void KTimer::release() {
  for (cudaEvent_t e : pool) cudaEventDestroy(e);
  pool.clear();
  pending.clear();
}
}  // namespace msckf

namespace {
thread_local std::string g_create_error;

struct Staging {          // one pinned + one device arena; reused round-robin
  char* h = nullptr;
  char* d = nullptr;
  size_t cap = 0;
  cudaEvent_t done = nullptr;
};
}  // namespace

struct msckf_ctx {
  msckf_config cfg;
  Dev d;
  Params hp;
  int device = 0;
  cudaStream_t st = nullptr;
  std::vector<void*> dev_allocs;
  Staging stg[2];
  int stg_next = 0;
  double* h_pub = nullptr;          // pinned [F][IMU_STRIDE]
  unsigned char* h_reset = nullptr; // pinned [F]
  unsigned* h_err = nullptr;        // pinned [8]
  std::string err;
  bool timing = false;
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  cudaEvent_t mark[2] = {nullptr, nullptr};
  float last_ms[5] = {0, 0, 0, 0, 0};
  // frames enqueued by msckf_feature_callback_async and not collected yet (at most two): results land in per-slot pinned buffers
  struct FrameSlot { double* pub = nullptr; unsigned char* reset = nullptr; unsigned* err = nullptr; cudaEvent_t done = nullptr; } slot[2];
  int n_pending = 0, head = 0;
  cudaStream_t copy_st = nullptr;    // H2D of the next frame's inputs overlaps the kernels of the current one
  // the kernel sequence of a whole frame, captured once (forks onto the auxiliary streams included) and replayed
  cudaGraphExec_t frame_graph = nullptr;
  bool graph_enabled = true;
  int eager_frames = 0;              // frames run launch by launch before the capture (function attributes get set there)
  long long graph_launches[KID_COUNT] = {};
  cudaEvent_t h2d_done = nullptr;
  KTimer kt;                        // launch counts / optional per-kernel event timing
  Lanes lanes;                      // auxiliary streams for the independent launches of one step
};

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      char buf_[512];                                                                              \
      snprintf(buf_, sizeof buf_, "%s failed: %s (%s:%d)%s%s", #call, cudaGetErrorString(e_), __FILE__, __LINE__, \
               ctx->kt.bad_launch >= 0 ? " first rejected launch: " : "", ctx->kt.bad_launch >= 0 ? kKernelNames[ctx->kt.bad_launch] : ""); \
      ctx->err = buf_;                                                                             \
      return MSCKF_ERR_CUDA;                                                                       \
    }                                                                                              \
  } while (0)

#define ARG(cond, msg)                  \
  do {                                  \
    if (!(cond)) {                      \
      ctx->err = std::string("bad argument: ") + msg; \
      return MSCKF_ERR_ARG;             \
    }                                   \
  } while (0)

template <class T>
static int dalloc(msckf_ctx* ctx, T** p, size_t count) {
  void* q = nullptr;
  CK(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
  CK(cudaMemsetAsync(q, 0, std::max<size_t>(count, 1) * sizeof(T), ctx->st));
  ctx->dev_allocs.push_back(q);
  *p = (T*)q;
  return 0;
}

static int check_device_errors(msckf_ctx* ctx, unsigned* h_err = nullptr) {
  // caller has synchronised the stream after copying d.err to h_err
  if (!h_err) h_err = ctx->h_err;
  const unsigned e = *h_err;
  if (!e) return MSCKF_OK;
  cudaMemsetAsync(ctx->d.err, 0, 8 * sizeof(unsigned), ctx->st);
  const unsigned* info = h_err;
  char detail[160];
  snprintf(detail, sizeof detail, " [first at kernel class %u, filter %u, sizes %u %u %u]", info[1], info[2], info[3], info[4], info[5]);
  const std::string det = detail;
  *h_err = 0;
  std::string s = "device reported:";
  int rc = MSCKF_ERR_CAPACITY;
  if (e & ERRBIT_FEATURE_CAP) s += " feature_capacity exceeded;";
  if (e & ERRBIT_POOL) s += " jacobian pool exhausted (raise jacobian_pool_bytes);";
  if (e & ERRBIT_ROWS) s += " stacked-Jacobian rows exceed max_rows;";
  if (e & ERRBIT_CAM_CAP) s += " camera-state window full (augment without prune);";
  if (e & ERRBIT_DUP_ID) { s += " duplicate feature id inside one message (ids must be unique, msckf_add_observations);"; rc = MSCKF_ERR_ARG; }
  if (e & ERRBIT_NOT_SPD) { s += " innovation covariance not positive definite;" + det; rc = MSCKF_ERR_NUMERIC; }
  ctx->err = s;
  return rc;
}

static int sync_and_check(msckf_ctx* ctx) {
  CK(cudaMemcpyAsync(ctx->h_err, ctx->d.err, 8 * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->st));
  CK(cudaStreamSynchronize(ctx->st));
  CK(cudaGetLastError());
  return check_device_errors(ctx);
}

// ---- staging: pack several host arrays into one pinned arena, one H2D copy
struct Packer {
  msckf_ctx* ctx;
  Staging* s;
  size_t off = 0;
  std::vector<std::pair<size_t, std::pair<const void*, size_t>>> parts;
  size_t add(const void* src, size_t bytes) {
    const size_t o = off;
    parts.push_back({o, {src, bytes}});
    off += (bytes + 255) & ~size_t(255);
    return o;
  }
};

static int stage_begin(msckf_ctx* ctx, Packer& p) {
  p.ctx = ctx;
  p.s = &ctx->stg[ctx->stg_next];
  ctx->stg_next ^= 1;
  return 0;
}

// Caller memory that is already page-locked (msckf_host_alloc, cudaHostRegister, torch pin_memory) is copied to the
// device straight from where it lies; everything else goes through the pinned arena first.
static bool is_pinned(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}
static int stage_commit(msckf_ctx* ctx, Packer& p, bool overlap = false) {
  Staging* s = p.s;
  CK(cudaEventSynchronize(s->done));       // previous user of this arena has finished
  if (p.off > s->cap) {
    if (s->h) cudaFreeHost(s->h);
    if (s->d) cudaFree(s->d);
    s->cap = std::max<size_t>(p.off * 2, 1 << 20);
    CK(cudaMallocHost((void**)&s->h, s->cap));
    CK(cudaMalloc((void**)&s->d, s->cap));
  }
  cudaStream_t cs = overlap ? ctx->copy_st : ctx->st;
  for (auto& part : p.parts) {
    const size_t bytes = part.second.second;
    if (!bytes) continue;
    const void* src = part.second.first;
    if (bytes >= (64u << 10) && is_pinned(src)) {
      CK(cudaMemcpyAsync(s->d + part.first, src, bytes, cudaMemcpyHostToDevice, cs));
    } else {
      memcpy(s->h + part.first, src, bytes);
      CK(cudaMemcpyAsync(s->d + part.first, s->h + part.first, bytes, cudaMemcpyHostToDevice, cs));
    }
  }
  if (overlap) {
    CK(cudaEventRecord(ctx->h2d_done, cs));
    CK(cudaStreamWaitEvent(ctx->st, ctx->h2d_done, 0));
  }
  return 0;
}
static int stage_end(msckf_ctx* ctx, Packer& p) {
  CK(cudaEventRecord(p.s->done, ctx->st));
  return 0;
}

extern "C" {

const char* msckf_version(void) { return "msckf_b200 sm_100a cuda " "12.9"; }

const char* msckf_last_error(const msckf_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int msckf_create(const msckf_config* cfg, msckf_ctx** out) {
  if (!cfg || !out) { g_create_error = "null argument"; return MSCKF_ERR_ARG; }
  *out = nullptr;
  if (cfg->n_filters < 1 || cfg->max_cam_states < 5 || cfg->max_cam_states > 64 || cfg->feature_capacity < 8) {
    g_create_error = "bad argument: n_filters >= 1, 5 <= max_cam_states <= 64, feature_capacity >= 8";
    return MSCKF_ERR_ARG;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0 || cfg->device < 0 || cfg->device >= ndev) {
    g_create_error = std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range");
    return MSCKF_ERR_CUDA;
  }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, cfg->device);
  if (prop.major != 10) {
    g_create_error = std::string("device is not sm_100-class: ") + prop.name + " (this library ships sm_100a code only, no fallback)";
    return MSCKF_ERR_CUDA;
  }
  msckf_ctx* ctx = new msckf_ctx();
  ctx->cfg = *cfg;
  ctx->device = cfg->device;
  auto fail = [&](int rc) {
    g_create_error = ctx->err;
    msckf_destroy(ctx);
    return rc;
  };
#define CKC(call) do { int rc_ = [&]() -> int { CK(call); return 0; }(); if (rc_) return fail(rc_); } while (0)
  CKC(cudaSetDevice(cfg->device));
  CKC(cudaStreamCreateWithFlags(&ctx->st, cudaStreamNonBlocking));
  for (auto& s : ctx->stg) CKC(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  for (auto& evn : ctx->ev) CKC(cudaEventCreate(&evn));
  for (auto& evn : ctx->mark) CKC(cudaEventCreate(&evn));
  for (auto& s : ctx->lanes.aux) CKC(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  for (auto& evn : ctx->lanes.join) CKC(cudaEventCreateWithFlags(&evn, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->lanes.fork, cudaEventDisableTiming));
  CKC(cudaEventCreateWithFlags(&ctx->lanes.built, cudaEventDisableTiming));
  ctx->lanes.enabled = getenv("MSCKF_NO_LANES") == nullptr;
  ctx->graph_enabled = getenv("MSCKF_NO_GRAPH") == nullptr;

  Dev& d = ctx->d;
  memset(&d, 0, sizeof d);
  d.host_timer = &ctx->kt;
  d.host_lanes = &ctx->lanes;
  d.F = cfg->n_filters;
  d.Nmax = cfg->max_cam_states;
  d.nmax = NI + 6 * d.Nmax;
  d.ld = (d.nmax + 7) & ~7;
  d.Fcap = cfg->feature_capacity;
  int hs = 16;
  while (hs < 2 * d.Fcap) hs <<= 1;
  d.Hs = hs;
  const int row_cap = cfg->row_cap;
  d.max_rows = cfg->max_rows > 0 ? cfg->max_rows : ((row_cap > 0 ? row_cap : 1500) + 4 * d.Nmax + 8);
  d.Tk = d.nmax;
  d.Sld = (d.Tk + 7) & ~7;
  d.Uinv_stride = (long long)((d.Sld + 31) / 32) * 32 * 32;
  d.A_elems = (long long)((d.Tk + 3) / 4 * 4) * (d.nmax + 1);
  d.Adense_elems = (((long long)d.max_rows + 3) / 4 * 4) * (d.nmax + 1);
  // Jacobian records: (4M-3) x (6M+1) doubles per landmark that reaches the gate in one stage
  long long pool;
  if (cfg->jacobian_pool_bytes > 0) {
    pool = cfg->jacobian_pool_bytes / 8;
  } else {    // typical tracks for every filter, and up to 2 GB towards the worst case (every landmark seen by the whole window)
    const long long base = (long long)d.F * ((long long)d.Fcap * 96 + 4 * jac_record_doubles(d.Nmax));
    const long long worst = (long long)d.F * d.Fcap * jac_record_doubles(d.Nmax);
    pool = std::max(base, std::min<long long>(worst, 256LL << 20));
  }
  d.pool_cap = pool;
  {
    const long long M = d.Nmax;
    d.jac_big_smem_doubles = 0;                             // (sized by the launcher, k_jacobian.cu)
    d.jac_big_grid = prop.multiProcessorCount * 2;
    const long long Rp = (4 * M - 3 + 7) & ~7LL;
    d.jac_scratch_stride = Rp * (Rp + 4);                    // jb_S_doubles(Nmax): the innovation covariance of the longest track
    {   // stacked rows: the capped lost-landmark update, or the pruning update (5 rows per live landmark); everything when the cap is lifted
      const long long all = (long long)d.Fcap * (4 * M - 3);
      const long long capped = std::max<long long>(d.max_rows, 5LL * d.Fcap + 8);
      d.rowmap_cap = (int)(row_cap > 0 ? std::min(all, capped) : all);
    }
    d.jac_rows_ok = getenv("MSCKF_NO_JAC_ROWS") ? 0 : 1;
    d.qr_tall_ok = getenv("MSCKF_NO_QR_TALL") ? 0 : 1;
    {
      const int kc[UPD_DMMA_CLASSES] = {16, 32, 48, 64, 80};
      for (int c = 0; c < UPD_DMMA_CLASSES; ++c)      // MSCKF_NO_DMMA: debugging knob, every update through the tiled kernels
        d.upd_dmma_ok[c] = (!getenv("MSCKF_NO_DMMA") && upd_dmma_smem_bytes(d.nmax, kc[c]) <= 220 * 1024) ? 1 : 0;
    }
    {   // TSQR of wide / tall stacks (k_qr_wide): up to leaf_cap leaves per filter so that a single filter's stack
        // spreads over the GPU, none when the batch already fills it; a tree of ceil(log2(leaf_cap)) levels
      d.qr_capcols = (int)((6 * M + 1 + 7) & ~7LL);
      long long lc = std::max<long long>(1, (2LL * prop.multiProcessorCount) / d.F);
      lc = std::min<long long>(lc, prop.multiProcessorCount);
      lc = std::min<long long>(lc, std::max<long long>(1, ((long long)d.rowmap_cap + 511) / 512));
      d.qr_leaf_cap = (int)lc;
      d.qr_levels = 0;
      while ((1 << d.qr_levels) < d.qr_leaf_cap) d.qr_levels += 1;
      d.Rnode_stride = (long long)d.qr_capcols * d.qr_capcols;
    }
  }

  Params& hp = ctx->hp;
  memset(&hp, 0, sizeof hp);
  hp.Qc[0] = cfg->gyro_noise; hp.Qc[1] = cfg->gyro_bias_noise; hp.Qc[2] = cfg->acc_noise; hp.Qc[3] = cfg->acc_bias_noise;
  hp.obs_noise = cfg->observation_noise;
  hp.P0[0] = cfg->gyro_bias_cov; hp.P0[1] = cfg->velocity_cov; hp.P0[2] = cfg->acc_bias_cov;
  hp.P0[3] = cfg->extrinsic_rotation_cov; hp.P0[4] = cfg->extrinsic_translation_cov;
  hp.pos_std_thr = cfg->position_std_threshold;
  hp.tri_translation_thr = cfg->translation_threshold; hp.huber_eps = cfg->huber_epsilon;
  hp.est_precision = cfg->estimation_precision; hp.init_damping = cfg->initial_damping;
  hp.outer_max = cfg->outer_loop_max_iteration; hp.inner_max = cfg->inner_loop_max_iteration;
  hp.prune_rot_thr = cfg->prune_rotation_threshold; hp.prune_trans_thr = cfg->prune_translation_threshold;
  hp.prune_track_thr = cfg->prune_tracking_rate_threshold;
  memcpy(hp.R01, cfg->R_cam0_cam1, sizeof hp.R01);
  memcpy(hp.t01, cfg->t_cam0_cam1, sizeof hp.t01);
  memcpy(hp.chi2, cfg->chi2_table, sizeof hp.chi2);
  hp.max_cam = d.Nmax;
  hp.row_cap = row_cap;

  const size_t F = d.F, Fc = d.Fcap, Nm = d.Nmax;
  Params* dprm = nullptr;
#define DA(ptr, count) do { int rc_ = dalloc(ctx, &(ptr), (count)); if (rc_) return fail(rc_); } while (0)
  DA(dprm, 1);
  CKC(cudaMemcpyAsync(dprm, &hp, sizeof hp, cudaMemcpyHostToDevice, ctx->st));
  d.prm = dprm;
  DA(d.imu, F * IMU_STRIDE);
  DA(d.cam_id, F * Nm); DA(d.cam_phys, F * Nm); DA(d.cam, F * Nm * CAM_STRIDE); DA(d.ncam, F); DA(d.phys_used, F); DA(d.obs_N, F); DA(d.obs_phys, F); DA(d.cam_pre, F * Nm * CAM_PRE);
  DA(d.P, 2 * F * (size_t)d.ld * d.ld); DA(d.Pcur, F); DA(d.p_compacted, F);
  DA(d.f_id, F * Fc); DA(d.f_mask, F * Fc); DA(d.f_flags, F * Fc); DA(d.f_pos, F * Fc * 3); DA(d.f_obs, F * Fc * Nm * 4);
  DA(d.ord, F * Fc); DA(d.ord_tmp, F * Fc); DA(d.nlive, F); DA(d.free_stack, F * Fc); DA(d.nfree, F);
  DA(d.h_key, F * (size_t)d.Hs); DA(d.h_val, F * (size_t)d.Hs);
  DA(d.cand, F * Fc); DA(d.cand_status, F * Fc); DA(d.ncand, F);
  DA(d.item_cand, F * Fc); DA(d.item_mask, F * Fc); DA(d.item_rows, F * Fc); DA(d.item_off, F * Fc); DA(d.item_rowoff, F * Fc);
  DA(d.item_dof, F * Fc); DA(d.item_gamma, F * Fc); DA(d.item_pass, F * Fc); DA(d.nitem, F);
  DA(d.rm_mask, F); DA(d.rm_idx, F * 2);
  DA(d.pool, (size_t)d.pool_cap); DA(d.stage_ctr, SC_COUNT); DA(d.item_lmask, F * Fc); DA(d.item_slot, F * Fc);
  DA(d.wl_tri, F * Fc); DA(d.wl_jac_small, JAC_CLASSES * F * Fc); DA(d.wl_jac_big, F * Fc); DA(d.wl_jac_big2, F * Fc); DA(d.wl_lu, F * Fc);
  { const size_t Rm = 4 * Nm - 3; DA(d.lu_scratch, Rm * 6 * Nm + Rm * Rm + 3 * Rm + 8); }
  DA(d.jac_scratch, (size_t)d.jac_big_grid * d.jac_scratch_stride);
  DA(d.Rnode, F * (size_t)d.qr_leaf_cap * (size_t)d.Rnode_stride);
  DA(d.Adense, (size_t)d.Adense_elems);
  DA(d.A, F * (size_t)d.A_elems); DA(d.m_rows, F); DA(d.m_pad, F); DA(d.n_act, F); DA(d.k_rows, F); DA(d.colmap, F * (size_t)d.nmax); DA(d.upd_path, F); DA(d.wl_upd_big, F); DA(d.wl_upd_dmma, (size_t)UPD_DMMA_CLASSES * F);
  DA(d.rowmap, F * (size_t)d.rowmap_cap); DA(d.qr_cls, F); DA(d.wl_qr, (size_t)QR_DMMA_CLASSES * F); DA(d.wl_qrb, F); DA(d.wl_qrt, F);
  DA(d.T, F * (size_t)d.Tk * Tld(d)); DA(d.S, F * (size_t)d.Tk * d.Sld); DA(d.TactT, F * (size_t)d.Tk * d.Sld); DA(d.Uinv, F * (size_t)d.Uinv_stride); DA(d.dx, F * (size_t)d.nmax);
  DA(d.log_id, 2 * F * Fc); DA(d.log_status, 2 * F * Fc); DA(d.log_gamma, 2 * F * Fc); DA(d.log_dof, 2 * F * Fc);
  DA(d.log_n, 2 * F); DA(d.log_upd, 2 * F * 2); DA(d.log_rm, F * 2);
  DA(d.frame_in, 1); DA(d.counters, F * NCNT); DA(d.err, 8); DA(d.did_reset, F); DA(d.imu_pub, F * IMU_STRIDE);
#undef DA
  CKC(cudaMallocHost((void**)&ctx->h_pub, F * IMU_STRIDE * sizeof(double)));
  CKC(cudaMallocHost((void**)&ctx->h_reset, F));
  CKC(cudaMallocHost((void**)&ctx->h_err, 8 * sizeof(unsigned)));
  memset(ctx->h_err, 0, 8 * sizeof(unsigned));
  for (auto& s : ctx->slot) {
    CKC(cudaMallocHost((void**)&s.pub, F * IMU_STRIDE * sizeof(double)));
    CKC(cudaMallocHost((void**)&s.reset, F));
    CKC(cudaMallocHost((void**)&s.err, 8 * sizeof(unsigned)));
    memset(s.err, 0, 8 * sizeof(unsigned));
    CKC(cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  }
  CKC(cudaStreamCreateWithFlags(&ctx->copy_st, cudaStreamNonBlocking));
  CKC(cudaEventCreateWithFlags(&ctx->h2d_done, cudaEventDisableTiming));
  // initial state: identity orientation, gravity from the caller later (set_imu_state), reset covariance
  {
    std::vector<double> imu0(F * IMU_STRIDE, 0.0);
    for (size_t f = 0; f < F; ++f) {
      imu0[f * IMU_STRIDE + MSCKF_IMU_Q + 3] = 1.0;
      imu0[f * IMU_STRIDE + MSCKF_IMU_Q_NULL + 3] = 1.0;
      for (int i = 0; i < 3; ++i) imu0[f * IMU_STRIDE + MSCKF_IMU_R_IC + 4 * i] = 1.0;
      imu0[f * IMU_STRIDE + MSCKF_IMU_GRAVITY + 2] = -9.81;
    }
    CKC(cudaMemcpyAsync(d.imu, imu0.data(), imu0.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->st));
    CKC(cudaStreamSynchronize(ctx->st));
  }
  launch_reset(d, -1, 0, ctx->st);
  CKC(cudaStreamSynchronize(ctx->st));
  CKC(cudaGetLastError());
#undef CKC
  *out = ctx;
  return MSCKF_OK;
}

void msckf_destroy(msckf_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->st) cudaStreamSynchronize(ctx->st);
  for (void* p : ctx->dev_allocs) cudaFree(p);
  for (auto& s : ctx->stg) {
    if (s.h) cudaFreeHost(s.h);
    if (s.d) cudaFree(s.d);
    if (s.done) cudaEventDestroy(s.done);
  }
  for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
  ctx->kt.release();
  for (auto& e : ctx->mark) if (e) cudaEventDestroy(e);
  if (ctx->frame_graph) cudaGraphExecDestroy(ctx->frame_graph);
  for (auto& s : ctx->lanes.aux) if (s) { cudaStreamSynchronize(s); cudaStreamDestroy(s); }
  for (auto& e : ctx->lanes.join) if (e) cudaEventDestroy(e);
  if (ctx->lanes.fork) cudaEventDestroy(ctx->lanes.fork);
  if (ctx->lanes.built) cudaEventDestroy(ctx->lanes.built);
  if (ctx->h_pub) cudaFreeHost(ctx->h_pub);
  if (ctx->h_reset) cudaFreeHost(ctx->h_reset);
  if (ctx->h_err) cudaFreeHost(ctx->h_err);
  for (auto& s : ctx->slot) {
    if (s.pub) cudaFreeHost(s.pub);
    if (s.reset) cudaFreeHost(s.reset);
    if (s.err) cudaFreeHost(s.err);
    if (s.done) cudaEventDestroy(s.done);
  }
  if (ctx->copy_st) { cudaStreamSynchronize(ctx->copy_st); cudaStreamDestroy(ctx->copy_st); }
  if (ctx->h2d_done) cudaEventDestroy(ctx->h2d_done);
  if (ctx->st) cudaStreamDestroy(ctx->st);
  delete ctx;
}

int msckf_host_alloc(size_t bytes, void** out) {
  if (!out) return MSCKF_ERR_ARG;
  *out = nullptr;
  return cudaMallocHost(out, bytes ? bytes : 1) == cudaSuccess ? MSCKF_OK : MSCKF_ERR_CUDA;
}
int msckf_host_free(void* p) {
  if (!p) return MSCKF_OK;
  return cudaFreeHost(p) == cudaSuccess ? MSCKF_OK : MSCKF_ERR_CUDA;
}

int msckf_sync(msckf_ctx* ctx) {
  if (!ctx) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  return sync_and_check(ctx);
}

// ------------------------------------------------------------------ state I/O
int msckf_set_imu_state(msckf_ctx* ctx, int32_t f, const double* imu) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(f >= 0 && f < ctx->d.F && imu, "filter index / pointer");
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->st));
  CK(cudaMemcpy(ctx->d.imu + (size_t)f * IMU_STRIDE, imu, IMU_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  return MSCKF_OK;
}
int msckf_get_imu_state(msckf_ctx* ctx, int32_t f, double* imu) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(f >= 0 && f < ctx->d.F && imu, "filter index / pointer");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  CK(cudaMemcpy(imu, ctx->d.imu + (size_t)f * IMU_STRIDE, IMU_STRIDE * sizeof(double), cudaMemcpyDeviceToHost));
  return MSCKF_OK;
}
int msckf_get_imu_states(msckf_ctx* ctx, double* imu) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(imu, "pointer");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  CK(cudaMemcpy(imu, ctx->d.imu, (size_t)ctx->d.F * IMU_STRIDE * sizeof(double), cudaMemcpyDeviceToHost));
  return MSCKF_OK;
}
int msckf_get_cam_states(msckf_ctx* ctx, int32_t f, int32_t* n_cam, int64_t* ids, double* cams) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(f >= 0 && f < ctx->d.F && n_cam, "filter index / pointer");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  const Dev& d = ctx->d;
  int n = 0;
  CK(cudaMemcpy(&n, d.ncam + f, sizeof(int), cudaMemcpyDeviceToHost));
  *n_cam = n;
  if (ids && n) CK(cudaMemcpy(ids, d.cam_id + (size_t)f * d.Nmax, n * sizeof(int64_t), cudaMemcpyDeviceToHost));
  if (cams && n) CK(cudaMemcpy(cams, d.cam + (size_t)f * d.Nmax * CAM_STRIDE, (size_t)n * CAM_STRIDE * sizeof(double), cudaMemcpyDeviceToHost));
  return MSCKF_OK;
}
int msckf_set_cam_states(msckf_ctx* ctx, int32_t f, int32_t n, const int64_t* ids, const double* cams) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && n >= 0 && n <= d.Nmax && (n == 0 || (ids && cams)), "filter index / count / pointers");
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->st));
  // keeps the physical slots of ids that are already present (their observations stay valid)
  int n_old = 0;
  CK(cudaMemcpy(&n_old, d.ncam + f, sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<long long> old_ids(d.Nmax);
  std::vector<int> old_phys(d.Nmax), phys(d.Nmax);
  if (n_old) {
    CK(cudaMemcpy(old_ids.data(), d.cam_id + (size_t)f * d.Nmax, n_old * sizeof(long long), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(old_phys.data(), d.cam_phys + (size_t)f * d.Nmax, n_old * sizeof(int), cudaMemcpyDeviceToHost));
  }
  unsigned long long used = 0;
  std::vector<char> have(n, 0);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n_old; ++j)
      if (old_ids[j] == ids[i]) { phys[i] = old_phys[j]; used |= 1ull << phys[i]; have[i] = 1; }
  for (int i = 0; i < n; ++i)
    if (!have[i]) {
      int s = 0;
      while ((used >> s) & 1ull) ++s;
      phys[i] = s;
      used |= 1ull << s;
    }
  if (n) {
    CK(cudaMemcpy(d.cam_id + (size_t)f * d.Nmax, ids, n * sizeof(int64_t), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d.cam_phys + (size_t)f * d.Nmax, phys.data(), n * sizeof(int), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d.cam + (size_t)f * d.Nmax * CAM_STRIDE, cams, (size_t)n * CAM_STRIDE * sizeof(double), cudaMemcpyHostToDevice));
  }
  CK(cudaMemcpy(d.ncam + f, &n, sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.phys_used + f, &used, sizeof(used), cudaMemcpyHostToDevice));
  return MSCKF_OK;
}
int msckf_get_cov(msckf_ctx* ctx, int32_t f, double* P, int32_t n) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && P && n >= NI && n <= d.nmax, "filter index / size");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  double* Pf = nullptr;                  // the filter's current covariance buffer
  CK(cudaMemcpy(&Pf, d.Pcur + f, sizeof Pf, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy2D(P, (size_t)n * sizeof(double), Pf, (size_t)d.ld * sizeof(double),
                  (size_t)n * sizeof(double), n, cudaMemcpyDeviceToHost));
  return MSCKF_OK;
}
int msckf_set_cov(msckf_ctx* ctx, int32_t f, const double* P, int32_t n) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && P && n >= NI && n <= d.nmax, "filter index / size");
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->st));
  double* Pf = nullptr;
  CK(cudaMemcpy(&Pf, d.Pcur + f, sizeof Pf, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy2D(Pf, (size_t)d.ld * sizeof(double), P, (size_t)n * sizeof(double),
                  (size_t)n * sizeof(double), n, cudaMemcpyHostToDevice));
  return MSCKF_OK;
}

// host copy of one filter's landmark tables
struct StoreDump {
  int nlive = 0, ncam = 0;
  std::vector<int> ord, flags, phys;
  std::vector<long long> ids;
  std::vector<unsigned long long> mask;
};
static int dump_store(msckf_ctx* ctx, int f, StoreDump& s) {
  const Dev& d = ctx->d;
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  CK(cudaMemcpy(&s.nlive, d.nlive + f, sizeof(int), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&s.ncam, d.ncam + f, sizeof(int), cudaMemcpyDeviceToHost));
  s.ord.resize(d.Fcap); s.flags.resize(d.Fcap); s.ids.resize(d.Fcap); s.mask.resize(d.Fcap); s.phys.resize(d.Nmax);
  CK(cudaMemcpy(s.ord.data(), d.ord + (size_t)f * d.Fcap, d.Fcap * sizeof(int), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(s.flags.data(), d.f_flags + (size_t)f * d.Fcap, d.Fcap * sizeof(int), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(s.ids.data(), d.f_id + (size_t)f * d.Fcap, d.Fcap * sizeof(long long), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(s.mask.data(), d.f_mask + (size_t)f * d.Fcap, d.Fcap * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(s.phys.data(), d.cam_phys + (size_t)f * d.Nmax, d.Nmax * sizeof(int), cudaMemcpyDeviceToHost));
  return 0;
}
static int find_slot(const StoreDump& s, long long id) {
  for (int j = 0; j < s.nlive; ++j)
    if (s.ids[s.ord[j]] == id) return s.ord[j];
  return -1;
}

int msckf_get_features(msckf_ctx* ctx, int32_t f, int32_t* n_features, int64_t* ids, uint64_t* obs_mask, uint8_t* is_init,
                       double* position, double* obs) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && n_features, "filter index / pointer");
  CK(cudaSetDevice(ctx->device));
  StoreDump s;
  int rc = dump_store(ctx, f, s);
  if (rc) return rc;
  *n_features = s.nlive;
  std::vector<double> pos, ob;
  if (position) { pos.resize((size_t)d.Fcap * 3); CK(cudaMemcpy(pos.data(), d.f_pos + (size_t)f * d.Fcap * 3, pos.size() * 8, cudaMemcpyDeviceToHost)); }
  if (obs) { ob.resize((size_t)d.Fcap * d.Nmax * 4); CK(cudaMemcpy(ob.data(), d.f_obs + (size_t)f * d.Fcap * d.Nmax * 4, ob.size() * 8, cudaMemcpyDeviceToHost)); }
  for (int j = 0; j < s.nlive; ++j) {
    const int slot = s.ord[j];
    if (ids) ids[j] = s.ids[slot];
    if (is_init) is_init[j] = (s.flags[slot] & FF_INIT) ? 1 : 0;
    if (obs_mask) {
      uint64_t lm = 0;
      for (int c = 0; c < s.ncam; ++c)
        if ((s.mask[slot] >> s.phys[c]) & 1ull) lm |= 1ull << c;
      obs_mask[j] = lm;
    }
    if (position) for (int k = 0; k < 3; ++k) position[3 * j + k] = pos[(size_t)slot * 3 + k];
    if (obs)
      for (int c = 0; c < d.Nmax; ++c)
        for (int k = 0; k < 4; ++k)
          obs[((size_t)j * d.Nmax + c) * 4 + k] = (c < s.ncam) ? ob[((size_t)slot * d.Nmax + s.phys[c]) * 4 + k] : 0.0;
  }
  return MSCKF_OK;
}

int msckf_set_features(msckf_ctx* ctx, int32_t f, int32_t n, const int64_t* ids, const uint64_t* obs_mask, const uint8_t* is_init,
                       const double* position, const double* obs) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && n >= 0 && n <= d.Fcap && (n == 0 || (ids && obs_mask && obs)), "filter / count / pointers");
  CK(cudaSetDevice(ctx->device));
  CK(cudaStreamSynchronize(ctx->st));
  int ncam = 0;
  CK(cudaMemcpy(&ncam, d.ncam + f, sizeof(int), cudaMemcpyDeviceToHost));
  std::vector<int> phys(d.Nmax, 0);
  if (ncam) CK(cudaMemcpy(phys.data(), d.cam_phys + (size_t)f * d.Nmax, ncam * sizeof(int), cudaMemcpyDeviceToHost));
  const size_t Fc = d.Fcap, Nm = d.Nmax;
  std::vector<long long> hid(Fc, 0);
  std::vector<unsigned long long> hmask(Fc, 0ull);
  std::vector<int> hflags(Fc, 0), hord(Fc, 0), hfree(Fc, 0);
  std::vector<double> hpos(Fc * 3, 0.0), hobs(Fc * Nm * 4, 0.0);
  for (int j = 0; j < n; ++j) {            // landmark j takes slot j; map order = argument order
    hid[j] = ids[j];
    hflags[j] = FF_LIVE | ((is_init && is_init[j]) ? FF_INIT : 0);
    hord[j] = j;
    if (position) for (int k = 0; k < 3; ++k) hpos[(size_t)j * 3 + k] = position[(size_t)j * 3 + k];
    ARG((obs_mask[j] >> ncam) == 0 || ncam == 64, "obs_mask has bits beyond the camera-state window");
    for (int c = 0; c < ncam; ++c)
      if ((obs_mask[j] >> c) & 1ull) {
        hmask[j] |= 1ull << phys[c];
        for (int k = 0; k < 4; ++k) hobs[((size_t)j * Nm + phys[c]) * 4 + k] = obs[((size_t)j * Nm + c) * 4 + k];
      }
  }
  const int nfree = d.Fcap - n;
  for (int i = 0; i < nfree; ++i) hfree[i] = d.Fcap - 1 - i;
  CK(cudaMemcpy(d.f_id + f * Fc, hid.data(), Fc * sizeof(long long), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.f_mask + f * Fc, hmask.data(), Fc * sizeof(unsigned long long), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.f_flags + f * Fc, hflags.data(), Fc * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.ord + f * Fc, hord.data(), Fc * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.free_stack + f * Fc, hfree.data(), Fc * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.f_pos + f * Fc * 3, hpos.data(), Fc * 3 * sizeof(double), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.f_obs + f * Fc * Nm * 4, hobs.data(), Fc * Nm * 4 * sizeof(double), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.nlive + f, &n, sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.nfree + f, &nfree, sizeof(int), cudaMemcpyHostToDevice));
  return MSCKF_OK;
}

int msckf_set_feature_position(msckf_ctx* ctx, int32_t f, int64_t id, const double* position, int32_t is_initialized) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && position, "filter index / pointer");
  CK(cudaSetDevice(ctx->device));
  StoreDump s;
  int rc = dump_store(ctx, f, s);
  if (rc) return rc;
  const int slot = find_slot(s, id);
  ARG(slot >= 0, "unknown feature id");
  CK(cudaMemcpy(d.f_pos + ((size_t)f * d.Fcap + slot) * 3, position, 3 * sizeof(double), cudaMemcpyHostToDevice));
  int fl = (s.flags[slot] & ~FF_INIT) | (is_initialized ? FF_INIT : 0);
  CK(cudaMemcpy(d.f_flags + (size_t)f * d.Fcap + slot, &fl, sizeof(int), cudaMemcpyHostToDevice));
  return MSCKF_OK;
}

// ------------------------------------------------------------------ stage enqueue helpers
static void enqueue_lost_stage(msckf_ctx* ctx) {
  const Dev& d = ctx->d;
  cudaMemsetAsync(d.stage_ctr, 0, SC_COUNT * sizeof(unsigned long long), ctx->st);
  {   // independent bookkeeping of the stage, side by side
    Fork fk(d, ctx->st);
    launch_cam_prepare(d, fk.lane(1));
    launch_classify_lost(d, fk.lane(0));
  }
  launch_triangulate_store(d, ctx->st);
  launch_plan_items(d, 0, ctx->st);
  launch_jacobian_gate(d, 0xE, 1, 0, ctx->st);            // lost landmarks have at least three observations
  launch_plan_stack(d, 0, ctx->st);
  launch_compress_update(d, 0, ctx->st);
}
static void enqueue_prune_stage(msckf_ctx* ctx) {
  const Dev& d = ctx->d;
  cudaMemsetAsync(d.stage_ctr, 0, SC_COUNT * sizeof(unsigned long long), ctx->st);
  {
    Fork fk(d, ctx->st);
    launch_cam_prepare(d, fk.lane(1));
    launch_prune_select(d, fk.lane(0));
  }
  launch_triangulate_store(d, ctx->st);
  launch_plan_items(d, 1, ctx->st);
  launch_jacobian_gate(d, 0x1, 0, 0, ctx->st);            // the pruning update always involves exactly two camera states
  launch_plan_stack(d, 1, ctx->st);
  launch_compress_update(d, 1, ctx->st);
  launch_remove_cams(d, ctx->st);
}

int msckf_propagate(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro, const double* acc) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(imu_off, "imu_off");
  CK(cudaSetDevice(ctx->device));
  const int F = ctx->d.F, n = imu_off[F];
  ARG(n == 0 || (ts && gyro && acc), "sample pointers");
  Packer p;
  stage_begin(ctx, p);
  const size_t o0 = p.add(imu_off, (F + 1) * sizeof(int)), o1 = p.add(ts, n * sizeof(double));
  const size_t o2 = p.add(gyro, n * 3 * sizeof(double)), o3 = p.add(acc, n * 3 * sizeof(double));
  int rc = stage_commit(ctx, p);
  if (rc) return rc;
  char* b = p.s->d;
  launch_propagate(ctx->d, (const int*)(b + o0), (const double*)(b + o1), (const double*)(b + o2), (const double*)(b + o3), ctx->st);
  stage_end(ctx, p);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_augment(msckf_ctx* ctx, const int64_t* cam_ids) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(cam_ids, "cam_ids");
  CK(cudaSetDevice(ctx->device));
  Packer p;
  stage_begin(ctx, p);
  const size_t o0 = p.add(cam_ids, ctx->d.F * sizeof(int64_t));
  int rc = stage_commit(ctx, p);
  if (rc) return rc;
  launch_augment(ctx->d, (const long long*)(p.s->d + o0), ctx->st);
  stage_end(ctx, p);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_add_observations(msckf_ctx* ctx, const int32_t* feat_off, const int64_t* ids, const double* uv) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(feat_off, "feat_off");
  CK(cudaSetDevice(ctx->device));
  const int F = ctx->d.F, n = feat_off[F];
  ARG(n == 0 || (ids && uv), "feature pointers");
  Packer p;
  stage_begin(ctx, p);
  const size_t o0 = p.add(feat_off, (F + 1) * sizeof(int)), o1 = p.add(ids, n * sizeof(int64_t)), o2 = p.add(uv, (size_t)n * 4 * sizeof(double));
  int rc = stage_commit(ctx, p);
  if (rc) return rc;
  char* b = p.s->d;
  launch_observe(ctx->d, (const int*)(b + o0), (const long long*)(b + o1), (const double*)(b + o2), ctx->st);
  stage_end(ctx, p);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_remove_lost_features(msckf_ctx* ctx) {
  if (!ctx) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  enqueue_lost_stage(ctx);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_prune_cam_states(msckf_ctx* ctx) {
  if (!ctx) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  enqueue_prune_stage(ctx);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_online_reset(msckf_ctx* ctx, uint8_t* did_reset) {
  if (!ctx) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  launch_finish_frame(ctx->d, 1, ctx->st);
  CK(cudaMemcpyAsync(ctx->h_reset, ctx->d.did_reset, ctx->d.F, cudaMemcpyDeviceToHost, ctx->st));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  if (did_reset) memcpy(did_reset, ctx->h_reset, ctx->d.F);
  return MSCKF_OK;
}

int msckf_reset(msckf_ctx* ctx, int32_t f) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(f >= 0 && f < ctx->d.F, "filter index");
  CK(cudaSetDevice(ctx->device));
  launch_reset(ctx->d, f, 1, ctx->st);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

__global__ void k_set_frame_in(FrameIn* dst, FrameIn v) { *dst = v; }

static void enqueue_frame_eager(msckf_ctx* ctx, const FrameIn& in, const FrameIn* fin) {
  const Dev& d = ctx->d;
  const bool tm = ctx->timing && !fin;
  if (tm) cudaEventRecord(ctx->ev[0], ctx->st);
  // add_feature_observations only needs to know where the new camera state will sit (k_cam_slot): it runs on its own lane
  // beside propagation and augmentation
  launch_cam_slot(d, ctx->st);
  {
    Fork fk(d, ctx->st);
    launch_observe(d, in.feat_off, in.ids, in.uv, fk.lane(1), fin, 1);
    launch_propagate(d, in.imu_off, in.ts, in.gyro, in.acc, ctx->st, fin);
    launch_augment(d, in.cam_ids, ctx->st, fin, true);
  }
  if (tm) cudaEventRecord(ctx->ev[1], ctx->st);
  if (tm) cudaEventRecord(ctx->ev[2], ctx->st);
  enqueue_lost_stage(ctx);
  if (tm) cudaEventRecord(ctx->ev[3], ctx->st);
  enqueue_prune_stage(ctx);
  if (tm) cudaEventRecord(ctx->ev[4], ctx->st);
  launch_finish_frame(d, 1, ctx->st);
  if (tm) cudaEventRecord(ctx->ev[5], ctx->st);
}

// One frame's kernels on the context stream.  After two eager frames the sequence is captured into a CUDA graph whose
// message-consuming kernels read their pointers from d.frame_in; from then on a frame is one tiny kernel that stores
// the pointers plus one graph launch.  Stage timing / per-kernel timing run eagerly.
static int enqueue_frame(msckf_ctx* ctx, const FrameIn& in) {
  const Dev& d = ctx->d;
  const bool use_graph = ctx->graph_enabled && !ctx->timing && !ctx->kt.timing;
  if (!use_graph || ctx->eager_frames < 2) {
    ctx->eager_frames += 1;
    enqueue_frame_eager(ctx, in, nullptr);
    return MSCKF_OK;
  }
  k_set_frame_in<<<1, 1, 0, ctx->st>>>(d.frame_in, in);
  CK(cudaGetLastError());
  if (!ctx->frame_graph) {
    long long before[KID_COUNT];
    memcpy(before, ctx->kt.launches, sizeof before);
    cudaGraph_t g = nullptr;
    CK(cudaStreamBeginCapture(ctx->st, cudaStreamCaptureModeRelaxed));
    enqueue_frame_eager(ctx, in, d.frame_in);
    cudaError_t e = cudaStreamEndCapture(ctx->st, &g);
    for (int i = 0; i < KID_COUNT; ++i) {
      ctx->graph_launches[i] = ctx->kt.launches[i] - before[i];
      ctx->kt.launches[i] = before[i];
    }
    // a launch the runtime rejected during capture (KLAUNCH notes the first one) invalidates the graph as well
    if (e != cudaSuccess || !g || ctx->kt.bad_launch >= 0 || cudaGraphInstantiate(&ctx->frame_graph, g, 0) != cudaSuccess) {
      cudaGetLastError();
      if (g) cudaGraphDestroy(g);
      ctx->frame_graph = nullptr;
      ctx->kt.bad_launch = -1;
      ctx->graph_enabled = false;                   // capture not possible here: stay with plain launches
      enqueue_frame_eager(ctx, in, nullptr);
      return MSCKF_OK;
    }
    cudaGraphDestroy(g);
  }
  CK(cudaGraphLaunch(ctx->frame_graph, ctx->st));
  for (int i = 0; i < KID_COUNT; ++i) ctx->kt.launches[i] += ctx->graph_launches[i];
  return MSCKF_OK;
}

int msckf_feature_callback_async(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro, const double* acc,
                                 const int64_t* cam_ids, const int32_t* feat_off, const int64_t* ids, const double* uv) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(imu_off && cam_ids && feat_off, "offset / id pointers");
  ARG(ctx->n_pending < 2, "two frames already in flight: collect one first");
  CK(cudaSetDevice(ctx->device));
  const Dev& d = ctx->d;
  const int F = d.F, ni = imu_off[F], nf = feat_off[F];
  ARG((ni == 0 || (ts && gyro && acc)) && (nf == 0 || (ids && uv)), "sample / feature pointers");
  Packer p;
  stage_begin(ctx, p);
  const size_t o0 = p.add(imu_off, (F + 1) * sizeof(int)), o1 = p.add(ts, ni * sizeof(double));
  const size_t o2 = p.add(gyro, (size_t)ni * 3 * sizeof(double)), o3 = p.add(acc, (size_t)ni * 3 * sizeof(double));
  const size_t o4 = p.add(cam_ids, F * sizeof(int64_t)), o5 = p.add(feat_off, (F + 1) * sizeof(int));
  const size_t o6 = p.add(ids, nf * sizeof(int64_t)), o7 = p.add(uv, (size_t)nf * 4 * sizeof(double));
  int rc = stage_commit(ctx, p, true);
  if (rc) return rc;
  char* b = p.s->d;
  FrameIn in;
  in.imu_off = (const int*)(b + o0); in.ts = (const double*)(b + o1); in.gyro = (const double*)(b + o2); in.acc = (const double*)(b + o3);
  in.cam_ids = (const long long*)(b + o4); in.feat_off = (const int*)(b + o5); in.ids = (const long long*)(b + o6); in.uv = (const double*)(b + o7);
  rc = enqueue_frame(ctx, in);
  if (rc) return rc;
  stage_end(ctx, p);
  msckf_ctx::FrameSlot& fs = ctx->slot[(ctx->head + ctx->n_pending) & 1];
  CK(cudaMemcpyAsync(fs.pub, d.imu_pub, (size_t)F * IMU_STRIDE * sizeof(double), cudaMemcpyDeviceToHost, ctx->st));
  CK(cudaMemcpyAsync(fs.reset, d.did_reset, F, cudaMemcpyDeviceToHost, ctx->st));
  CK(cudaMemcpyAsync(fs.err, d.err, 8 * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->st));
  CK(cudaEventRecord(fs.done, ctx->st));
  CK(cudaGetLastError());
  ctx->n_pending += 1;
  return MSCKF_OK;
}

int msckf_feature_callback_dev(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro, const double* acc,
                               const int64_t* cam_ids, const int32_t* feat_off, const int64_t* ids, const double* uv,
                               double* imu_out_dev) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(imu_off && cam_ids && feat_off, "offset / id pointers");
  ARG(ctx->n_pending == 0, "previous frame not collected");
  CK(cudaSetDevice(ctx->device));
  const Dev& d = ctx->d;
  FrameIn in;
  in.imu_off = imu_off; in.ts = ts; in.gyro = gyro; in.acc = acc;
  in.cam_ids = (const long long*)cam_ids; in.feat_off = feat_off; in.ids = (const long long*)ids; in.uv = uv;
  int rc = enqueue_frame(ctx, in);
  if (rc) return rc;
  if (imu_out_dev)
    CK(cudaMemcpyAsync(imu_out_dev, d.imu_pub, (size_t)d.F * IMU_STRIDE * sizeof(double), cudaMemcpyDeviceToDevice, ctx->st));
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_collect(msckf_ctx* ctx, double* imu_out, uint8_t* did_reset) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(ctx->n_pending > 0, "no frame in flight");
  CK(cudaSetDevice(ctx->device));
  msckf_ctx::FrameSlot& fs = ctx->slot[ctx->head];
  ctx->head ^= 1;
  ctx->n_pending -= 1;
  CK(cudaEventSynchronize(fs.done));         // the oldest frame in flight; a younger one keeps running
  CK(cudaGetLastError());
  int rc = check_device_errors(ctx, fs.err);
  if (rc) return rc;
  if (imu_out) memcpy(imu_out, fs.pub, (size_t)ctx->d.F * IMU_STRIDE * sizeof(double));
  if (did_reset) memcpy(did_reset, fs.reset, ctx->d.F);
  if (ctx->timing) {
    for (int i = 0; i < 4; ++i) cudaEventElapsedTime(&ctx->last_ms[i], ctx->ev[i], ctx->ev[i + 1]);
    float fin = 0;
    cudaEventElapsedTime(&fin, ctx->ev[4], ctx->ev[5]);
    cudaEventElapsedTime(&ctx->last_ms[4], ctx->ev[0], ctx->ev[5]);
    ctx->last_ms[3] += fin;
  }
  return MSCKF_OK;
}

int msckf_feature_callback(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro, const double* acc,
                           const int64_t* cam_ids, const int32_t* feat_off, const int64_t* ids, const double* uv,
                           double* imu_out, uint8_t* did_reset) {
  int rc = msckf_feature_callback_async(ctx, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv);
  if (rc) return rc;
  return msckf_collect(ctx, imu_out, did_reset);
}

int msckf_set_timing(msckf_ctx* ctx, int32_t enabled) {
  if (!ctx) return MSCKF_ERR_ARG;
  ctx->timing = enabled != 0;
  return MSCKF_OK;
}
int msckf_last_timing(msckf_ctx* ctx, float* ms) {
  if (!ctx || !ms) return MSCKF_ERR_ARG;
  memcpy(ms, ctx->last_ms, sizeof ctx->last_ms);
  return MSCKF_OK;
}

// ------------------------------------------------------------------ logs / counters
int msckf_get_stage_log(msckf_ctx* ctx, int32_t f, int32_t stage, int32_t* n_entries, int64_t* feature_ids, int32_t* status,
                        double* gamma, int32_t* dof, int32_t* update_rows, int32_t* update_cols, int64_t* removed_cam_ids) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && (stage == 0 || stage == 1) && n_entries, "filter / stage / pointer");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  int n = 0;
  CK(cudaMemcpy(&n, d.log_n + stage * d.F + f, sizeof(int), cudaMemcpyDeviceToHost));
  *n_entries = n;
  const size_t base = ((size_t)stage * d.F + f) * d.Fcap;
  if (n && feature_ids) CK(cudaMemcpy(feature_ids, d.log_id + base, n * sizeof(int64_t), cudaMemcpyDeviceToHost));
  if (n && status) CK(cudaMemcpy(status, d.log_status + base, n * sizeof(int), cudaMemcpyDeviceToHost));
  if (n && gamma) CK(cudaMemcpy(gamma, d.log_gamma + base, n * sizeof(double), cudaMemcpyDeviceToHost));
  if (n && dof) CK(cudaMemcpy(dof, d.log_dof + base, n * sizeof(int), cudaMemcpyDeviceToHost));
  int upd[2] = {0, 0};
  CK(cudaMemcpy(upd, d.log_upd + ((size_t)stage * d.F + f) * 2, sizeof upd, cudaMemcpyDeviceToHost));
  if (update_rows) *update_rows = upd[0];
  if (update_cols) *update_cols = upd[1];
  if (removed_cam_ids && stage == 1) CK(cudaMemcpy(removed_cam_ids, d.log_rm + 2 * (size_t)f, 2 * sizeof(int64_t), cudaMemcpyDeviceToHost));
  return MSCKF_OK;
}

int msckf_get_counters(msckf_ctx* ctx, int64_t* counters) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(counters, "pointer");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  CK(cudaMemcpy(counters, ctx->d.counters, (size_t)ctx->d.F * NCNT * sizeof(int64_t), cudaMemcpyDeviceToHost));
  return MSCKF_OK;
}

int msckf_get_update_shapes(msckf_ctx* ctx, int32_t* shapes) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(shapes, "pointer");
  CK(cudaSetDevice(ctx->device));
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  const size_t F = (size_t)ctx->d.F;
  std::vector<int32_t> tmp(4 * F);
  const int* src[4] = {ctx->d.m_rows, ctx->d.n_act, ctx->d.k_rows, ctx->d.upd_path};
  for (int a = 0; a < 4; ++a) CK(cudaMemcpy(tmp.data() + a * F, src[a], F * sizeof(int32_t), cudaMemcpyDeviceToHost));
  for (size_t f = 0; f < F; ++f)
    for (int a = 0; a < 4; ++a) shapes[4 * f + a] = tmp[a * F + f];
  return MSCKF_OK;
}

// ------------------------------------------------------------------ directly callable pieces
static int triangulate_call(msckf_ctx* ctx, int mode, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q,
                            const double* cam_p, double* position, uint8_t* valid, int32_t* lm_solves) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(n >= 0 && obs_off, "count / pointers");
  if (n == 0) return MSCKF_OK;
  CK(cudaSetDevice(ctx->device));
  const int total = obs_off[n];
  ARG(total > 0 && uv && cam_q && cam_p, "observation pointers");
  int max_obs = 0;
  for (int i = 0; i < n; ++i) {
    const int M = obs_off[i + 1] - obs_off[i];
    ARG(M >= 1, "every landmark needs at least one observation");
    max_obs = std::max(max_obs, M);
  }
  ARG(max_obs <= 200, "at most 200 observations per landmark");
  Packer p;
  stage_begin(ctx, p);
  const size_t o0 = p.add(obs_off, (n + 1) * sizeof(int)), o1 = p.add(uv, (size_t)total * 4 * sizeof(double));
  const size_t o2 = p.add(cam_q, (size_t)total * 4 * sizeof(double)), o3 = p.add(cam_p, (size_t)total * 3 * sizeof(double));
  int rc = stage_commit(ctx, p);
  if (rc) return rc;
  double* dpos = nullptr;
  unsigned char* dval = nullptr;
  int* dsol = nullptr;
  CK(cudaMalloc((void**)&dpos, (size_t)n * 3 * sizeof(double)));
  CK(cudaMalloc((void**)&dval, n));
  CK(cudaMalloc((void**)&dsol, n * sizeof(int)));
  char* b = p.s->d;
  launch_triangulate_raw(ctx->d, n, (const int*)(b + o0), (const double*)(b + o1), (const double*)(b + o2), (const double*)(b + o3), dpos,
                         dval, dsol, max_obs, mode, ctx->st);
  stage_end(ctx, p);
  rc = sync_and_check(ctx);
  if (!rc) {
    if (position) cudaMemcpy(position, dpos, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost);
    if (valid) cudaMemcpy(valid, dval, n, cudaMemcpyDeviceToHost);
    if (lm_solves) cudaMemcpy(lm_solves, dsol, n * sizeof(int), cudaMemcpyDeviceToHost);
  }
  cudaFree(dpos); cudaFree(dval); cudaFree(dsol);
  return rc;
}

int msckf_triangulate(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q, const double* cam_p,
                      double* position, uint8_t* valid, int32_t* lm_solves) {
  if (ctx && !(position && valid)) { ctx->err = "bad argument: output pointers"; return MSCKF_ERR_ARG; }
  return triangulate_call(ctx, 0, n, obs_off, uv, cam_q, cam_p, position, valid, lm_solves);
}

int msckf_check_motion(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q, const double* cam_p,
                       uint8_t* moved) {
  if (ctx && !moved) { ctx->err = "bad argument: output pointer"; return MSCKF_ERR_ARG; }
  return triangulate_call(ctx, 1, n, obs_off, uv, cam_q, cam_p, nullptr, moved, nullptr);
}

int msckf_initial_guess(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q, const double* cam_p,
                        double* guess) {
  if (ctx && !guess) { ctx->err = "bad argument: output pointer"; return MSCKF_ERR_ARG; }
  return triangulate_call(ctx, 2, n, obs_off, uv, cam_q, cam_p, guess, nullptr, nullptr);
}

int msckf_initialize_position(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q,
                              const double* cam_p, double* position, uint8_t* valid, int32_t* lm_solves) {
  if (ctx && !(position && valid)) { ctx->err = "bad argument: output pointers"; return MSCKF_ERR_ARG; }
  return triangulate_call(ctx, 3, n, obs_off, uv, cam_q, cam_p, position, valid, lm_solves);
}

int msckf_kernel_launches(msckf_ctx* ctx, int64_t* n) {
  if (!ctx || !n) return MSCKF_ERR_ARG;
  long long t = 0;
  for (int i = 0; i < KID_COUNT; ++i) t += ctx->kt.launches[i];
  *n = t;
  return MSCKF_OK;
}

int msckf_mark(msckf_ctx* ctx, int32_t which) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(which == 0 || which == 1, "mark index");
  CK(cudaSetDevice(ctx->device));
  CK(cudaEventRecord(ctx->mark[which], ctx->st));
  return MSCKF_OK;
}
int msckf_mark_elapsed(msckf_ctx* ctx, float* ms) {
  if (!ctx || !ms) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  CK(cudaEventSynchronize(ctx->mark[1]));
  CK(cudaEventElapsedTime(ms, ctx->mark[0], ctx->mark[1]));
  return MSCKF_OK;
}

int msckf_mark_elapsed_between(msckf_ctx* first, msckf_ctx* last, float* ms) {
  if (!first || !last || !ms) return MSCKF_ERR_ARG;
  msckf_ctx* ctx = last;
  ARG(first->device == last->device, "contexts on different devices");
  CK(cudaSetDevice(ctx->device));
  CK(cudaEventSynchronize(last->mark[1]));
  CK(cudaEventElapsedTime(ms, first->mark[0], last->mark[1]));
  return MSCKF_OK;
}

int msckf_kernel_count(void) { return KID_COUNT; }
const char* msckf_kernel_name(int32_t i) { return (i >= 0 && i < KID_COUNT) ? kKernelNames[i] : ""; }

int msckf_set_kernel_timing(msckf_ctx* ctx, int32_t enabled) {
  if (!ctx) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  ctx->kt.flush(ctx->st);
  ctx->kt.timing = enabled != 0;
  return MSCKF_OK;
}

int msckf_kernel_times(msckf_ctx* ctx, int64_t* launches, int64_t* timed_launches, double* total_ms) {
  if (!ctx) return MSCKF_ERR_ARG;
  CK(cudaSetDevice(ctx->device));
  ctx->kt.flush(ctx->st);
  for (int i = 0; i < KID_COUNT; ++i) {
    if (launches) launches[i] = ctx->kt.launches[i];
    if (timed_launches) timed_launches[i] = ctx->kt.timed[i];
    if (total_ms) total_ms[i] = ctx->kt.ms[i];
  }
  return MSCKF_OK;
}

int msckf_feature_jacobian(msckf_ctx* ctx, int32_t f, int64_t feature_id, int32_t n_cam_ids, const int64_t* cam_ids, int32_t* rows_out,
                           double* H, double* r, int32_t n_state) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && cam_ids && rows_out && H && r && n_cam_ids > 0, "filter / pointers");
  CK(cudaSetDevice(ctx->device));
  StoreDump s;
  int rc = dump_store(ctx, f, s);
  if (rc) return rc;
  ARG(n_state == NI + 6 * s.ncam, "n_state must be 21 + 6 * (number of camera states)");
  const int slot = find_slot(s, feature_id);
  ARG(slot >= 0, "unknown feature id");
  std::vector<long long> ids(d.Nmax);
  if (s.ncam) CK(cudaMemcpy(ids.data(), d.cam_id + (size_t)f * d.Nmax, s.ncam * sizeof(long long), cudaMemcpyDeviceToHost));
  unsigned long long im = 0;
  for (int k = 0; k < n_cam_ids; ++k)
    for (int c = 0; c < s.ncam; ++c)
      if (ids[c] == cam_ids[k] && ((s.mask[slot] >> s.phys[c]) & 1ull)) im |= 1ull << s.phys[c];
  const int M = __builtin_popcountll(im);
  ARG(M >= 1, "landmark not observed by any of the given camera states");
  const int rows = 4 * M - 3;
  const long long need = jac_record_doubles(M);
  ARG(need <= d.pool_cap, "jacobian pool too small");
  // a one-item work list for filter f
  CK(cudaMemsetAsync(d.nitem, 0, d.F * sizeof(int), ctx->st));
  CK(cudaMemsetAsync(d.ncand, 0, d.F * sizeof(int), ctx->st));
  const size_t b = (size_t)f * d.Fcap;
  const int one = 1, zero = 0, dof = std::min(M, 99);
  const long long off0 = 0;
  CK(cudaMemcpyAsync(d.cand + b, &slot, sizeof(int), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.item_cand + b, &zero, sizeof(int), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.item_mask + b, &im, sizeof(im), cudaMemcpyHostToDevice, ctx->st));
  unsigned long long lm = 0;                 // the same mask in logical camera order, and the slot (k_plan_items' part)
  for (int c = 0; c < s.ncam; ++c) lm |= ((im >> s.phys[c]) & 1ull) << c;
  CK(cudaMemcpyAsync(d.item_lmask + b, &lm, sizeof(lm), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.item_slot + b, &slot, sizeof(int), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.item_rows + b, &rows, sizeof(int), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.item_dof + b, &dof, sizeof(int), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.item_off + b, &off0, sizeof(off0), cudaMemcpyHostToDevice, ctx->st));
  CK(cudaMemcpyAsync(d.nitem + f, &one, sizeof(int), cudaMemcpyHostToDevice, ctx->st));
  {                                          // a one-entry work list for the matching Jacobian kernel
    unsigned long long ctr[SC_COUNT] = {0};
    ctr[M <= JAC_SMALL_M ? SC_NJAC_SMALL + jac_class(M) : (M <= JAC_BIG_SPLIT_M ? SC_NJAC_BIG : SC_NJAC_BIG2)] = 1;
    const int2 entry = make_int2(f, 0);
    CK(cudaMemcpyAsync(d.stage_ctr, ctr, sizeof ctr, cudaMemcpyHostToDevice, ctx->st));
    CK(cudaMemcpyAsync(M <= JAC_SMALL_M ? d.wl_jac_small + (size_t)jac_class(M) * d.F * d.Fcap : (M <= JAC_BIG_SPLIT_M ? d.wl_jac_big : d.wl_jac_big2), &entry,
                       sizeof entry, cudaMemcpyHostToDevice, ctx->st));
    CK(cudaStreamSynchronize(ctx->st));      // the small host sources above live on this stack frame
  }
  launch_cam_prepare(d, ctx->st);
  launch_jacobian_gate(d, 0xF, 1, 1, ctx->st);
  double* dH = nullptr;
  CK(cudaMalloc((void**)&dH, ((size_t)rows * n_state + rows) * sizeof(double)));
  launch_export_item_dense(d, f, dH, dH + (size_t)rows * n_state, n_state, ctx->st);
  rc = sync_and_check(ctx);
  if (!rc) {
    cudaMemcpy(H, dH, (size_t)rows * n_state * sizeof(double), cudaMemcpyDeviceToHost);
    cudaMemcpy(r, dH + (size_t)rows * n_state, rows * sizeof(double), cudaMemcpyDeviceToHost);
    *rows_out = rows;
  }
  cudaFree(dH);
  cudaMemsetAsync(d.nitem, 0, d.F * sizeof(int), ctx->st);
  return rc;
}

int msckf_measurement_jacobian(msckf_ctx* ctx, int32_t f, int64_t cam_id, int64_t feature_id, double* H_x, double* H_f, double* r) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && H_x && H_f && r, "filter / pointers");
  CK(cudaSetDevice(ctx->device));
  StoreDump s;
  int rc = dump_store(ctx, f, s);
  if (rc) return rc;
  const int slot = find_slot(s, feature_id);
  ARG(slot >= 0, "unknown feature id");
  std::vector<long long> ids(d.Nmax);
  if (s.ncam) CK(cudaMemcpy(ids.data(), d.cam_id + (size_t)f * d.Nmax, s.ncam * sizeof(long long), cudaMemcpyDeviceToHost));
  int c = -1;
  for (int k = 0; k < s.ncam; ++k)
    if (ids[k] == cam_id) c = k;
  ARG(c >= 0 && ((s.mask[slot] >> s.phys[c]) & 1ull), "camera state unknown or does not observe the landmark");
  double* dout = nullptr;
  CK(cudaMalloc((void**)&dout, 40 * sizeof(double)));
  launch_cam_prepare(d, ctx->st);
  launch_measurement_jacobian(d, f, c, slot, dout, ctx->st);
  rc = sync_and_check(ctx);
  double out[40];
  if (!rc) {
    cudaMemcpy(out, dout, sizeof out, cudaMemcpyDeviceToHost);
    memcpy(H_x, out, 24 * sizeof(double));
    memcpy(H_f, out + 24, 12 * sizeof(double));
    memcpy(r, out + 36, 4 * sizeof(double));
  }
  cudaFree(dout);
  return rc;
}

int msckf_gating_test(msckf_ctx* ctx, int32_t f, const double* H, const double* r, int32_t m, int32_t n_state, int32_t dof,
                      double* gamma, int32_t* pass) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && H && r && m > 0 && dof >= 1 && dof <= 99 && n_state >= NI && n_state <= d.nmax, "filter / sizes / dof");
  CK(cudaSetDevice(ctx->device));
  double* buf = nullptr;
  const size_t nH = (size_t)m * n_state, nscr = nH + (size_t)m * m + m;
  CK(cudaMalloc((void**)&buf, (nH + m + nscr + 2) * sizeof(double)));
  CK(cudaMemcpy(buf, H, nH * sizeof(double), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(buf + nH, r, m * sizeof(double), cudaMemcpyHostToDevice));
  double* dg = buf + nH + m + nscr;
  int* dp = (int*)(dg + 1);
  launch_gate_dense(d, f, buf, buf + nH, m, n_state, dof, buf + nH + m, dg, dp, ctx->st);
  int rc = sync_and_check(ctx);
  if (!rc || rc == MSCKF_ERR_NUMERIC) {
    double g;
    int pz;
    cudaMemcpy(&g, dg, sizeof g, cudaMemcpyDeviceToHost);
    cudaMemcpy(&pz, dp, sizeof pz, cudaMemcpyDeviceToHost);
    if (gamma) *gamma = g;
    if (pass) *pass = pz;
  }
  cudaFree(buf);
  return rc;
}

// ------------------------------------------------------------------ gating test without any allocation
// The scratch of k_gate_dense (T1 = H P: m x n, S: m x m, y: m doubles) in a file-scope device array instead of a
// cudaMalloc per call.  Every GPU that loads the library holds the array until the process exits.
static __device__ __align__(256) double g_gate_scratch[MSCKF_GATE_NOALLOC_MAX_DOUBLES];
static double* g_gate_scratch_ptr[64];      // device address of the array, per device (the name above is a host-side stand-in)
static int g_gate_scratch_rc[64];
static DevOnce g_gate_scratch_once;

int msckf_gating_test_noalloc(msckf_ctx* ctx, int32_t f, const double* H_dev, const double* r_dev, int32_t m, int32_t n_state,
                              int32_t dof, double* gamma_dev, int32_t* pass_dev) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && H_dev && r_dev && gamma_dev && pass_dev && m > 0 && dof >= 1 && dof <= 99 && n_state >= NI && n_state <= d.nmax,
      "filter / pointers / sizes / dof");
  const unsigned long long need = (unsigned long long)m * n_state + (unsigned long long)m * m + (unsigned long long)m;
  if (need > (unsigned long long)MSCKF_GATE_NOALLOC_MAX_DOUBLES) {
    ctx->err = "msckf_gating_test_noalloc: m = " + std::to_string(m) + ", n_state = " + std::to_string(n_state) + " need " +
               std::to_string(need) + " doubles of scratch (m n_state + m m + m), the static array holds " +
               std::to_string((long long)MSCKF_GATE_NOALLOC_MAX_DOUBLES) + "; use msckf_gating_test";
    return MSCKF_ERR_CAPACITY;
  }
  ARG(ctx->device >= 0 && ctx->device < 64, "device index");
  CK(cudaSetDevice(ctx->device));
  g_gate_scratch_once.run([&] {          // once per device, with that device current
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_gate_scratch);
    g_gate_scratch_ptr[ctx->device] = (e == cudaSuccess) ? p : nullptr;
    g_gate_scratch_rc[ctx->device] = (int)e;
  });
  double* scratch = g_gate_scratch_ptr[ctx->device];
  if (!scratch) {
    ctx->err = std::string("cudaGetSymbolAddress(g_gate_scratch): ") + cudaGetErrorString((cudaError_t)g_gate_scratch_rc[ctx->device]);
    return MSCKF_ERR_CUDA;
  }
  // the kernel writes T1, S and y in full before it reads them: nothing in the array has to start at zero
  launch_gate_dense(d, f, H_dev, r_dev, m, n_state, dof, scratch, gamma_dev, pass_dev, ctx->st);
  CK(cudaGetLastError());
  return MSCKF_OK;
}

int msckf_measurement_update(msckf_ctx* ctx, int32_t f, const double* H, const double* r, int32_t m, int32_t n_state) {
  if (!ctx) return MSCKF_ERR_ARG;
  const Dev& d = ctx->d;
  ARG(f >= 0 && f < d.F && n_state >= NI && n_state <= d.nmax, "filter / size");
  if (m <= 0) return MSCKF_OK;                       // msckf.py:544-545
  ARG(H && r, "pointers");
  CK(cudaSetDevice(ctx->device));
  const long long mp = std::max(4, (m + 3) / 4 * 4);
  ARG(mp * (n_state + 1) <= d.Adense_elems, "H has more rows than max_rows allows");
  int ncam = 0;
  int rc = sync_and_check(ctx);
  if (rc) return rc;
  CK(cudaMemcpy(&ncam, d.ncam + f, sizeof(int), cudaMemcpyDeviceToHost));
  ARG(n_state == NI + 6 * ncam, "n_state must be 21 + 6 * (number of camera states)");
  double* buf = nullptr;
  const size_t nH = (size_t)m * n_state;
  CK(cudaMalloc((void**)&buf, (nH + m) * sizeof(double)));
  CK(cudaMemcpy(buf, H, nH * sizeof(double), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(buf + nH, r, m * sizeof(double), cudaMemcpyHostToDevice));
  launch_load_dense_H(d, f, buf, buf + nH, m, n_state, ctx->st);
  launch_qr_dense(d, f, (int)mp, ctx->st);
  launch_update(d, 0, ctx->st);
  rc = sync_and_check(ctx);
  cudaFree(buf);
  return rc;
}

int msckf_find_redundant_cam_states(msckf_ctx* ctx, int32_t f, int64_t* cam_ids) {
  if (!ctx) return MSCKF_ERR_ARG;
  ARG(f >= 0 && f < ctx->d.F && cam_ids, "filter / pointer");
  CK(cudaSetDevice(ctx->device));
  long long* dout = nullptr;
  CK(cudaMalloc((void**)&dout, 2 * sizeof(long long)));
  launch_find_redundant(ctx->d, f, dout, ctx->st);
  int rc = sync_and_check(ctx);
  if (!rc) cudaMemcpy(cam_ids, dout, 2 * sizeof(long long), cudaMemcpyDeviceToHost);
  cudaFree(dout);
  return rc;
}

}  // extern "C"
