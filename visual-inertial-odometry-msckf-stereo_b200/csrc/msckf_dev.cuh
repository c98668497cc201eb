// Device-side data model and small geometry helpers of the B200 MSCKF back end.
// All arithmetic is FP64 (the reference computes in numpy float64).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/msckf_b200.h"

namespace msckf {

// ------------------------------------------------------------------ constants
constexpr int IMU_STRIDE = MSCKF_IMU_STRIDE;
constexpr int CAM_STRIDE = MSCKF_CAM_STRIDE;   // q[4] p[3] q_null[4] p_null[3] pad[2]
constexpr int CAM_PRE = 32;                     // R0[9] R1[9] t0[3] t1[3] R(q_null) g [3] p_null[3] pad[2]
constexpr int NI = 21;                          // IMU error-state size
constexpr int NCNT = MSCKF_N_COUNTERS;

// feature flags
constexpr int FF_LIVE = 1;
constexpr int FF_INIT = 2;

// candidate status while a stage runs (final values are the MSCKF_FEAT_* log codes)
constexpr int CS_NEED_TRI = 100;   // lost / involved, not initialised yet
constexpr int CS_READY = 101;      // initialised: will be taken through jacobian + gate

// error bits accumulated on the device (one word per context)
constexpr unsigned ERRBIT_FEATURE_CAP = 1u;
constexpr unsigned ERRBIT_POOL = 2u;
constexpr unsigned ERRBIT_ROWS = 4u;
constexpr unsigned ERRBIT_NOT_SPD = 8u;
constexpr unsigned ERRBIT_CAM_CAP = 16u;
constexpr unsigned ERRBIT_DUP_ID = 32u;     // a feature id occurs twice in one message

// per-stage device counters (one memset per stage): pool bump pointer and work-list lengths
constexpr int SC_POOL = 0, SC_NTRI = 1, SC_NJAC_BIG = 2, SC_NUPD_BIG = 3, SC_NJAC_SMALL = 4 /* .. 7: one per size class */, SC_NQR = 12 /* .. 16: tensor-pipe QR classes */, SC_NUPD_DMMA = 17 /* .. 21: tensor-pipe update classes */, SC_NQR_BUILD = 22 /* wide / uncompressed stacks (k_qr_build) */, SC_NQR_TALL = 23 /* pruning stacks (k_qr_tall) */, SC_NJAC_BIG2 = 24 /* tracks longer than JAC_BIG_SPLIT_M */, SC_NLU = 25 /* gates to be redone with an LU solve */, SC_COUNT = 26;
// stacked-Jacobian compression classes: 0..3 = compressed (m > n_a) with n_a + 1 <= 16 / 32 / 48 / 64 columns (k_qr_dmma), 4 = general CTA kernel
constexpr int QR_DMMA_CLASSES = 5;
// one-CTA tensor-pipe update classes, aligned with the compression classes: max(k, n_a) + 1 <= 16 / 32 / 48 / 64
constexpr int UPD_DMMA_CLASSES = 5;
// size classes of the group-per-landmark Jacobian kernel (shared memory is sized per class): M <= 2, M = 3, M = 4, M <= 6
constexpr int JAC_CLASSES = 4;
__host__ __device__ inline int jac_class(int M) { return M <= 2 ? 0 : (M == 3 ? 1 : (M == 4 ? 2 : 3)); }
__host__ __device__ inline int jac_class_maxM(int c) { return c == 0 ? 2 : (c == 1 ? 3 : (c == 2 ? 4 : 6)); }
// landmarks with at most this many observations go through the warp-per-landmark Jacobian kernel
constexpr int JAC_SMALL_M = 6;
// longer tracks go through the CTA-per-landmark kernel, in two size classes split at this track length
constexpr int JAC_BIG_SPLIT_M = 16;

struct Params {
  double Qc[4];            // gyro, gyro-bias, acc, acc-bias noise variances (msckf.py:132-137)
  double obs_noise;        // config.py:76
  double P0[5];            // gyro_bias, velocity, acc_bias, extr. rotation, extr. translation (msckf.py:820-830)
  double pos_std_thr;      // config.py:64
  double tri_translation_thr, huber_eps, est_precision, init_damping;   // config.py:10-13
  int outer_max, inner_max;                                              // config.py:14-15
  double prune_rot_thr, prune_trans_thr, prune_track_thr;                // msckf.py:717
  double R01[9], t01[3];   // cam0 -> cam1 (config.py:110-114)
  double chi2[100];        // msckf.py:121-123
  int max_cam;             // config.py:57
  int row_cap;             // msckf.py:678
};

// Message pointers of the frame being processed (device addresses).  The kernels that consume the message read them
// from here when the frame runs as a replayed CUDA graph: the graph is static, only this block changes per frame.
struct FrameIn {
  const int* imu_off; const double* ts; const double* gyro; const double* acc;
  const long long* cam_ids; const int* feat_off; const long long* ids; const double* uv;
};

// Everything a kernel needs, passed by value.
struct Dev {
  int F, Nmax, nmax, ld, Fcap, Hs;
  int max_rows;            // row capacity of the stacked Jacobian per filter
  long long A_elems;       // elements of A per filter
  int Tk;                  // row capacity of T / S per filter (= 6*Nmax, or more for user H)
  const Params* prm;
  // ---- filter state
  double* imu;             // [F][IMU_STRIDE]
  long long* cam_id;       // [F][Nmax]   logical (insertion) order
  int* cam_phys;           // [F][Nmax]   physical observation slot of each camera state
  double* cam;             // [F][Nmax][CAM_STRIDE]
  int* ncam;               // [F]
  double* cam_pre;         // [F][Nmax][CAM_PRE] per-stage derived camera quantities (k_cam_prepare)
  unsigned long long* phys_used;   // [F] bit mask of physical slots in use
  int* obs_N;              // [F] camera-state count after this frame's augmentation (k_cam_slot), what k_observe sees
  int* obs_phys;           // [F] physical observation slot of the camera state being added (-1: window full)
  double* P;               // [2][F][ld*ld] two covariance buffers per filter, row-major, logical layout, n = 21+6*ncam
  double** Pcur;           // [F] the buffer that holds filter f's covariance now: the pruning update writes its result
                           // compacted (the two removed camera states dropped) into the other one and flips this pointer
  unsigned char* p_compacted;      // [F] set by that update: k_remove_cams then only compacts the camera table
  // ---- landmark store (map_server)
  long long* f_id;         // [F][Fcap]
  unsigned long long* f_mask;      // [F][Fcap] bit = physical cam slot
  int* f_flags;            // [F][Fcap]
  double* f_pos;           // [F][Fcap][3]
  double* f_obs;           // [F][Fcap][Nmax][4]
  int* ord;                // [F][Fcap] slots in insertion order
  int* ord_tmp;            // [F][Fcap]
  int* nlive;              // [F]
  int* free_stack;         // [F][Fcap]
  int* nfree;              // [F]
  long long* h_key;        // [F][Hs] open-addressing table rebuilt per frame
  int* h_val;              // [F][Hs]
  // ---- stage work lists
  int* cand;               // [F][Fcap] slots of lost / involved landmarks in map order
  int* cand_status;        // [F][Fcap]
  int* ncand;              // [F]
  int* item_cand;          // [F][Fcap] index into cand of the landmarks that reach the Jacobian
  unsigned long long* item_mask;   // [F][Fcap] physical-slot mask of the observations used
  int* item_rows;          // [F][Fcap] 4M-3
  long long* item_off;     // [F][Fcap] element offset in pool
  int* item_rowoff;        // [F][Fcap] row offset in the stacked matrix (-1 = not stacked)
  int* item_dof;           // [F][Fcap]
  double* item_gamma;      // [F][Fcap]
  int* item_pass;          // [F][Fcap]
  int* nitem;              // [F]
  unsigned long long* rm_mask;     // [F] physical mask of the camera states being pruned
  int* rm_idx;             // [F][2] logical indices, -1 when not pruning
  unsigned long long* item_lmask;  // [F][Fcap] the same mask in logical camera order (k_plan_items)
  int* item_slot;          // [F][Fcap] map slot of the item's landmark (= cand[item_cand])
  double* pool;            // compact per-landmark Jacobian records (58 M doubles each, bump allocated per stage)
  unsigned long long* stage_ctr;   // [SC_COUNT] pool bump pointer + work-list lengths, zeroed per stage
  long long pool_cap;      // elements
  int2* wl_tri;            // [F*Fcap] (filter, candidate) pairs that need triangulation
  int2* wl_jac_small;      // [JAC_CLASSES][F*Fcap] (filter, item) pairs, M <= JAC_SMALL_M, binned by size class
  int2* wl_jac_big;        // [F*Fcap] (filter, item) pairs, tracks of JAC_SMALL_M < M <= JAC_BIG_SPLIT_M observations
  int2* wl_jac_big2;       // [F*Fcap] (filter, item) pairs, longer tracks
  int2* wl_lu;             // [F*Fcap] (filter, item) pairs whose innovation covariance was not positive definite to rounding: the
                           // gate is redone with an LU solve as in the reference (msckf.py:609)
  double* lu_scratch;      // T1 | S | y | x of one such landmark
  double* jac_scratch;     // innovation covariance of tracks too long for shared memory, per CTA of the long-track kernel
  long long jac_scratch_stride;
  int jac_big_smem_doubles, jac_big_grid;
  int jac_rows_ok;         // short tracks go through the lane-per-row kernel (debug knob MSCKF_NO_JAC_ROWS)
  int qr_capcols;          // columns of a TSQR node: 6 Nmax + 1 rounded up to 8 (k_qr_wide)
  int qr_leaf_cap;         // leaves per filter at most
  int qr_levels;           // tree levels launched per stage: ceil(log2(qr_leaf_cap))
  double* Rnode;           // [F][qr_leaf_cap][Rnode_stride] triangular factors of the TSQR nodes, row-major, ld = qr_capcols
  long long Rnode_stride;
  double* A;               // [F][A_elems] compressed Jacobian H_thin | r_thin, column-major, ld = m_pad
  double* Adense;          // [Adense_elems] caller-supplied dense H of the measurement_update API (one filter at a time)
  long long Adense_elems;
  int* m_rows;             // [F] rows stacked
  int* m_pad;              // [F] leading dimension of A (k_rows rounded up to 4)
  int* n_act;              // [F] active columns (without r)
  int* k_rows;             // [F] rows of H_thin after compression
  int* colmap;             // [F][nmax] active column -> state index
  int* rowmap;             // [F][rowmap_cap] stacked row -> (item << 8 | row inside the item's record)
  int rowmap_cap;
  int* qr_cls;             // [F] compression class of this stage's stacked Jacobian (-1: nothing to do)
  int* wl_qr;              // [QR_DMMA_CLASSES][F] filters of the tensor-pipe QR classes
  int* wl_qrt;             // [F] filters whose pruning stack (n_a = 12, every landmark on the same two cameras) goes through k_qr_tall
  int qr_tall_ok;          // 0: debugging knob MSCKF_NO_QR_TALL, those stacks take k_qr_dmma class 0
  int* wl_qrb;             // [F] filters whose stack goes through k_qr_build (wider than 64 columns, or not compressed)
  int* upd_path;           // [F] 0 = no update, 2 = tiled kernels (queued in wl_upd_big), 5 + c = DMMA kernel of class c
  int* wl_upd_big;         // [F]
  int* wl_upd_dmma;        // [UPD_DMMA_CLASSES][F] filters of the tensor-pipe update classes
  int upd_dmma_ok[UPD_DMMA_CLASSES];   // class fits shared memory for this window size
  double* T;               // [F][Tk][ld+8]  H_thin P | r  -> W | y
  double* S;               // [F][Tk][Sld]  innovation covariance -> its upper Cholesky factor (wide updates)
  int Sld;                 // Tk rounded up to 8: rows are 64-byte aligned for the bulk copies
  double* TactT;           // [F][Tk][Sld]  (H_thin P[act,act])^T, the operand of S
  double* Uinv;            // [F][Uinv_stride] inverses of the 32 x 32 diagonal blocks of the factor
  long long Uinv_stride;
  double* dx;              // [F][nmax]
  // ---- logs / counters / errors
  long long* log_id;       // [2][F][Fcap]
  int* log_status;         // [2][F][Fcap]
  double* log_gamma;       // [2][F][Fcap]
  int* log_dof;            // [2][F][Fcap]
  int* log_n;              // [2][F]
  int* log_upd;            // [2][F][2] rows, cols of the update
  long long* log_rm;       // [F][2]
  long long* counters;     // [F][MSCKF_N_COUNTERS]
  unsigned* err;           // [1] error bits, followed by [7] ints of detail about the first numeric failure
  unsigned char* did_reset;        // [F]
  double* imu_pub;         // [F][IMU_STRIDE] snapshot for publish()
  FrameIn* frame_in;       // [1] device copy of the current frame's message pointers (graph replay)
  void* host_timer;        // KTimer* of the owning context (host only; never dereferenced on the device)
  void* host_lanes;        // Lanes* of the owning context (host only): auxiliary streams for independent launches of one step
};

__host__ __device__ inline int Tld(const Dev& d) { return d.ld + 8; }
// covariance of filter f
__device__ __forceinline__ double* cov_of(const Dev& d, int f) { return d.Pcur[f]; }
__device__ __forceinline__ double* cov_other(const Dev& d, int f) {
  double* b0 = d.P + (size_t)f * d.ld * d.ld;
  return d.Pcur[f] == b0 ? b0 + (size_t)d.F * d.ld * d.ld : b0;
}
// Pool record of a landmark that passed the gate: its projected Jacobian rows, row-major [4M-3][6M+1]
// (six columns per observation in logical camera order, last column = projected residual).
__host__ __device__ inline long long jac_record_doubles(int M) { return (long long)(4 * M - 3) * (6 * M + 1); }

// A gate whose Cholesky factorisation met a non-positive pivot is queued for k_gate_lu
__device__ __forceinline__ void queue_gate_lu(int2* wl_lu, unsigned long long* stage_ctr, int f, int k) {
  wl_lu[atomicAdd(stage_ctr + SC_NLU, 1ull)] = make_int2(f, k);
}

// records where the first non-positive-definite innovation covariance was met (where: 1 gate, 2 dense gate, 3 tiled update,
// 4 fused update), then raises the error bit
__device__ __forceinline__ void raise_not_spd(const Dev& d, int where, int f, int a, int b, int c) {
  if (atomicCAS(d.err + 1, 0u, (unsigned)where) == 0u) {
    d.err[2] = (unsigned)f; d.err[3] = (unsigned)a; d.err[4] = (unsigned)b; d.err[5] = (unsigned)c;
  }
  atomicOr(d.err, ERRBIT_NOT_SPD);
}

// ------------------------------------------------------------------ warp helpers
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------ 3x3 / quaternion algebra (utils.py:4-128)
__device__ __forceinline__ void mat3_mul(const double* A, const double* B, double* C) {   // C = A B
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
__device__ __forceinline__ void mat3_mul_nt(const double* A, const double* B, double* C) {  // C = A B^T
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[3 * j] + A[3 * i + 1] * B[3 * j + 1] + A[3 * i + 2] * B[3 * j + 2];
}
__device__ __forceinline__ void mat3_mul_tn(const double* A, const double* B, double* C) {  // C = A^T B
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
}
__device__ __forceinline__ void mat3_vec(const double* A, const double* x, double* y) {     // y = A x
#pragma unroll
  for (int i = 0; i < 3; ++i) y[i] = A[3 * i] * x[0] + A[3 * i + 1] * x[1] + A[3 * i + 2] * x[2];
}
__device__ __forceinline__ void mat3t_vec(const double* A, const double* x, double* y) {    // y = A^T x
#pragma unroll
  for (int i = 0; i < 3; ++i) y[i] = A[i] * x[0] + A[3 + i] * x[1] + A[6 + i] * x[2];
}
__device__ __forceinline__ void skew3(const double* v, double* S) {   // utils.py:4-12
  S[0] = 0; S[1] = -v[2]; S[2] = v[1];
  S[3] = v[2]; S[4] = 0; S[5] = -v[0];
  S[6] = -v[1]; S[7] = v[0]; S[8] = 0;
}
__device__ __forceinline__ void cross3(const double* a, const double* b, double* c) {   // skew(a) b
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double norm4(const double* q) { return sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]); }
__device__ __forceinline__ double norm3(const double* v) { return sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }

// R from a quaternion WITHOUT normalising, with an explicit vector for the skew term
// (jit_utils.py:86-104 re-uses a stale skew there).
__device__ __forceinline__ void rot_from_parts(const double* vec, double w, const double* skew_vec, double* R) {
  const double a = 2 * w * w - 1, b = 2 * w;
  R[0] = a + 2 * vec[0] * vec[0];              R[1] = b * skew_vec[2] + 2 * vec[0] * vec[1];  R[2] = -b * skew_vec[1] + 2 * vec[0] * vec[2];
  R[3] = -b * skew_vec[2] + 2 * vec[1] * vec[0]; R[4] = a + 2 * vec[1] * vec[1];              R[5] = b * skew_vec[0] + 2 * vec[1] * vec[2];
  R[6] = b * skew_vec[1] + 2 * vec[2] * vec[0];  R[7] = -b * skew_vec[0] + 2 * vec[2] * vec[1]; R[8] = a + 2 * vec[2] * vec[2];
}
// utils.py:14-27
__device__ __forceinline__ void to_rotation(const double* q_in, double* R) {
  const double n = norm4(q_in);
  const double q[4] = {q_in[0] / n, q_in[1] / n, q_in[2] / n, q_in[3] / n};
  rot_from_parts(q, q[3], q, R);
}
// utils.py:29-53
__device__ __forceinline__ void to_quaternion(const double* R, double* q) {
  double t;
  if (R[8] < 0) {
    if (R[0] > R[4]) {
      t = 1 + R[0] - R[4] - R[8];
      q[0] = t; q[1] = R[1] + R[3]; q[2] = R[6] + R[2]; q[3] = R[5] - R[7];
    } else {
      t = 1 - R[0] + R[4] - R[8];
      q[0] = R[1] + R[3]; q[1] = t; q[2] = R[7] + R[5]; q[3] = R[6] - R[2];
    }
  } else {
    if (R[0] < -R[4]) {
      t = 1 - R[0] - R[4] + R[8];
      q[0] = R[2] + R[6]; q[1] = R[7] + R[5]; q[2] = t; q[3] = R[1] - R[3];
    } else {
      t = 1 + R[0] + R[4] + R[8];
      q[0] = R[5] - R[7]; q[1] = R[6] - R[2]; q[2] = R[1] - R[3]; q[3] = t;
    }
  }
  const double n = norm4(q);
  q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}
// utils.py:67-82
__device__ __forceinline__ void quat_mul(const double* q1, const double* q2, double* out) {
  const double n1 = norm4(q1), n2 = norm4(q2);
  const double a[4] = {q1[0] / n1, q1[1] / n1, q1[2] / n1, q1[3] / n1};
  const double b[4] = {q2[0] / n2, q2[1] / n2, q2[2] / n2, q2[3] / n2};
  double q[4];
  q[0] = a[3] * b[0] + a[2] * b[1] - a[1] * b[2] + a[0] * b[3];
  q[1] = -a[2] * b[0] + a[3] * b[1] + a[0] * b[2] + a[1] * b[3];
  q[2] = a[1] * b[0] - a[0] * b[1] + a[3] * b[2] + a[2] * b[3];
  q[3] = -a[0] * b[0] - a[1] * b[1] - a[2] * b[2] + a[3] * b[3];
  const double n = norm4(q);
  out[0] = q[0] / n; out[1] = q[1] / n; out[2] = q[2] / n; out[3] = q[3] / n;
}
// utils.py:85-101
__device__ __forceinline__ void small_angle_quat(const double* dtheta, double* q) {
  const double d[3] = {dtheta[0] / 2., dtheta[1] / 2., dtheta[2] / 2.};
  const double s = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
  if (s <= 1) {
    q[0] = d[0]; q[1] = d[1]; q[2] = d[2]; q[3] = sqrt(1 - s);
  } else {
    const double n = sqrt(1 + s);
    q[0] = d[0] / n; q[1] = d[1] / n; q[2] = d[2] / n; q[3] = 1. / n;
  }
}

// state injection of the correction dx (msckf.py:565-595), cooperatively by `nthreads` threads of one CTA
__device__ __forceinline__ void inject_state(const Dev& d, int f, const double* dx, int tid, int nthreads) {
  const int N = d.ncam[f];
  if (tid == 0) {
    double* imu = d.imu + (size_t)f * IMU_STRIDE;
    double dq[4], q[4];
    small_angle_quat(dx, dq);
    quat_mul(dq, imu + MSCKF_IMU_Q, q);
    for (int i = 0; i < 4; ++i) imu[MSCKF_IMU_Q + i] = q[i];
    for (int i = 0; i < 3; ++i) {
      imu[MSCKF_IMU_BG + i] += dx[3 + i];
      imu[MSCKF_IMU_V + i] += dx[6 + i];
      imu[MSCKF_IMU_BA + i] += dx[9 + i];
      imu[MSCKF_IMU_P + i] += dx[12 + i];
    }
    if (imu[MSCKF_IMU_NULL_ALIASED] != 0.0) {       // the reference's "+=" moves the aliased null members too
      for (int i = 0; i < 3; ++i) {
        imu[MSCKF_IMU_V_NULL + i] = imu[MSCKF_IMU_V + i];
        imu[MSCKF_IMU_P_NULL + i] = imu[MSCKF_IMU_P + i];
      }
    }
    double dqe[4], Re[9], Rn[9];
    small_angle_quat(dx + 15, dqe);
    to_rotation(dqe, Re);
    mat3_mul(Re, imu + MSCKF_IMU_R_IC, Rn);
    for (int i = 0; i < 9; ++i) imu[MSCKF_IMU_R_IC + i] = Rn[i];
    for (int i = 0; i < 3; ++i) imu[MSCKF_IMU_T_CI + i] += dx[18 + i];
  }
  for (int c = tid; c < N; c += nthreads) {
    double* cs = d.cam + ((size_t)f * d.Nmax + c) * CAM_STRIDE;
    const double* dc = dx + NI + 6 * c;
    double dq[4], q[4];
    small_angle_quat(dc, dq);
    quat_mul(dq, cs, q);
    for (int i = 0; i < 4; ++i) cs[i] = q[i];
    for (int i = 0; i < 3; ++i) {
      cs[4 + i] += dc[3 + i];
      cs[11 + i] = cs[4 + i];      // position_null aliases _vio_position__ (msckf.py:400, 595)
    }
  }
}

}  // namespace msckf
