/*
 * msckf_b200.h -- C ABI of libmsckf_b200.so, the B200 (sm_100a) MSCKF back end.
 *
 * This is the drop-in boundary for the hot path of
 *   NonStopEagle137/Visual-Inertial-Odometry-MSCKF-Stereo  MSCKF/msckf.py, feature.py, jit_utils.py
 * A context owns a BATCH of independent filters (Monte-Carlo seeds / sequences).  All filter
 * state lives in device memory: IMU state, the camera-state window, the covariance P and the
 * landmark/observation store (the reference's map_server).  The host only feeds messages and
 * reads results.  Plain pointers and sizes only; every array is contiguous, row-major
 * ("C order"), float64 / int32 / int64 as declared.  Every call returns 0 on success or a
 * negative msckf_status; msckf_last_error() gives the text.  There is no CPU fallback: every
 * entry point fails with MSCKF_ERR_CUDA when no sm_100 device is usable.
 *
 * Threading: one CUDA stream per context; calls on one context must be serialised by the
 * caller (the reference drives its filter from a single thread, vio.py:56-65).
 *
 * Error-state layout (msckf.py:569-595): theta(0:3) b_g(3:6) v(6:9) b_a(9:12) p(12:15)
 * theta_ext(15:18) t_ext(18:21), then [theta_c(3) p_c(3)] per camera state in insertion order.
 * Quaternions are JPL [x y z w].
 */
#ifndef MSCKF_B200_H
#define MSCKF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct msckf_ctx msckf_ctx;

enum msckf_status {
  MSCKF_OK = 0,
  MSCKF_ERR_ARG = -1,       /* bad argument */
  MSCKF_ERR_CUDA = -2,      /* CUDA runtime error / no device */
  MSCKF_ERR_CAPACITY = -3,  /* feature store, Jacobian pool or stacked-H row capacity exceeded */
  MSCKF_ERR_NUMERIC = -4    /* innovation covariance of a measurement update not positive definite (reference: LinAlgError) */
};

/* Layout of the per-filter IMU block used by msckf_get_imu_state / msckf_set_imu_state
 * (IMUState, msckf.py:16-59; gravity is per filter here, a class global in the reference). */
enum {
  MSCKF_IMU_Q = 0,        /* orientation[4]                       */
  MSCKF_IMU_BG = 4,       /* _vio_gyro_bias__[3]                  */
  MSCKF_IMU_V = 7,        /* velocity[3]                          */
  MSCKF_IMU_BA = 10,      /* acc_bias[3]                          */
  MSCKF_IMU_P = 13,       /* _vio_position__[3]                   */
  MSCKF_IMU_Q_NULL = 16,  /* orientation_null[4]                  */
  MSCKF_IMU_V_NULL = 20,  /* velocity_null[3]                     */
  MSCKF_IMU_P_NULL = 23,  /* position_null[3]                     */
  MSCKF_IMU_R_IC = 26,    /* R_imu_cam0[9]                        */
  MSCKF_IMU_T_CI = 35,    /* t_cam0_imu[3]                        */
  MSCKF_IMU_GRAVITY = 38, /* IMUState._vio_gravity__[3]           */
  MSCKF_IMU_TIME = 41,    /* timestamp                            */
  MSCKF_IMU_NULL_ALIASED = 42, /* 1.0 once process_model ran: velocity_null / position_null then
                                  alias the live arrays (msckf.py:366-368) and move with "+="    */
  MSCKF_IMU_TRACKING_RATE = 43,
  MSCKF_IMU_STRIDE = 48
};

/* Per camera state: orientation[4] position[3] orientation_null[4] position_null[3] pad[2]
 * (CAMState, msckf.py:62-86). */
enum { MSCKF_CAM_STRIDE = 16 };

/* Status of a feature in a stage log (msckf_get_stage_log). */
enum {
  MSCKF_FEAT_FEW_OBS = 1,     /* lost with < 3 observations -> removed (msckf.py:627-629)       */
  MSCKF_FEAT_TRI_FAILED = 2,  /* check_motion / initialize_position failed (msckf.py:634-643)   */
  MSCKF_FEAT_GATE_PASS = 3,   /* Jacobian stacked                                               */
  MSCKF_FEAT_GATE_FAIL = 4,   /* rejected by gating_test                                        */
  MSCKF_FEAT_AFTER_CAP = 5    /* after the row-cap break (msckf.py:678): never gated            */
};

/* Per-filter counters since creation (msckf_get_counters).  The *_FLOPS / *_BYTES entries accumulate the
 * ALGORITHMIC work of the corresponding kernel family with the formulas of SURVEY.md section 8d (DESIGN.md
 * restates them); bench.py divides them by measured kernel time for the roofline figures. */
enum {
  MSCKF_CNT_FEATURE_UPDATES = 0, /* landmarks taken through feature_jacobian + gating_test          */
  MSCKF_CNT_TRIANGULATIONS = 1,  /* initialize_position calls                                        */
  MSCKF_CNT_UPDATES = 2,         /* measurement updates with at least one row                        */
  MSCKF_CNT_FRAMES = 3,          /* feature_callback / online_reset calls                            */
  MSCKF_CNT_TRI_OBS = 4,         /* sum of M over triangulated landmarks                             */
  MSCKF_CNT_TRI_SOLVES = 5,      /* damped 3x3 solves                                                */
  MSCKF_CNT_JAC_OBS = 6,         /* sum of M over landmarks that reached the Jacobian                */
  MSCKF_CNT_K2_FLOPS = 7, MSCKF_CNT_K2_BYTES = 8,   /* Jacobian + null space + gate                 */
  MSCKF_CNT_QR_FLOPS = 9, MSCKF_CNT_QR_BYTES = 10,  /* stacked-H compression                        */
  MSCKF_CNT_K4_FLOPS = 11, MSCKF_CNT_K4_BYTES = 12, /* gain + correction + covariance               */
  MSCKF_CNT_IMU_STEPS = 13,      /* IMU samples integrated                                           */
  MSCKF_CNT_GATE_NOT_SPD = 14,   /* landmarks rejected because rounding made their innovation covariance indefinite */
  MSCKF_CNT_K4_EXEC_FLOPS = 15,  /* flops the update kernels actually execute: triangular H_thin, active columns only,
                                    upper triangles of S and of the downdate (the K4 contract counts dense GEMMs)    */
  MSCKF_N_COUNTERS = 16
};

typedef struct msckf_config {
  int32_t n_filters;          /* batch size F >= 1                                              */
  int32_t max_cam_states;     /* ConfigEuRoC._vio_max_cam_state_size__ (config.py:57), 4..64    */
  int32_t feature_capacity;   /* live landmarks per filter (map_server size bound)              */
  int32_t max_rows;           /* stacked-Jacobian row capacity per filter; 0 = row_cap + 4*max_cam + 8 */
  int32_t row_cap;            /* 1500 in the reference (msckf.py:678); <= 0 lifts the cap       */
  int32_t device;             /* CUDA device ordinal                                            */
  int32_t outer_loop_max_iteration, inner_loop_max_iteration;      /* config.py:14-15           */
  int64_t jacobian_pool_bytes;/* scratch for per-feature Jacobians; 0 = automatic               */
  double gyro_noise, gyro_bias_noise, acc_noise, acc_bias_noise;   /* variances, config.py:72-75 */
  double observation_noise;                                        /* config.py:76              */
  double gyro_bias_cov, velocity_cov, acc_bias_cov;                /* config.py:84-86           */
  double extrinsic_rotation_cov, extrinsic_translation_cov;        /* config.py:87-88           */
  double position_std_threshold;                                   /* config.py:64              */
  double translation_threshold, huber_epsilon, estimation_precision, initial_damping; /* config.py:10-13 */
  double prune_rotation_threshold, prune_translation_threshold, prune_tracking_rate_threshold; /* msckf.py:717 literals */
  double R_cam0_cam1[9], t_cam0_cam1[3];                           /* config.py:110-114         */
  double chi2_table[100];     /* [dof] = chi2.ppf(0.05, dof), dof 1..99 (msckf.py:121-123)      */
} msckf_config;

/* ---- life cycle (replaces MSCKF.__init__, msckf.py:105-164) */
int msckf_create(const msckf_config* cfg, msckf_ctx** out);
void msckf_destroy(msckf_ctx* ctx);
const char* msckf_last_error(const msckf_ctx* ctx);   /* ctx may be NULL for create errors */
int msckf_sync(msckf_ctx* ctx);                       /* waits for the stream; surfaces device-side errors */

/* Page-locked host memory for message buffers.  Arrays passed to msckf_feature_callback* that lie in page-locked
 * memory (from here, cudaHostRegister or torch pin_memory) are copied to the device directly, without the
 * intermediate copy into the library's own staging arena. */
int msckf_host_alloc(size_t bytes, void** out);
int msckf_host_free(void* p);

/* ---- whole-filter state I/O (snapshots for parity harnesses; StateServer msckf.py:89-100) */
int msckf_set_imu_state(msckf_ctx* ctx, int32_t filter, const double* imu /*[MSCKF_IMU_STRIDE]*/);
int msckf_get_imu_state(msckf_ctx* ctx, int32_t filter, double* imu /*[MSCKF_IMU_STRIDE]*/);
int msckf_get_imu_states(msckf_ctx* ctx, double* imu /*[F][MSCKF_IMU_STRIDE]*/);
int msckf_get_cam_states(msckf_ctx* ctx, int32_t filter, int32_t* n_cam, int64_t* ids /*[max_cam]*/,
                         double* cams /*[max_cam][MSCKF_CAM_STRIDE]*/);
int msckf_set_cam_states(msckf_ctx* ctx, int32_t filter, int32_t n_cam, const int64_t* ids, const double* cams);
int msckf_get_cov(msckf_ctx* ctx, int32_t filter, double* P /*[n][n], n = 21+6*n_cam*/, int32_t n);
int msckf_set_cov(msckf_ctx* ctx, int32_t filter, const double* P, int32_t n);
/* map_server view: live landmarks in insertion order. obs_mask bit j = observed by the j-th camera
 * state (insertion order); obs is [n][max_cam][4]. Any output pointer may be NULL. */
int msckf_get_features(msckf_ctx* ctx, int32_t filter, int32_t* n_features, int64_t* ids, uint64_t* obs_mask,
                       uint8_t* is_initialized, double* position /*[n][3]*/, double* obs);
/* Replaces the whole map of one filter (snapshot restore; the inverse of msckf_get_features, same layouts).
 * The camera states must have been set first.  is_initialized / position may be NULL (false / zeros). */
int msckf_set_features(msckf_ctx* ctx, int32_t filter, int32_t n_features, const int64_t* ids, const uint64_t* obs_mask,
                       const uint8_t* is_initialized, const double* position /*[n][3]*/, const double* obs /*[n][max_cam][4]*/);
/* Overwrites position / is_initialized of one live landmark (tests; Feature._vio_position__). */
int msckf_set_feature_position(msckf_ctx* ctx, int32_t filter, int64_t feature_id, const double* position,
                               int32_t is_initialized);

/* ---- pipeline stages, each batched over all F filters ------------------------------------ */

/* batch_imu_processing + process_model + predict_new_state (msckf.py:262-380, jit_utils.py:6-135).
 * The host selects the messages (state time <= ts <= image time); filter f integrates samples
 * [imu_off[f], imu_off[f+1]). */
int msckf_propagate(msckf_ctx* ctx, const int32_t* imu_off /*[F+1]*/, const double* ts,
                    const double* gyro /*[.][3]*/, const double* acc /*[.][3]*/);

/* state_augmentation (msckf.py:385-407, jit_utils.py:137-167): appends camera state cam_ids[f]. */
int msckf_augment(msckf_ctx* ctx, const int64_t* cam_ids /*[F]*/);

/* add_feature_observations (msckf.py:409-427): filter f receives features [feat_off[f], feat_off[f+1]);
 * ids must be unique inside one message. uv is [.][4] = u0 v0 u1 v1 (normalised). */
int msckf_add_observations(msckf_ctx* ctx, const int32_t* feat_off /*[F+1]*/, const int64_t* ids,
                           const double* uv);

/* remove_lost_features (msckf.py:616-689) incl. check_motion / initialize_position
 * (feature.py:124-295), feature_jacobian (msckf.py:500-541), gating_test (:606-614) and
 * measurement_update (:543-604). */
int msckf_remove_lost_features(msckf_ctx* ctx);

/* prune_cam_state_buffer (msckf.py:730-818) incl. find_redundant_cam_states (:691-727). */
int msckf_prune_cam_states(msckf_ctx* ctx);

/* online_reset (msckf.py:859-886); did_reset[f] (may be NULL) = 1 when filter f was reset. */
int msckf_online_reset(msckf_ctx* ctx, uint8_t* did_reset /*[F]*/);

/* reset (msckf.py:832-857): back to the initial covariance, empty window and map. */
int msckf_reset(msckf_ctx* ctx, int32_t filter);

/* One full feature_callback (msckf.py:180-233) for every filter: propagate, augment, observe,
 * remove lost, prune, snapshot for publish, online reset.  imu_out (may be NULL) receives the
 * IMU blocks as they were BEFORE online_reset (what publish() reads); did_reset may be NULL.
 * Returns after the results are in host memory. */
int msckf_feature_callback(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro,
                           const double* acc, const int64_t* cam_ids, const int32_t* feat_off,
                           const int64_t* ids, const double* uv, double* imu_out /*[F][MSCKF_IMU_STRIDE]*/,
                           uint8_t* did_reset /*[F]*/);
/* Same, but only enqueues the work; msckf_collect waits for the OLDEST frame in flight and fetches its results.
 * Up to two frames may be in flight: the host->device copy of frame k+1 (on a copy stream) overlaps the kernels
 * of frame k.  Pageable inputs are copied to pinned staging before returning; page-locked inputs are read by the
 * DMA engine and must stay untouched until that frame has been collected. */
int msckf_feature_callback_async(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro,
                                 const double* acc, const int64_t* cam_ids, const int32_t* feat_off,
                                 const int64_t* ids, const double* uv);
int msckf_collect(msckf_ctx* ctx, double* imu_out, uint8_t* did_reset);
/* Same stages with every input already in DEVICE memory (same layouts; the pointers are device pointers
 * that must stay valid until the work has run).  Nothing is copied and nothing is waited for: frames can be
 * enqueued back to back.  imu_out_dev (device, [F][MSCKF_IMU_STRIDE], may be NULL) receives the publish()
 * snapshot of this frame.  Errors surface at the next msckf_sync. */
int msckf_feature_callback_dev(msckf_ctx* ctx, const int32_t* imu_off, const double* ts, const double* gyro,
                               const double* acc, const int64_t* cam_ids, const int32_t* feat_off, const int64_t* ids,
                               const double* uv, double* imu_out_dev);

/* ---- per-stage decision logs (bit-exact outputs: ids, gate decisions, removed cam ids) ---- */
/* stage: 0 = last remove_lost_features, 1 = last prune_cam_states.  Entries are in map order.
 * removed_cam_ids is only written for stage 1 ([2], -1 when no pruning happened). */
int msckf_get_stage_log(msckf_ctx* ctx, int32_t filter, int32_t stage, int32_t* n_entries,
                        int64_t* feature_ids, int32_t* status, double* gamma, int32_t* dof,
                        int32_t* update_rows, int32_t* update_cols, int64_t* removed_cam_ids);
/* Counters since creation, per filter: counters[F][MSCKF_N_COUNTERS] (MSCKF_CNT_* above). */
int msckf_get_counters(msckf_ctx* ctx, int64_t* counters);
/* Shapes of the most recent measurement update of every filter (the last stage that ran): shapes[F][4] =
 * stacked rows m, active columns n_a, rows after compression k = min(m, n_a) (msckf.py:549-556), and the kernel
 * path taken (0 none, 2 tiled kernels, 5 + c one-CTA tensor-pipe class c).  Diagnostic: tests and profiling. */
int msckf_get_update_shapes(msckf_ctx* ctx, int32_t* shapes);

/* ---- directly callable pieces of the reference API ---------------------------------------- */

/* Feature.check_motion + initialize_position for n landmarks given explicit camera poses
 * (feature.py:124-295).  Landmark i owns observations [obs_off[i], obs_off[i+1]); each has
 * uv[4] and the cam0 pose of its camera state: cam_q[4] (world->cam, JPL) and cam_p[3].
 * out: position[n][3] (written even when invalid, like the reference), valid[n] (is_initialized),
 * lm_solves[n] (may be NULL): number of damped 3x3 solves taken. */
int msckf_triangulate(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q,
                      const double* cam_p, double* position, uint8_t* valid, int32_t* lm_solves);

/* Feature.initialize_position (feature.py:167-295) alone, i.e. msckf_triangulate without the check_motion test. */
int msckf_initialize_position(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q,
                              const double* cam_p, double* position, uint8_t* valid, int32_t* lm_solves);

/* Feature.check_motion (feature.py:124-165) alone: moved[i] = 1 when the translation between the first and
 * the last observing camera, orthogonal to the first bearing, exceeds translation_threshold. */
int msckf_check_motion(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q,
                       const double* cam_p, uint8_t* moved);

/* Feature.generate_initial_guess (feature.py:99-122) alone, between the first and the last cam0 view of each
 * landmark: guess[n][3] in the frame of the first cam0 (what initialize_position starts from, feature.py:216-218). */
int msckf_initial_guess(msckf_ctx* ctx, int32_t n, const int32_t* obs_off, const double* uv, const double* cam_q,
                        const double* cam_p, double* guess);

/* feature_jacobian (msckf.py:500-541) of a stored landmark over the given camera-state ids:
 * H is [(4M-3)][n_state] dense, r [(4M-3)].  The null-space basis is the Householder one, so H and
 * r equal the reference's up to an orthogonal row transform (H^T H, H^T r and gamma are equal). */
int msckf_feature_jacobian(msckf_ctx* ctx, int32_t filter, int64_t feature_id, int32_t n_cam_ids,
                           const int64_t* cam_ids, int32_t* rows, double* H, double* r, int32_t n_state);

/* measurement_jacobian (msckf.py:429-498): H_x[4][6], H_f[4][3], r[4] for one (cam state, feature). */
int msckf_measurement_jacobian(msckf_ctx* ctx, int32_t filter, int64_t cam_id, int64_t feature_id,
                               double* H_x, double* H_f, double* r);

/* gating_test (msckf.py:606-614) on a caller-supplied dense H [m][n_state], r [m]. */
int msckf_gating_test(msckf_ctx* ctx, int32_t filter, const double* H, const double* r, int32_t m,
                      int32_t n_state, int32_t dof, double* gamma, int32_t* pass);

/* The gating test for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same kernel and arithmetic as msckf_gating_test, but H_dev [m][n_state], r_dev [m], gamma_dev [1] and pass_dev [1] are
 * DEVICE pointers, the kernel is enqueued on the context's stream (results are valid after msckf_sync or any later
 * blocking call) and its scratch (m n_state + m m + m doubles) lives in a static device array of
 * MSCKF_GATE_NOALLOC_MAX_DOUBLES doubles (534 KiB, held by every GPU that loads the library until the process exits);
 * the call performs no cudaMalloc / cudaFree.  Larger shapes return MSCKF_ERR_CAPACITY (text in msckf_last_error).
 * All calls on a device share the array: calls of different contexts (streams) that overlap on the GPU race; calls on one
 * context are ordered by its stream.  The array keeps the previous call's contents; the kernel overwrites every entry
 * it reads. */
#define MSCKF_GATE_NOALLOC_MAX_DOUBLES (128 * 405 + 128 * 128 + 128)   /* e.g. m = 128 rows at n_state = 405 */
int msckf_gating_test_noalloc(msckf_ctx* ctx, int32_t filter, const double* H_dev, const double* r_dev, int32_t m,
                              int32_t n_state, int32_t dof, double* gamma_dev, int32_t* pass_dev);

/* measurement_update (msckf.py:543-604) with a caller-supplied dense H [m][n_state], r [m]. */
int msckf_measurement_update(msckf_ctx* ctx, int32_t filter, const double* H, const double* r, int32_t m,
                             int32_t n_state);

/* find_redundant_cam_states (msckf.py:691-727): the two camera-state ids that would be removed. */
int msckf_find_redundant_cam_states(msckf_ctx* ctx, int32_t filter, int64_t* cam_ids /*[2]*/);

/* ---- the step in front of the filter: point undistortion of the feature message (context free) ---- */
/* undistort_points (image.py:640-674; cv2.undistortPoints / cv2.fisheye.undistortPoints): pts_in [n][2] pixel
 * coordinates -> pts_out [n][2].  intrinsics = fx fy cx cy; model 0 = radtan (k1 k2 p1 p2), 1 = equidistant (k1..k4);
 * rectification [9] row-major or NULL (identity); new_intrinsics [4] or NULL (1 1 0 0: normalised coordinates).
 * As in the reference's positional cv2.fisheye call, the equidistant model ignores the rectification. */
int msckf_undistort_points(int32_t device, int32_t n, const double* pts_in, const double* intrinsics, int32_t model,
                           const double* dist, const double* rectification, const double* new_intrinsics,
                           double* pts_out);
/* publish (image.py:406-438): raw pixel tracks of both cameras -> the rows [u0 v0 u1 v1] of the feature message,
 * uv_out [n][4], in the layout msckf_add_observations / msckf_feature_callback take. */
int msckf_undistort_stereo(int32_t device, int32_t n, const double* cam0_pts, const double* cam1_pts,
                           const double* intr0, int32_t model0, const double* dist0, const double* intr1,
                           int32_t model1, const double* dist1, double* uv_out);

/* ---- the step before that: synthetic message generator (replaces the host generator of the benchmark) ---- */
/* One lock-step frame for n_filters Monte-Carlo filters, written in DEVICE memory in the layout msckf_feature_callback_dev
 * takes.  Filter f adds the noise stream of seed (seed0 + f) to the frame's clean structure (host arrays: the n_imu samples
 * ts / gyro_clean / acc_clean [.][3] selected for this frame, the n_features tracked ids and their noise-free normalised
 * coordinates uv_clean [.][4]); every random number is a pure function of (seed, sample index or (frame, feature id)), so
 * any filter or frame can be regenerated alone.  Noise model of SURVEY 8d: white IMU noise (discrete sigmas) on top of
 * constant biases; pixel noise sigma_z plus, for outlier_frac of the observations, a gross error of outlier_sigma.
 * sample0 = global index of the frame's first IMU sample.  Outputs (device): imu_off [F+1], ts_out [F*n_imu],
 * gyro / acc [F*n_imu][3], cam_ids [F] (= cam_id), feat_off [F+1], ids_out [F*n_features], uv [F*n_features][4]. */
int msckf_synth_frame(int32_t device, int32_t n_filters, uint64_t seed0, int64_t frame, int32_t n_imu, int64_t sample0,
                      const double* ts, const double* gyro_clean, const double* acc_clean, const double* gyro_bias /*[3]*/,
                      const double* acc_bias /*[3]*/, double sigma_gyro, double sigma_acc, int32_t n_features,
                      const int64_t* ids, const double* uv_clean, double sigma_z, double outlier_frac, double outlier_sigma,
                      int64_t cam_id, int32_t* imu_off, double* ts_out, double* gyro, double* acc, int64_t* cam_ids,
                      int32_t* feat_off, int64_t* ids_out, double* uv);

/* Timing of the last msckf_feature_callback*: device milliseconds per stage (CUDA events on the
 * context's stream): [0] propagate+augment [1] observe [2] remove_lost [3] prune [4] total. */
int msckf_last_timing(msckf_ctx* ctx, float* ms /*[5]*/);
/* Enable (1) / disable (0) per-stage event timing (off by default: events cost launches). */
int msckf_set_timing(msckf_ctx* ctx, int32_t enabled);

/* Number of CUDA kernels of this library enqueued on the context so far (bench.py: gpu_launches). */
int msckf_kernel_launches(msckf_ctx* ctx, int64_t* n);
/* Per-kernel accounting.  msckf_kernel_count() kernels, named by msckf_kernel_name(i).  With
 * msckf_set_kernel_timing(ctx, 1) every launch is bracketed by two CUDA events on the context's stream;
 * msckf_kernel_times waits for the stream and returns, per kernel, the launches counted since creation,
 * the number of timed launches and their total device milliseconds (arrays of msckf_kernel_count()). */
int msckf_kernel_count(void);
const char* msckf_kernel_name(int32_t i);
int msckf_set_kernel_timing(msckf_ctx* ctx, int32_t enabled);
int msckf_kernel_times(msckf_ctx* ctx, int64_t* launches, int64_t* timed_launches, double* total_ms);

/* Measures the FP64 peaks of the device with two register-resident microbenchmarks: DFMA chains (vector
 * pipe) and mma.sync m8n8k4 f64 (DMMA, the FP64 tensor op), in TFLOP/s.  Roofline denominators for the FP64
 * kernels; independent of any context. */
int msckf_fp64_peak(int32_t device, double* dfma_tflops, double* dmma_tflops);

/* Generic event pair on the context's stream for callers that time several calls from outside:
 * msckf_mark(ctx, 0) ... msckf_mark(ctx, 1); msckf_mark_elapsed waits for mark 1 and returns the milliseconds. */
int msckf_mark(msckf_ctx* ctx, int32_t which);
int msckf_mark_elapsed(msckf_ctx* ctx, float* ms);
/* The same across two contexts of one device (a batch split over several contexts, each with its own stream):
 * waits for mark 1 of `last` and returns the milliseconds since mark 0 of `first`. */
int msckf_mark_elapsed_between(msckf_ctx* first, msckf_ctx* last, float* ms);

/* Library build info, e.g. "msckf_b200 sm_100a cuda 12.9". */
const char* msckf_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MSCKF_B200_H */
