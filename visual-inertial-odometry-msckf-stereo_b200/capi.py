"""ctypes binding of ``libmsckf_b200.so`` (C ABI in ``include/msckf_b200.h``).

This is the only route from Python to the CUDA kernels.  There is no CPU fallback: if the
shared library is missing, or the machine has no sm_100-class GPU, construction raises
``MsckfCudaError``.  Status codes map to exceptions the way the reference's failures surface:
a non-SPD innovation covariance (``np.linalg.LinAlgError`` in the reference, msckf.py:562/609)
raises ``MsckfNumericError`` which derives from ``np.linalg.LinAlgError``.
"""
import ctypes as C
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, 'libmsckf_b200.so')

# enum mirrors (include/msckf_b200.h)
IMU_Q, IMU_BG, IMU_V, IMU_BA, IMU_P = 0, 4, 7, 10, 13
IMU_Q_NULL, IMU_V_NULL, IMU_P_NULL = 16, 20, 23
IMU_R_IC, IMU_T_CI, IMU_GRAVITY, IMU_TIME = 26, 35, 38, 41
IMU_NULL_ALIASED, IMU_TRACKING_RATE, IMU_STRIDE = 42, 43, 48
CAM_STRIDE = 16
N_COUNTERS = 16
(CNT_FEATURE_UPDATES, CNT_TRIANGULATIONS, CNT_UPDATES, CNT_FRAMES, CNT_TRI_OBS, CNT_TRI_SOLVES, CNT_JAC_OBS, CNT_K2_FLOPS,
 CNT_K2_BYTES, CNT_QR_FLOPS, CNT_QR_BYTES, CNT_K4_FLOPS, CNT_K4_BYTES, CNT_IMU_STEPS, CNT_GATE_NOT_SPD, CNT_K4_EXEC_FLOPS) = range(16)
FEAT_FEW_OBS, FEAT_TRI_FAILED, FEAT_GATE_PASS, FEAT_GATE_FAIL, FEAT_AFTER_CAP = 1, 2, 3, 4, 5

ERR_ARG, ERR_CUDA, ERR_CAPACITY, ERR_NUMERIC = -1, -2, -3, -4
GATE_NOALLOC_MAX_DOUBLES = 128 * 405 + 128 * 128 + 128     # include/msckf_b200.h MSCKF_GATE_NOALLOC_MAX_DOUBLES


class MsckfCudaError(RuntimeError):
    """Any failure reported by libmsckf_b200 (bad argument, CUDA error, capacity)."""

    def __init__(self, code, text):
        RuntimeError.__init__(self, 'libmsckf_b200 error %d: %s' % (code, text))
        self.code = code


class MsckfNumericError(MsckfCudaError, np.linalg.LinAlgError):
    """Innovation covariance not positive definite (the reference raises LinAlgError here)."""


class CConfig(C.Structure):
    _fields_ = [
        ('n_filters', C.c_int32), ('max_cam_states', C.c_int32), ('feature_capacity', C.c_int32),
        ('max_rows', C.c_int32), ('row_cap', C.c_int32), ('device', C.c_int32),
        ('outer_loop_max_iteration', C.c_int32), ('inner_loop_max_iteration', C.c_int32),
        ('jacobian_pool_bytes', C.c_int64),
        ('gyro_noise', C.c_double), ('gyro_bias_noise', C.c_double), ('acc_noise', C.c_double),
        ('acc_bias_noise', C.c_double), ('observation_noise', C.c_double),
        ('gyro_bias_cov', C.c_double), ('velocity_cov', C.c_double), ('acc_bias_cov', C.c_double),
        ('extrinsic_rotation_cov', C.c_double), ('extrinsic_translation_cov', C.c_double),
        ('position_std_threshold', C.c_double),
        ('translation_threshold', C.c_double), ('huber_epsilon', C.c_double),
        ('estimation_precision', C.c_double), ('initial_damping', C.c_double),
        ('prune_rotation_threshold', C.c_double), ('prune_translation_threshold', C.c_double),
        ('prune_tracking_rate_threshold', C.c_double),
        ('R_cam0_cam1', C.c_double * 9), ('t_cam0_cam1', C.c_double * 3),
        ('chi2_table', C.c_double * 100),
    ]


_P = C.c_void_p
_I = C.c_int32
_L = C.c_int64
# name -> argument types (all return int unless noted); must list every symbol of the header
SIGNATURES = {
    'msckf_create': [C.POINTER(CConfig), C.POINTER(_P)],
    'msckf_destroy': [_P],
    'msckf_last_error': [_P],
    'msckf_sync': [_P],
    'msckf_set_imu_state': [_P, _I, _P],
    'msckf_get_imu_state': [_P, _I, _P],
    'msckf_get_imu_states': [_P, _P],
    'msckf_get_cam_states': [_P, _I, _P, _P, _P],
    'msckf_set_cam_states': [_P, _I, _I, _P, _P],
    'msckf_get_cov': [_P, _I, _P, _I],
    'msckf_set_cov': [_P, _I, _P, _I],
    'msckf_get_features': [_P, _I, _P, _P, _P, _P, _P, _P],
    'msckf_set_features': [_P, _I, _I, _P, _P, _P, _P, _P],
    'msckf_set_feature_position': [_P, _I, _L, _P, _I],
    'msckf_propagate': [_P, _P, _P, _P, _P],
    'msckf_augment': [_P, _P],
    'msckf_add_observations': [_P, _P, _P, _P],
    'msckf_remove_lost_features': [_P],
    'msckf_prune_cam_states': [_P],
    'msckf_online_reset': [_P, _P],
    'msckf_reset': [_P, _I],
    'msckf_feature_callback': [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    'msckf_feature_callback_async': [_P, _P, _P, _P, _P, _P, _P, _P, _P],
    'msckf_collect': [_P, _P, _P],
    'msckf_feature_callback_dev': [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    'msckf_get_stage_log': [_P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P],
    'msckf_get_counters': [_P, _P],
    'msckf_get_update_shapes': [_P, _P],
    'msckf_triangulate': [_P, _I, _P, _P, _P, _P, _P, _P, _P],
    'msckf_initialize_position': [_P, _I, _P, _P, _P, _P, _P, _P, _P],
    'msckf_check_motion': [_P, _I, _P, _P, _P, _P, _P],
    'msckf_initial_guess': [_P, _I, _P, _P, _P, _P, _P],
    'msckf_feature_jacobian': [_P, _I, _L, _I, _P, _P, _P, _P, _I],
    'msckf_measurement_jacobian': [_P, _I, _L, _L, _P, _P, _P],
    'msckf_gating_test': [_P, _I, _P, _P, _I, _I, _I, _P, _P],
    'msckf_gating_test_noalloc': [_P, _I, _P, _P, _I, _I, _I, _P, _P],
    'msckf_measurement_update': [_P, _I, _P, _P, _I, _I],
    'msckf_find_redundant_cam_states': [_P, _I, _P],
    'msckf_last_timing': [_P, _P],
    'msckf_set_timing': [_P, _I],
    'msckf_kernel_launches': [_P, _P],
    'msckf_kernel_count': [],
    'msckf_kernel_name': [_I],
    'msckf_set_kernel_timing': [_P, _I],
    'msckf_kernel_times': [_P, _P, _P, _P],
    'msckf_fp64_peak': [_I, _P, _P],
    'msckf_mark': [_P, _I],
    'msckf_mark_elapsed': [_P, _P],
    'msckf_mark_elapsed_between': [_P, _P, _P],
    'msckf_version': [],
    'msckf_undistort_points': [_I, _I, _P, _P, _I, _P, _P, _P, _P],
    'msckf_undistort_stereo': [_I, _I, _P, _P, _P, _I, _P, _P, _I, _P, _P],
    'msckf_synth_frame': [_I, _I, C.c_uint64, _L, _I, _L, _P, _P, _P, _P, _P, C.c_double, C.c_double, _I, _P, _P, C.c_double,
                          C.c_double, C.c_double, _L, _P, _P, _P, _P, _P, _P, _P, _P],
    'msckf_host_alloc': [C.c_size_t, _P],
    'msckf_host_free': [_P],
}
_RESTYPE = {'msckf_destroy': None, 'msckf_last_error': C.c_char_p, 'msckf_version': C.c_char_p,
            'msckf_kernel_name': C.c_char_p}

_lib = None


def load_library():
    """dlopen the CUDA library (once).  Fails loudly: there is no other implementation."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MsckfCudaError(ERR_CUDA, 'shared library %s not built (run __graft_entry__.build() or make -C %s)'
                             % (LIB_PATH, os.path.join(PKG_DIR, 'csrc')))
    lib = C.CDLL(LIB_PATH)
    for name, args in SIGNATURES.items():
        fn = getattr(lib, name)           # AttributeError if the header and the library diverge
        fn.argtypes = args
        fn.restype = _RESTYPE.get(name, C.c_int)
    _lib = lib
    return lib


def pinned_empty(shape, dtype=np.float64):
    """numpy array in page-locked host memory (msckf_host_alloc): message buffers built here go to the
    device without the extra copy through the library's staging arena.  Freed when the array is collected."""
    import weakref
    lib = load_library()
    dtype = np.dtype(dtype)
    shape = (shape,) if np.isscalar(shape) else tuple(shape)
    nbytes = int(np.prod(shape, dtype=np.int64)) * dtype.itemsize
    p = C.c_void_p(0)
    rc = lib.msckf_host_alloc(max(nbytes, 1), C.addressof(p))
    if rc != 0 or not p.value:
        raise MsckfCudaError(rc, 'msckf_host_alloc(%d bytes) failed' % nbytes)
    buf = (C.c_char * max(nbytes, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape, dtype=np.int64))).reshape(shape)
    weakref.finalize(buf, lib.msckf_host_free, p.value)
    return arr


def pinned_copy(a):
    out = pinned_empty(np.shape(a), np.asarray(a).dtype)
    out[...] = a
    return out


def _ptr(a):
    return None if a is None else a.ctypes.data


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def chi2_table(tail=0.05):
    """msckf.py:121-123 -- computed with scipy on the host so the thresholds are bit-identical."""
    from scipy.stats import chi2
    t = np.zeros(100)
    for i in range(1, 100):
        t[i] = chi2.ppf(tail, i)
    return t


def make_cconfig(config, n_filters=1, feature_capacity=None, max_rows=0, device=0, jacobian_pool_bytes=0,
                 row_cap=None, chi2_tail=None):
    """Fill the C configuration from a reference-style ``ConfigEuRoC`` object (config.py:17-124)."""
    oc = config._vio_optimization_config__
    c = CConfig()
    c.n_filters = int(n_filters)
    c.max_cam_states = int(config._vio_max_cam_state_size__)
    c.feature_capacity = int(feature_capacity if feature_capacity else 2048)
    c.max_rows = int(max_rows)
    rc = getattr(config, 'row_cap', 1500) if row_cap is None else row_cap
    c.row_cap = int(rc)
    c.device = int(device)
    c.outer_loop_max_iteration = int(oc._vio_outer_loop_max_iteration__)
    c.inner_loop_max_iteration = int(oc._vio_inner_loop_max_iteration__)
    c.jacobian_pool_bytes = int(jacobian_pool_bytes)
    c.gyro_noise = config._vio_gyro_noise__
    c.gyro_bias_noise = config._vio_gyro_bias_noise__
    c.acc_noise = config._vio_acc_noise__
    c.acc_bias_noise = config._vio_acc_bias_noise__
    c.observation_noise = config._vio_observation_noise__
    c.gyro_bias_cov = config._vio_gyro_bias_cov__
    c.velocity_cov = config._vio_velocity_cov__
    c.acc_bias_cov = config._vio_acc_bias_cov__
    c.extrinsic_rotation_cov = config._vio_extrinsic_rotation_cov__
    c.extrinsic_translation_cov = config._vio_extrinsic_translation_cov__
    c.position_std_threshold = config._vio_position_std_threshold__
    c.translation_threshold = oc._vio_translation_threshold__
    c.huber_epsilon = oc._vio_huber_epsilon__
    c.estimation_precision = oc._vio_estimation_precision__
    c.initial_damping = oc._vio_initial_damping__
    # msckf.py:717 hard-codes these three (config.py:67-69 holds the same numbers, unused there)
    c.prune_rotation_threshold = 0.2618
    c.prune_translation_threshold = 0.4
    c.prune_tracking_rate_threshold = 0.5
    T = np.asarray(config._vio_T_cn_cnm1__, dtype=np.float64)
    c.R_cam0_cam1[:] = list(T[:3, :3].ravel())
    c.t_cam0_cam1[:] = list(T[:3, 3])
    tail = getattr(config, 'chi2_tail', 0.05) if chi2_tail is None else chi2_tail
    c.chi2_table[:] = list(chi2_table(tail))
    return c


class Context(object):
    """Owner of one ``msckf_ctx`` (a batch of F device-resident filters)."""

    def __init__(self, cconfig):
        self.lib = load_library()
        self.cfg = cconfig
        self.F = cconfig.n_filters
        self.max_cam = cconfig.max_cam_states
        self.feature_capacity = cconfig.feature_capacity
        h = _P()
        rc = self.lib.msckf_create(C.byref(cconfig), C.byref(h))
        if rc != 0:
            raise self._error(rc, None)
        self.h = h

    # ---- plumbing
    def _error(self, rc, h):
        text = self.lib.msckf_last_error(h)
        text = text.decode('utf-8', 'replace') if text else ''
        if rc == ERR_NUMERIC:
            return MsckfNumericError(rc, text)
        return MsckfCudaError(rc, text)

    def _ck(self, rc):
        if rc != 0:
            raise self._error(rc, self.h)

    def close(self):
        if getattr(self, 'h', None):
            self.lib.msckf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        self._ck(self.lib.msckf_sync(self.h))

    # ---- state I/O
    def set_imu_state(self, f, imu):
        imu = _f64(imu)
        assert imu.size == IMU_STRIDE
        self._ck(self.lib.msckf_set_imu_state(self.h, f, _ptr(imu)))

    def get_imu_state(self, f=0):
        out = np.zeros(IMU_STRIDE)
        self._ck(self.lib.msckf_get_imu_state(self.h, f, _ptr(out)))
        return out

    def get_imu_states(self):
        out = np.zeros((self.F, IMU_STRIDE))
        self._ck(self.lib.msckf_get_imu_states(self.h, _ptr(out)))
        return out

    def get_cam_states(self, f=0):
        n = C.c_int32(0)
        ids = np.zeros(self.max_cam, dtype=np.int64)
        cams = np.zeros((self.max_cam, CAM_STRIDE))
        self._ck(self.lib.msckf_get_cam_states(self.h, f, C.addressof(n), _ptr(ids), _ptr(cams)))
        return ids[:n.value].copy(), cams[:n.value].copy()

    def set_cam_states(self, f, ids, cams):
        ids = _i64(ids)
        cams = _f64(cams).reshape(len(ids), CAM_STRIDE) if len(ids) else np.zeros((0, CAM_STRIDE))
        self._ck(self.lib.msckf_set_cam_states(self.h, f, len(ids), _ptr(ids), _ptr(cams)))

    def n_state(self, f=0):
        n = C.c_int32(0)
        self._ck(self.lib.msckf_get_cam_states(self.h, f, C.addressof(n), None, None))
        return 21 + 6 * n.value

    def get_cov(self, f=0, n=None):
        if n is None:
            n = self.n_state(f)
        P = np.zeros((n, n))
        self._ck(self.lib.msckf_get_cov(self.h, f, _ptr(P), n))
        return P

    def set_cov(self, f, P):
        P = _f64(P)
        self._ck(self.lib.msckf_set_cov(self.h, f, _ptr(P), P.shape[0]))

    def get_features(self, f=0, with_obs=True):
        cap = self.feature_capacity
        n = C.c_int32(0)
        ids = np.zeros(cap, dtype=np.int64)
        mask = np.zeros(cap, dtype=np.uint64)
        init = np.zeros(cap, dtype=np.uint8)
        pos = np.zeros((cap, 3))
        obs = np.zeros((cap, self.max_cam, 4)) if with_obs else None
        self._ck(self.lib.msckf_get_features(self.h, f, C.addressof(n), _ptr(ids), _ptr(mask), _ptr(init), _ptr(pos),
                                             _ptr(obs)))
        k = n.value
        return ids[:k].copy(), mask[:k].copy(), init[:k].astype(bool), pos[:k].copy(), (obs[:k].copy() if with_obs else None)

    def set_features(self, f, ids, obs_mask, is_initialized, position, obs):
        ids = _i64(ids)
        n = len(ids)
        mask = np.ascontiguousarray(obs_mask, dtype=np.uint64)
        init = None if is_initialized is None else np.ascontiguousarray(is_initialized, dtype=np.uint8)
        pos = None if position is None else _f64(position)
        obs = _f64(obs)
        assert obs.size == n * self.max_cam * 4
        self._ck(self.lib.msckf_set_features(self.h, f, n, _ptr(ids), _ptr(mask), _ptr(init), _ptr(pos), _ptr(obs)))

    def set_feature_position(self, f, fid, position, is_initialized):
        p = _f64(position)
        self._ck(self.lib.msckf_set_feature_position(self.h, f, int(fid), _ptr(p), int(bool(is_initialized))))

    # ---- stages
    def propagate(self, imu_off, ts, gyro, acc):
        imu_off, ts, gyro, acc = _i32(imu_off), _f64(ts), _f64(gyro), _f64(acc)
        self._ck(self.lib.msckf_propagate(self.h, _ptr(imu_off), _ptr(ts), _ptr(gyro), _ptr(acc)))

    def augment(self, cam_ids):
        cam_ids = _i64(cam_ids)
        self._ck(self.lib.msckf_augment(self.h, _ptr(cam_ids)))

    def add_observations(self, feat_off, ids, uv):
        feat_off, ids, uv = _i32(feat_off), _i64(ids), _f64(uv)
        self._ck(self.lib.msckf_add_observations(self.h, _ptr(feat_off), _ptr(ids), _ptr(uv)))

    def remove_lost_features(self):
        self._ck(self.lib.msckf_remove_lost_features(self.h))

    def prune_cam_states(self):
        self._ck(self.lib.msckf_prune_cam_states(self.h))

    def online_reset(self):
        out = np.zeros(self.F, dtype=np.uint8)
        self._ck(self.lib.msckf_online_reset(self.h, _ptr(out)))
        return out.astype(bool)

    def reset(self, f=0):
        self._ck(self.lib.msckf_reset(self.h, f))

    def feature_callback(self, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv):
        self.feature_callback_async(imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv)
        return self.collect()

    def feature_callback_async(self, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv):
        imu_off, ts, gyro, acc = _i32(imu_off), _f64(ts), _f64(gyro), _f64(acc)
        cam_ids, feat_off, ids, uv = _i64(cam_ids), _i32(feat_off), _i64(ids), _f64(uv)
        self._ck(self.lib.msckf_feature_callback_async(self.h, _ptr(imu_off), _ptr(ts), _ptr(gyro), _ptr(acc), _ptr(cam_ids),
                                                       _ptr(feat_off), _ptr(ids), _ptr(uv)))

    def collect(self):
        imu = np.zeros((self.F, IMU_STRIDE))
        rst = np.zeros(self.F, dtype=np.uint8)
        self._ck(self.lib.msckf_collect(self.h, _ptr(imu), _ptr(rst)))
        return imu, rst.astype(bool)

    # ---- logs
    def stage_log(self, f, stage):
        cap = self.feature_capacity
        n = C.c_int32(0)
        ids = np.zeros(cap, dtype=np.int64)
        status = np.zeros(cap, dtype=np.int32)
        gamma = np.zeros(cap)
        dof = np.zeros(cap, dtype=np.int32)
        ur, uc = C.c_int32(0), C.c_int32(0)
        rm = np.full(2, -1, dtype=np.int64)
        self._ck(self.lib.msckf_get_stage_log(self.h, f, stage, C.addressof(n), _ptr(ids), _ptr(status), _ptr(gamma), _ptr(dof),
                                              C.addressof(ur), C.addressof(uc), _ptr(rm)))
        k = n.value
        return {'ids': ids[:k].copy(), 'status': status[:k].copy(), 'gamma': gamma[:k].copy(), 'dof': dof[:k].copy(),
                'update_rows': ur.value, 'update_cols': uc.value, 'removed_cam_ids': rm}

    def counters(self):
        out = np.zeros((self.F, N_COUNTERS), dtype=np.int64)
        self._ck(self.lib.msckf_get_counters(self.h, _ptr(out)))
        return out

    def update_shapes(self):
        out = np.zeros((self.F, 4), dtype=np.int32)
        self._ck(self.lib.msckf_get_update_shapes(self.h, _ptr(out)))
        return out

    def feature_callback_dev(self, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv, imu_out=0):
        """All arguments are raw DEVICE addresses (ints), e.g. torch ``tensor.data_ptr()``."""
        self._ck(self.lib.msckf_feature_callback_dev(self.h, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv, imu_out or None))

    def mark(self, which):
        self._ck(self.lib.msckf_mark(self.h, int(which)))

    def mark_elapsed_ms(self):
        ms = C.c_float(0)
        self._ck(self.lib.msckf_mark_elapsed(self.h, C.addressof(ms)))
        return ms.value

    def set_kernel_timing(self, on):
        self._ck(self.lib.msckf_set_kernel_timing(self.h, int(bool(on))))

    def kernel_times(self):
        """{kernel name: (launches, timed launches, total ms of the timed launches)}"""
        k = self.lib.msckf_kernel_count()
        la, tl, ms = np.zeros(k, dtype=np.int64), np.zeros(k, dtype=np.int64), np.zeros(k)
        self._ck(self.lib.msckf_kernel_times(self.h, _ptr(la), _ptr(tl), _ptr(ms)))
        return dict((self.lib.msckf_kernel_name(i).decode(), (int(la[i]), int(tl[i]), float(ms[i]))) for i in range(k))

    def kernel_launches(self):
        n = C.c_int64(0)
        self._ck(self.lib.msckf_kernel_launches(self.h, C.addressof(n)))
        return n.value

    # ---- direct pieces
    def triangulate(self, obs_off, uv, cam_q, cam_p, check_motion=True):
        obs_off, uv, cam_q, cam_p = _i32(obs_off), _f64(uv), _f64(cam_q), _f64(cam_p)
        n = len(obs_off) - 1
        pos = np.zeros((n, 3))
        valid = np.zeros(n, dtype=np.uint8)
        solves = np.zeros(n, dtype=np.int32)
        fn = self.lib.msckf_triangulate if check_motion else self.lib.msckf_initialize_position
        self._ck(fn(self.h, n, _ptr(obs_off), _ptr(uv), _ptr(cam_q), _ptr(cam_p), _ptr(pos), _ptr(valid), _ptr(solves)))
        return pos, valid.astype(bool), solves

    def check_motion(self, obs_off, uv, cam_q, cam_p):
        obs_off, uv, cam_q, cam_p = _i32(obs_off), _f64(uv), _f64(cam_q), _f64(cam_p)
        n = len(obs_off) - 1
        out = np.zeros(n, dtype=np.uint8)
        self._ck(self.lib.msckf_check_motion(self.h, n, _ptr(obs_off), _ptr(uv), _ptr(cam_q), _ptr(cam_p), _ptr(out)))
        return out.astype(bool)

    def initial_guess(self, obs_off, uv, cam_q, cam_p):
        obs_off, uv, cam_q, cam_p = _i32(obs_off), _f64(uv), _f64(cam_q), _f64(cam_p)
        n = len(obs_off) - 1
        out = np.zeros((n, 3))
        self._ck(self.lib.msckf_initial_guess(self.h, n, _ptr(obs_off), _ptr(uv), _ptr(cam_q), _ptr(cam_p), _ptr(out)))
        return out

    def feature_jacobian(self, f, fid, cam_ids):
        cam_ids = _i64(cam_ids)
        n = self.n_state(f)
        cap_rows = 4 * len(cam_ids)
        H = np.zeros((cap_rows, n))
        r = np.zeros(cap_rows)
        rows = C.c_int32(0)
        self._ck(self.lib.msckf_feature_jacobian(self.h, f, int(fid), len(cam_ids), _ptr(cam_ids), C.addressof(rows), _ptr(H),
                                                 _ptr(r), n))
        k = rows.value
        return H.ravel()[:k * n].reshape(k, n).copy(), r[:k].copy()

    def measurement_jacobian(self, f, cam_id, fid):
        Hx, Hf, r = np.zeros((4, 6)), np.zeros((4, 3)), np.zeros(4)
        self._ck(self.lib.msckf_measurement_jacobian(self.h, f, int(cam_id), int(fid), _ptr(Hx), _ptr(Hf), _ptr(r)))
        return Hx, Hf, r

    def gating_test(self, f, H, r, dof):
        H, r = _f64(H), _f64(r)
        g = C.c_double(0)
        p = C.c_int32(0)
        self._ck(self.lib.msckf_gating_test(self.h, f, _ptr(H), _ptr(r), H.shape[0], H.shape[1], int(dof), C.addressof(g),
                                            C.addressof(p)))
        return g.value, bool(p.value)

    def gating_test_noalloc(self, f, H_dev, r_dev, m, n_state, dof, gamma_dev, pass_dev):
        """msckf_gating_test_noalloc: all pointers are raw DEVICE addresses (ints); enqueued on the context's stream, no
        allocation.  Results are valid after sync()."""
        self._ck(self.lib.msckf_gating_test_noalloc(self.h, int(f), H_dev, r_dev, int(m), int(n_state), int(dof), gamma_dev, pass_dev))

    def measurement_update(self, f, H, r):
        H, r = _f64(H), _f64(r)
        if H.shape[0] == 0:
            return
        self._ck(self.lib.msckf_measurement_update(self.h, f, _ptr(H), _ptr(r), H.shape[0], H.shape[1]))

    def find_redundant_cam_states(self, f=0):
        out = np.zeros(2, dtype=np.int64)
        self._ck(self.lib.msckf_find_redundant_cam_states(self.h, f, _ptr(out)))
        return [int(out[0]), int(out[1])]

    def set_timing(self, on):
        self._ck(self.lib.msckf_set_timing(self.h, int(bool(on))))

    def last_timing(self):
        ms = np.zeros(5, dtype=np.float32)
        self._ck(self.lib.msckf_last_timing(self.h, _ptr(ms)))
        return ms


class ContextGroup(object):
    """F filters spread over G contexts (each with its own CUDA stream, work lists and frame graph) behind the interface of
    one ``Context``: per-filter calls go to the context that owns the filter, batch calls are split at the group
    boundaries (ragged offsets rebased) and enqueued on every group before anything is waited for, so the kernels of the
    groups interleave on the GPU: the tail of one group's launch (a launch lasts as long as its longest filter) is filled
    by the other group's CTAs.  Filters are independent, so the results do not depend on the grouping."""

    PER_FILTER = ('set_imu_state', 'get_imu_state', 'get_cam_states', 'set_cam_states', 'n_state', 'get_cov', 'set_cov',
                  'get_features', 'set_features', 'set_feature_position', 'reset', 'stage_log', 'feature_jacobian',
                  'measurement_jacobian', 'gating_test', 'measurement_update', 'find_redundant_cam_states')

    def __init__(self, cconfigs):
        self.parts = []
        try:
            for c in cconfigs:
                self.parts.append(Context(c))
        except Exception:
            self.close()
            raise
        self.lib = self.parts[0].lib
        self.bounds = np.concatenate([[0], np.cumsum([p.F for p in self.parts])]).astype(np.int64)
        self.F = int(self.bounds[-1])
        self.max_cam = self.parts[0].max_cam
        self.feature_capacity = self.parts[0].feature_capacity
        for name in self.PER_FILTER:
            setattr(self, name, self._route(name))

    def _route(self, name):
        def call(f=0, *a, **kw):
            g = int(np.searchsorted(self.bounds, f, side='right')) - 1
            if not (0 <= f < self.F):
                raise MsckfCudaError(ERR_ARG, 'filter index %d outside 0..%d' % (f, self.F - 1))
            return getattr(self.parts[g], name)(int(f - self.bounds[g]), *a, **kw)
        call.__name__ = name
        return call

    def ranges(self):
        return [(int(self.bounds[g]), int(self.bounds[g + 1])) for g in range(len(self.parts))]

    def close(self):
        for p in self.parts:
            p.close()

    def sync(self):
        for p in self.parts:
            p.sync()

    # ---- splitting of ragged batch arguments
    def _split_imu(self, imu_off, ts, gyro, acc):
        ts = np.asarray(ts, dtype=np.float64).reshape(-1)
        gyro = np.asarray(gyro, dtype=np.float64).reshape(-1, 3)
        acc = np.asarray(acc, dtype=np.float64).reshape(-1, 3)
        o = np.asarray(imu_off)
        out = []
        for a, b in self.ranges():
            lo, hi = int(o[a]), int(o[b])
            out.append(((o[a:b + 1] - lo).astype(np.int32), ts[lo:hi], gyro[lo:hi], acc[lo:hi]))
        return out

    def _split_feat(self, feat_off, ids, uv):
        ids = np.asarray(ids, dtype=np.int64).reshape(-1)
        uv = np.asarray(uv, dtype=np.float64).reshape(-1, 4)
        o = np.asarray(feat_off)
        out = []
        for a, b in self.ranges():
            lo, hi = int(o[a]), int(o[b])
            out.append(((o[a:b + 1] - lo).astype(np.int32), ids[lo:hi], uv[lo:hi]))
        return out

    # ---- stages
    def propagate(self, imu_off, ts, gyro, acc):
        for p, a in zip(self.parts, self._split_imu(imu_off, ts, gyro, acc)):
            p.propagate(*a)

    def augment(self, cam_ids):
        cam_ids = np.asarray(cam_ids, dtype=np.int64)
        for p, (a, b) in zip(self.parts, self.ranges()):
            p.augment(cam_ids[a:b])

    def add_observations(self, feat_off, ids, uv):
        for p, a in zip(self.parts, self._split_feat(feat_off, ids, uv)):
            p.add_observations(*a)

    def remove_lost_features(self):
        for p in self.parts:
            p.remove_lost_features()

    def prune_cam_states(self):
        for p in self.parts:
            p.prune_cam_states()

    def online_reset(self):
        return np.concatenate([p.online_reset() for p in self.parts])

    def feature_callback(self, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv):
        self.feature_callback_async(imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv)
        return self.collect()

    def feature_callback_async(self, imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv):
        cam_ids = np.asarray(cam_ids, dtype=np.int64)
        imu = self._split_imu(imu_off, ts, gyro, acc)
        feat = self._split_feat(feat_off, ids, uv)
        for g, (p, (a, b)) in enumerate(zip(self.parts, self.ranges())):
            p.feature_callback_async(imu[g][0], imu[g][1], imu[g][2], imu[g][3], cam_ids[a:b], feat[g][0], feat[g][1], feat[g][2])

    def collect(self):
        r = [p.collect() for p in self.parts]
        return np.concatenate([x[0] for x in r]), np.concatenate([x[1] for x in r])

    def feature_callback_dev(self, *a, **kw):
        raise MsckfCudaError(ERR_ARG, 'a context group takes one set of device addresses per group: feature_callback_dev_parts')

    def feature_callback_dev_parts(self, args_per_group, imu_out_per_group=None):
        """``args_per_group[g]`` = the eight device addresses of group g's messages (offsets rebased to the group),
        ``imu_out_per_group[g]`` = where its IMU blocks go (0: nowhere)."""
        for g, p in enumerate(self.parts):
            p.feature_callback_dev(*args_per_group[g], imu_out=(imu_out_per_group[g] if imu_out_per_group else 0))

    # ---- read-backs over all filters
    def get_imu_states(self):
        return np.concatenate([p.get_imu_states() for p in self.parts])

    def counters(self):
        return np.concatenate([p.counters() for p in self.parts])

    def update_shapes(self):
        return np.concatenate([p.update_shapes() for p in self.parts])

    def kernel_launches(self):
        return sum(p.kernel_launches() for p in self.parts)

    def set_kernel_timing(self, on):
        for p in self.parts:
            p.set_kernel_timing(on)

    def kernel_times(self):
        out = {}
        for p in self.parts:
            for k, v in p.kernel_times().items():
                o = out.get(k, (0, 0, 0.0))
                out[k] = (o[0] + v[0], o[1] + v[1], o[2] + v[2])
        return out

    # ---- timing marks: the groups are idle when mark 0 is taken (host sync), group 0's mark comes first; the elapsed
    # time runs from it to the LAST group's mark 1
    def mark(self, which):
        if which == 0:
            self.sync()
        for p in self.parts:
            p.mark(which)

    def mark_elapsed_ms(self):
        worst = 0.0
        for p in self.parts:
            ms = C.c_float(0)
            p._ck(self.lib.msckf_mark_elapsed_between(self.parts[0].h, p.h, C.addressof(ms)))
            worst = max(worst, ms.value)
        return worst


def fp64_peak(device=0):
    """(DFMA TFLOP/s, DMMA TFLOP/s) measured on the device."""
    a, b = C.c_double(0), C.c_double(0)
    rc = load_library().msckf_fp64_peak(int(device), C.addressof(a), C.addressof(b))
    if rc != 0:
        raise MsckfCudaError(rc, 'msckf_fp64_peak failed')
    return a.value, b.value


def synth_frame(device, n_filters, seed0, frame, sample0, ts, gyro_clean, acc_clean, gyro_bias, acc_bias, sigma_gyro, sigma_acc,
                ids, uv_clean, sigma_z, outlier_frac, outlier_sigma, cam_id, out_ptrs):
    """msckf_synth_frame: ``out_ptrs`` = device addresses (ints) of imu_off, ts, gyro, acc, cam_ids, feat_off, ids, uv."""
    ts, gyro_clean, acc_clean = _f64(ts), _f64(gyro_clean), _f64(acc_clean)
    gyro_bias, acc_bias, ids, uv_clean = _f64(gyro_bias), _f64(acc_bias), _i64(ids), _f64(uv_clean)
    rc = load_library().msckf_synth_frame(int(device), int(n_filters), int(seed0), int(frame), len(ts), int(sample0), _ptr(ts),
                                          _ptr(gyro_clean), _ptr(acc_clean), _ptr(gyro_bias), _ptr(acc_bias), float(sigma_gyro),
                                          float(sigma_acc), len(ids), _ptr(ids), _ptr(uv_clean), float(sigma_z), float(outlier_frac),
                                          float(outlier_sigma), int(cam_id), *[int(p) for p in out_ptrs])
    if rc != 0:
        raise MsckfCudaError(rc, 'msckf_synth_frame failed')


def version():
    return load_library().msckf_version().decode()
