"""Drop-in ``MSCKF`` back end running on the B200 through ``libmsckf_b200.so``.

Mirrors the public surface of the reference's ``MSCKF`` class (msckf.py:104-908): same method
names, argument meaning and attribute names (including the renamed ``_vio_*__`` ones), so a
harness written against the reference reads this object unchanged.  What differs is where the
state lives: IMU state, camera-state window, covariance and the landmark map (``map_server``)
are device resident; ``state_server`` and ``map_server`` are *snapshots* downloaded on access.

The host keeps only what the reference keeps outside its numeric path: the IMU message buffer
(msckf.py:166-178, 262-287 -- message selection by timestamp), the running IMU-state id and the
first-image / gravity flags.  Every numeric stage is a CUDA launch; there is no CPU fallback.
"""
from collections import namedtuple

import numpy as np

from . import capi
from .capi import (IMU_Q, IMU_BG, IMU_V, IMU_BA, IMU_P, IMU_Q_NULL, IMU_V_NULL, IMU_P_NULL, IMU_R_IC, IMU_T_CI,
                   IMU_GRAVITY, IMU_TIME, IMU_NULL_ALIASED, IMU_TRACKING_RATE, IMU_STRIDE, CAM_STRIDE)
from .utils import Isometry3d, to_rotation, from_two_vectors

VioResult = namedtuple('vio_result', ['timestamp', 'pose', 'velocity', 'cam0_pose'])


class IMUState(object):
    """msckf.py:16-59 (snapshot; ``gravity`` is per filter on the device, a class global in the reference)."""
    _vio_next_id__ = 0
    _vio_gravity__ = np.array([0., 0., -9.81])
    _vio_T_imu_body__ = Isometry3d(np.identity(3), np.zeros(3))

    def __init__(self, new_id=None):
        self.id = new_id
        self.timestamp = None
        self.orientation = np.array([0., 0., 0., 1.])
        self._vio_position__ = np.zeros(3)
        self.velocity = np.zeros(3)
        self._vio_gyro_bias__ = np.zeros(3)
        self.acc_bias = np.zeros(3)
        self.orientation_null = np.array([0., 0., 0., 1.])
        self.position_null = np.zeros(3)
        self.velocity_null = np.zeros(3)
        self.R_imu_cam0 = np.identity(3)
        self.t_cam0_imu = np.zeros(3)

    @staticmethod
    def from_block(b, new_id=None):
        s = IMUState(new_id)
        s.timestamp = float(b[IMU_TIME])
        s.orientation = b[IMU_Q:IMU_Q + 4].copy()
        s._vio_gyro_bias__ = b[IMU_BG:IMU_BG + 3].copy()
        s.velocity = b[IMU_V:IMU_V + 3].copy()
        s.acc_bias = b[IMU_BA:IMU_BA + 3].copy()
        s._vio_position__ = b[IMU_P:IMU_P + 3].copy()
        s.orientation_null = b[IMU_Q_NULL:IMU_Q_NULL + 4].copy()
        s.velocity_null = b[IMU_V_NULL:IMU_V_NULL + 3].copy()
        s.position_null = b[IMU_P_NULL:IMU_P_NULL + 3].copy()
        s.R_imu_cam0 = b[IMU_R_IC:IMU_R_IC + 9].reshape(3, 3).copy()
        s.t_cam0_imu = b[IMU_T_CI:IMU_T_CI + 3].copy()
        s.gravity = b[IMU_GRAVITY:IMU_GRAVITY + 3].copy()
        return s


class CAMState(object):
    """msckf.py:62-86 (snapshot)"""
    _vio_R_cam0_cam1__ = None
    _vio_t_cam0_cam1__ = None

    def __init__(self, new_id=None):
        self.id = new_id
        self.timestamp = None
        self.orientation = np.array([0., 0., 0., 1.])
        self._vio_position__ = np.zeros(3)
        self.orientation_null = np.array([0., 0., 0., 1.])
        self.position_null = np.zeros(3)

    @staticmethod
    def from_block(cid, b):
        c = CAMState(int(cid))
        c.orientation = b[0:4].copy()
        c._vio_position__ = b[4:7].copy()
        c.orientation_null = b[7:11].copy()
        c.position_null = b[11:14].copy()
        return c

    def to_block(self):
        b = np.zeros(CAM_STRIDE)
        b[0:4], b[4:7], b[7:11], b[11:14] = self.orientation, self._vio_position__, self.orientation_null, self.position_null
        return b


class StateServer(object):
    """msckf.py:89-100 as a view of the device state.  ``_vio_state_cov__`` is downloaded on first
    access and uploaded when assigned."""

    def __init__(self, owner):
        self._owner = owner
        ctx, f = owner.ctx, owner.filter_index
        self._vio_imu_state__ = IMUState.from_block(ctx.get_imu_state(f), owner._imu_id)
        ids, cams = ctx.get_cam_states(f)
        self._vio_cam_states__ = dict((int(i), CAMState.from_block(i, c)) for i, c in zip(ids, cams))
        self._vio_continuous_noise_cov__ = owner._continuous_noise_cov
        self._P = None

    @property
    def _vio_state_cov__(self):
        if self._P is None:
            self._P = self._owner.ctx.get_cov(self._owner.filter_index, 21 + 6 * len(self._vio_cam_states__))
        return self._P

    @_vio_state_cov__.setter
    def _vio_state_cov__(self, P):
        P = np.ascontiguousarray(P, dtype=np.float64)
        self._owner.ctx.set_cov(self._owner.filter_index, P)
        self._P = P


class FeatureView(object):
    """One landmark of the device map (the reference's Feature fields, feature.py:15-31)."""
    __slots__ = ('id', 'observations', '_vio_position__', 'is_initialized', 'optimization_config')

    def __init__(self, fid, observations, position, is_initialized, optimization_config):
        self.id = fid
        self.observations = observations
        self._vio_position__ = position
        self.is_initialized = is_initialized
        self.optimization_config = optimization_config


def unpack_features(feature_msg):
    """``vio_feature_msg__`` (image.py:425-438) -> (ids int64[n], uv float64[n][4]).  Messages that
    already carry ``ids`` / ``uv`` arrays (synth.PackedFeatureMsg) skip the per-object loop."""
    ids = getattr(feature_msg, 'ids', None)
    if ids is not None:
        return np.ascontiguousarray(ids, dtype=np.int64), np.ascontiguousarray(feature_msg.uv, dtype=np.float64)
    feats = feature_msg.vio_features
    ids = np.fromiter((f.id for f in feats), dtype=np.int64, count=len(feats))
    uv = np.array([(f.u0, f.v0, f.u1, f.v1) for f in feats], dtype=np.float64).reshape(len(feats), 4)
    return ids, uv


class MSCKF(object):
    """Same constructor and methods as the reference's ``MSCKF(config)`` (msckf.py:104-164).

    Extra keyword arguments size the device buffers: ``feature_capacity`` (live landmarks),
    ``max_rows`` (stacked-Jacobian rows, default row cap + slack), ``device`` (CUDA ordinal).
    """

    def __init__(self, config, device=0, feature_capacity=2048, max_rows=0, jacobian_pool_bytes=0):
        self.config = config
        self.optimization_config = config._vio_optimization_config__
        self.imu_msg_buffer = []
        self.filter_index = 0
        self.cconfig = capi.make_cconfig(config, n_filters=1, feature_capacity=feature_capacity, max_rows=max_rows,
                                         device=device, jacobian_pool_bytes=jacobian_pool_bytes)
        self.ctx = capi.Context(self.cconfig)
        self.chi_squared_test_table = dict((i, self.cconfig.chi2_table[i]) for i in range(1, 100))
        Qc = np.identity(12)
        Qc[:3, :3] *= config._vio_gyro_noise__
        Qc[3:6, 3:6] *= config._vio_gyro_bias_noise__
        Qc[6:9, 6:9] *= config._vio_acc_noise__
        Qc[9:, 9:] *= config._vio_acc_bias_noise__
        self._continuous_noise_cov = Qc
        self.T_imu_body = Isometry3d(np.asarray(config._vio_T_imu_body__)[:3, :3], np.asarray(config._vio_T_imu_body__)[:3, 3])
        # class-level extrinsics, as the reference sets them (msckf.py:148-155)
        from .feature import Feature
        T01 = np.asarray(config._vio_T_cn_cnm1__, dtype=np.float64)
        CAMState._vio_R_cam0_cam1__ = Feature._vio_R_cam0_cam1__ = T01[:3, :3]
        CAMState._vio_t_cam0_cam1__ = Feature._vio_t_cam0_cam1__ = T01[:3, 3]
        IMUState._vio_gravity__ = config._vio_gravity__
        IMUState._vio_T_imu_body__ = self.T_imu_body
        self._imu_id = None          # IMUState.id (msckf.py:283)
        self._next_imu_id = 0        # IMUState._vio_next_id__ (a class global in the reference)
        self._imu_time = None        # IMUState.timestamp, mirrored on the host for message selection
        self.tracking_rate = None
        self.is_gravity_set = False
        self.is_first_img = True
        self.did_reset = False
        self.ctx.set_imu_state(0, initial_imu_block(config))

    # ------------------------------------------------------------------ snapshots
    @property
    def state_server(self):
        return StateServer(self)

    @property
    def map_server(self):
        ids, mask, init, pos, obs = self.ctx.get_features(self.filter_index)
        cam_ids, _ = self.ctx.get_cam_states(self.filter_index)
        out = dict()
        for k in range(len(ids)):
            o = dict()
            m = int(mask[k])
            for c, cid in enumerate(cam_ids):
                if (m >> c) & 1:
                    o[int(cid)] = obs[k, c].copy()
            out[int(ids[k])] = FeatureView(int(ids[k]), o, pos[k].copy(), bool(init[k]), self.optimization_config)
        return out

    # ------------------------------------------------------------------ msckf.py:166-178
    def imu_callback(self, imu_msg):
        self.imu_msg_buffer.append(imu_msg)
        if not self.is_gravity_set:
            if len(self.imu_msg_buffer) >= 200:
                self.initialize_gravity_and_bias()
                self.is_gravity_set = True

    # ------------------------------------------------------------------ msckf.py:235-258 (one-off, host)
    def initialize_gravity_and_bias(self):
        sum_w = np.zeros(3)
        sum_a = np.zeros(3)
        for m in self.imu_msg_buffer:
            sum_w += m.angular_velocity
            sum_a += m.linear_acceleration
        b = self.ctx.get_imu_state(self.filter_index)
        b[IMU_BG:IMU_BG + 3] = sum_w / len(self.imu_msg_buffer)
        g_imu = sum_a / len(self.imu_msg_buffer)
        gravity = np.array([0., 0., -np.linalg.norm(g_imu)])
        b[IMU_GRAVITY:IMU_GRAVITY + 3] = gravity
        b[IMU_Q:IMU_Q + 4] = from_two_vectors(-gravity, g_imu)
        self.ctx.set_imu_state(self.filter_index, b)

    # ------------------------------------------------------------------ message selection (msckf.py:262-287)
    def _select_imu(self, time_bound):
        used = 0
        ts, gy, ac = [], [], []
        t_state = self._imu_time
        for m in self.imu_msg_buffer:
            t = m.vio_timestamp__
            if t < t_state:
                used += 1
                continue
            if t > time_bound:
                break
            ts.append(t)
            gy.append(m.angular_velocity)
            ac.append(m.linear_acceleration)
            used += 1
            t_state = t
        self.imu_msg_buffer = self.imu_msg_buffer[used:]
        n = len(ts)
        return (np.array(ts, dtype=np.float64), np.array(gy, dtype=np.float64).reshape(n, 3),
                np.array(ac, dtype=np.float64).reshape(n, 3), t_state)

    def _begin_frame(self, timestamp):
        if self.is_first_img:
            self.is_first_img = False
            self._imu_time = timestamp
            b = self.ctx.get_imu_state(self.filter_index)
            b[IMU_TIME] = timestamp
            self.ctx.set_imu_state(self.filter_index, b)

    # ------------------------------------------------------------------ msckf.py:180-233
    def feature_callback(self, feature_msg):
        if not self.is_gravity_set:
            return None
        self._begin_frame(feature_msg.timestamp)
        ts, gy, ac, t_state = self._select_imu(feature_msg.timestamp)
        self._imu_time = t_state
        self._imu_id = self._next_imu_id
        self._next_imu_id += 1
        ids, uv = unpack_features(feature_msg)
        imu, did_reset = self.ctx.feature_callback([0, len(ts)], ts, gy, ac, [self._imu_id], [0, len(ids)], ids, uv)
        self.tracking_rate = float(imu[0, IMU_TRACKING_RATE])
        self.did_reset = bool(did_reset[0])          # online_reset fired at the end of this frame (msckf.py:859-886)
        return self._publish_from_block(feature_msg.timestamp, imu[0])

    # ------------------------------------------------------------------ stage-by-stage API (same names as the reference)
    def batch_imu_processing(self, time_bound):
        ts, gy, ac, t_state = self._select_imu(time_bound)
        if len(ts):
            self.ctx.propagate([0, len(ts)], ts, gy, ac)
        self._imu_time = t_state
        self._imu_id = self._next_imu_id
        self._next_imu_id += 1

    def process_model(self, time, m_gyro, m_acc):
        """msckf.py:291-368 for one IMU sample (the state time advances to ``time``)."""
        self.ctx.propagate([0, 1], [time], np.asarray(m_gyro, dtype=np.float64).reshape(1, 3),
                           np.asarray(m_acc, dtype=np.float64).reshape(1, 3))
        self._imu_time = time

    def state_augmentation(self, time):
        self.ctx.augment([self._imu_id])

    def add_feature_observations(self, feature_msg):
        ids, uv = unpack_features(feature_msg)
        self.ctx.add_observations([0, len(ids)], ids, uv)
        self.tracking_rate = float(self.ctx.get_imu_state(self.filter_index)[IMU_TRACKING_RATE])

    def measurement_jacobian(self, cam_state_id, feature_id):
        return self.ctx.measurement_jacobian(self.filter_index, cam_state_id, feature_id)

    def feature_jacobian(self, feature_id, cam_state_ids):
        return self.ctx.feature_jacobian(self.filter_index, feature_id, list(cam_state_ids))

    def measurement_update(self, H, r):
        if len(H) == 0 or len(r) == 0:
            return
        self.ctx.measurement_update(self.filter_index, H, r)

    def gating_test(self, H, r, dof):
        return self.ctx.gating_test(self.filter_index, H, r, dof)[1]

    def remove_lost_features(self):
        self.ctx.remove_lost_features()
        self.ctx.sync()

    def find_redundant_cam_states(self):
        return self.ctx.find_redundant_cam_states(self.filter_index)

    def prune_cam_state_buffer(self):
        self.ctx.prune_cam_states()
        self.ctx.sync()

    def reset(self):
        self.ctx.reset(self.filter_index)
        self.ctx.sync()
        self.imu_msg_buffer.clear()
        self.is_gravity_set = False
        self.is_first_img = True

    def online_reset(self):
        return bool(self.ctx.online_reset()[self.filter_index])

    def publish(self, time):
        return self._publish_from_block(time, self.ctx.get_imu_state(self.filter_index))

    # ------------------------------------------------------------------ msckf.py:888-908
    def _publish_from_block(self, time, b):
        R_i_w = to_rotation(b[IMU_Q:IMU_Q + 4]).T
        p = b[IMU_P:IMU_P + 3].copy()
        T_i_w = Isometry3d(R_i_w, p)
        T_b_w = self.T_imu_body * T_i_w * self.T_imu_body.inverse()
        body_velocity = self.T_imu_body._vio_R__ @ b[IMU_V:IMU_V + 3]
        R_w_c = b[IMU_R_IC:IMU_R_IC + 9].reshape(3, 3) @ R_i_w.T
        t_c_w = p + R_i_w @ b[IMU_T_CI:IMU_T_CI + 3]
        return VioResult(time, T_b_w, body_velocity, Isometry3d(R_w_c.T, t_c_w))

    # ------------------------------------------------------------------ decision logs (tests, diagnostics)
    def stage_log(self, stage):
        """stage 0 = last remove_lost_features, 1 = last prune_cam_state_buffer."""
        return self.ctx.stage_log(self.filter_index, stage)

    def close(self):
        self.ctx.close()


def initial_imu_block(config):
    """IMU block at construction (msckf.py:127-145): identity orientation, configured velocity,
    gravity and the IMU-cam0 extrinsics exactly as the reference derives them."""
    b = np.zeros(IMU_STRIDE)
    b[IMU_Q + 3] = 1.0
    b[IMU_Q_NULL + 3] = 1.0
    b[IMU_V:IMU_V + 3] = config._vio_velocity__
    T_cam0_imu = np.linalg.inv(config._vio_T_imu_cam0__)
    b[IMU_R_IC:IMU_R_IC + 9] = T_cam0_imu[:3, :3].T.ravel()
    b[IMU_T_CI:IMU_T_CI + 3] = T_cam0_imu[:3, 3]
    b[IMU_GRAVITY:IMU_GRAVITY + 3] = config._vio_gravity__
    return b


class MSCKFBatch(object):
    """F independent filters (Monte-Carlo seeds / sequences) stepped together on one GPU.

    This is the throughput form of the same back end: one context owns all F filters and every
    stage is one launch over the batch (grid.y / blockIdx = filter).  Filters may receive
    different numbers of IMU samples and features per frame (ragged offsets)."""

    GROUP_MIN_FILTERS = 512     # from this batch size on the filters are spread over two contexts (see __init__)

    def __init__(self, config, n_filters, device=0, feature_capacity=1024, max_rows=0, jacobian_pool_bytes=0, groups=None):
        """``groups``: number of contexts (CUDA streams) the filters are spread over.  A launch over a batch lasts as long as
        its longest filter; with two half-batches in flight one half's CTAs fill the other's tails (measured on B200: 1.10 ->
        0.97 ms per frame at 1024 filters, 0.65 -> 0.62 at 512; 0.48 -> 0.50 at 256, and three or four groups gain nothing).
        None = 2 from GROUP_MIN_FILTERS filters on, else 1."""
        self.config = config
        self.F = int(n_filters)
        if groups is None:
            groups = 2 if self.F >= self.GROUP_MIN_FILTERS else 1
        groups = max(1, min(int(groups), self.F))
        self.groups = groups
        sizes = [self.F // groups + (1 if g < self.F % groups else 0) for g in range(groups)]
        mk = lambda nf: capi.make_cconfig(config, n_filters=nf, feature_capacity=feature_capacity, max_rows=max_rows,   # noqa: E731
                                          device=device, jacobian_pool_bytes=jacobian_pool_bytes)
        self.cconfig = mk(sizes[0])
        self.ctx = capi.Context(self.cconfig) if groups == 1 else capi.ContextGroup([mk(nf) for nf in sizes])
        b0 = initial_imu_block(config)
        for f in range(self.F):
            self.ctx.set_imu_state(f, b0)
        self.next_imu_id = 0
        self.imu_time = None
        self.is_gravity_set = False
        self.is_first_img = True
        self.T_imu_body = Isometry3d(np.asarray(config._vio_T_imu_body__)[:3, :3], np.asarray(config._vio_T_imu_body__)[:3, 3])

    def initialize_gravity_and_bias(self, gyro, acc):
        """msckf.py:235-258 for every filter: gyro / acc are [F][K][3] static samples (the reference
        uses the first 200 buffered messages, msckf.py:175-178)."""
        gyro = np.asarray(gyro, dtype=np.float64)
        acc = np.asarray(acc, dtype=np.float64)
        blocks = self.ctx.get_imu_states()
        for f in range(self.F):
            sum_w = np.zeros(3)
            sum_a = np.zeros(3)
            for k in range(gyro.shape[1]):
                sum_w += gyro[f, k]
                sum_a += acc[f, k]
            b = blocks[f]
            b[IMU_BG:IMU_BG + 3] = sum_w / gyro.shape[1]
            g_imu = sum_a / gyro.shape[1]
            gravity = np.array([0., 0., -np.linalg.norm(g_imu)])
            b[IMU_GRAVITY:IMU_GRAVITY + 3] = gravity
            b[IMU_Q:IMU_Q + 4] = from_two_vectors(-gravity, g_imu)
            self.ctx.set_imu_state(f, b)
        self.is_gravity_set = True

    def select_imu(self, frame_time, imu_ts):
        """batch_imu_processing's message selection (msckf.py:267-281) on a timestamp array shared
        by all filters: boolean mask of the samples to integrate; advances the state time."""
        imu_ts = np.asarray(imu_ts, dtype=np.float64)
        if self.is_first_img:
            self.is_first_img = False
            self.imu_time = frame_time
            blocks = self.ctx.get_imu_states()
            for f in range(self.F):
                blocks[f, IMU_TIME] = frame_time
                self.ctx.set_imu_state(f, blocks[f])
        take = (imu_ts >= self.imu_time) & (imu_ts <= frame_time)
        if take.any():
            self.imu_time = float(imu_ts[take][-1])
        return take

    def feature_callback(self, frame_time, imu_ts, gyro, acc, ids, uv):
        """Lock-step frame: all filters share the buffered IMU timestamps ``imu_ts`` [K] (every not yet
        consumed sample up to ``frame_time``, like the reference's imu_msg_buffer) and the feature ids
        ``ids`` [n]; gyro/acc are [F][K][3], uv is [F][n][4].
        Returns (imu blocks [F][48] as publish() sees them, did_reset [F])."""
        self.feature_callback_async(frame_time, imu_ts, gyro, acc, ids, uv)
        return self.ctx.collect()

    def feature_callback_async(self, frame_time, imu_ts, gyro, acc, ids, uv):
        F = self.F
        take = self.select_imu(frame_time, imu_ts)
        ts = np.asarray(imu_ts, dtype=np.float64)[take]
        gyro = np.asarray(gyro, dtype=np.float64)[:, take]
        acc = np.asarray(acc, dtype=np.float64)[:, take]
        K, n = len(ts), len(ids)
        imu_off = np.arange(F + 1, dtype=np.int32) * K
        feat_off = np.arange(F + 1, dtype=np.int32) * n
        cam_ids = np.full(F, self.next_imu_id, dtype=np.int64)
        self.next_imu_id += 1
        self.ctx.feature_callback_async(imu_off, np.tile(ts, F), gyro, acc, cam_ids, feat_off,
                                        np.tile(np.asarray(ids, dtype=np.int64), F), uv)

    def feature_callback_ragged(self, frame_time, imu_ts, gyro, acc, ids, uv):
        """One frame with per-filter messages (filters that follow different sequences): ``imu_ts[f]`` [K_f], ``gyro[f]`` /
        ``acc[f]`` [K_f][3] are filter f's not yet consumed IMU samples up to ``frame_time[f]`` (a scalar serves all),
        ``ids[f]`` [n_f] / ``uv[f]`` [n_f][4] its feature message.  The message selection of batch_imu_processing
        (msckf.py:267-281) runs per filter.  Returns (imu blocks [F][48], did_reset [F])."""
        F = self.F
        ft = np.broadcast_to(np.asarray(frame_time, dtype=np.float64), (F,))
        if self.is_first_img:
            self.is_first_img = False
            self.imu_time = None
            self._imu_time_f = ft.copy()
            blocks = self.ctx.get_imu_states()
            for f in range(F):
                blocks[f, IMU_TIME] = ft[f]
                self.ctx.set_imu_state(f, blocks[f])
        elif getattr(self, '_imu_time_f', None) is None:
            self._imu_time_f = np.full(F, self.imu_time, dtype=np.float64)
        imu_off, feat_off, ts_l, gy_l, ac_l = [0], [0], [], [], []
        for f in range(F):
            t = np.asarray(imu_ts[f], dtype=np.float64)
            take = (t >= self._imu_time_f[f]) & (t <= ft[f])
            if take.any():
                self._imu_time_f[f] = float(t[take][-1])
            ts_l.append(t[take])
            gy_l.append(np.asarray(gyro[f], dtype=np.float64).reshape(-1, 3)[take])
            ac_l.append(np.asarray(acc[f], dtype=np.float64).reshape(-1, 3)[take])
            imu_off.append(imu_off[-1] + int(take.sum()))
            feat_off.append(feat_off[-1] + len(ids[f]))
        cam_ids = np.full(F, self.next_imu_id, dtype=np.int64)
        self.next_imu_id += 1
        cat = lambda xs, shape: np.concatenate(xs) if len(xs) and sum(len(x) for x in xs) else np.zeros(shape)   # noqa: E731
        return self.ctx.feature_callback(imu_off, cat(ts_l, (0,)), cat(gy_l, (0, 3)), cat(ac_l, (0, 3)), cam_ids, feat_off,
                                         np.concatenate([np.asarray(i, dtype=np.int64) for i in ids]),
                                         np.concatenate([np.asarray(u, dtype=np.float64).reshape(-1, 4) for u in uv]))

    def collect(self):
        return self.ctx.collect()

    def device_frame_parts(self, frame, keys):
        """For messages resident in HBM (``frame[k]``: device tensors in the layout of feature_callback_async, per-filter
        arrays with the filter as leading dimension or flat): the device addresses each context takes, precomputed once
        per frame -- [(addresses of group g in the order of ``keys``)].  The ragged offsets of a group are rebased on the
        device."""
        if self.groups == 1:
            return [tuple(frame[k].data_ptr() for k in keys)]
        io, fo = frame['imu_off'], frame['feat_off']
        ioh, foh = io.cpu().numpy(), fo.cpu().numpy()
        keep = frame.setdefault('_group_offsets', [])
        out = []
        for a, b in self.ctx.ranges():
            gi, gf = (io[a:b + 1] - io[a]).contiguous(), (fo[a:b + 1] - fo[a]).contiguous()
            keep.append((gi, gf))                          # the rebased offsets live as long as the frame
            i0, i1, f0, f1 = int(ioh[a]), int(ioh[b]), int(foh[a]), int(foh[b])
            part = {'imu_off': gi, 'feat_off': gf, 'ts': frame['ts'].reshape(-1)[i0:i1], 'gyro': frame['gyro'].reshape(-1, 3)[i0:i1],
                    'acc': frame['acc'].reshape(-1, 3)[i0:i1], 'cam_ids': frame['cam_ids'][a:b],
                    'ids': frame['ids'].reshape(-1)[f0:f1], 'uv': frame['uv'].reshape(-1, 4)[f0:f1]}
            out.append(tuple(part[k].data_ptr() for k in keys))
        return out

    def feature_callback_dev(self, parts, imu_out=None):
        """Enqueues one frame from device-resident messages (``parts`` from device_frame_parts); ``imu_out``: device tensor
        [F][48] that receives the IMU blocks after the frame, or None."""
        if self.groups == 1:
            self.ctx.feature_callback_dev(*parts[0], imu_out=(imu_out.data_ptr() if imu_out is not None else 0))
        else:
            outs = [imu_out[a:b].data_ptr() for a, b in self.ctx.ranges()] if imu_out is not None else None
            self.ctx.feature_callback_dev_parts(parts, outs)

    def close(self):
        self.ctx.close()
