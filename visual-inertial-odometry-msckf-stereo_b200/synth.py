"""Synthetic EuRoC-MH_01-shaped stream: 200 Hz IMU + 20 Hz stereo feature
messages generated from a known trajectory and landmark field.  The dataset itself is
not available offline, so every test and benchmark in this repo runs on this stream.

The *structure* (trajectory, landmark field, which landmark is tracked in which frame,
persistent feature ids) depends on ``structure_seed`` only; the *noise* (IMU noise,
pixel noise, gross outliers) depends on ``noise_seed`` only, so Monte-Carlo filter k is
``realize(structure, noise_seed=k)`` and all filters of a batch share timestamps and ids.

Message shapes follow the reference: ``imu_msg(vio_timestamp__, angular_velocity,
linear_acceleration)`` (dataset.py:56-57) and ``vio_feature_msg__(timestamp,
vio_features[{id,u0,v0,u1,v1}])`` (image.py:425-438), coordinates normalised.
"""
from collections import namedtuple

import numpy as np

from .config import ConfigEuRoC
from .utils import to_rotation

ImuMsg = namedtuple('imu_msg', ['vio_timestamp__', 'angular_velocity', 'linear_acceleration'])
FeatureMsg = namedtuple('vio_feature_msg__', ['timestamp', 'vio_features'])


class FeatureMeasurement(object):
    __slots__ = ('id', 'u0', 'v0', 'u1', 'v1')

    def __init__(self, fid, u0, v0, u1, v1):
        self.id, self.u0, self.v0, self.u1, self.v1 = fid, u0, v0, u1, v1


IMU_DT = 0.005
FRAME_DT = 0.05
T_START = 1.5          # static lead-in (>= 200 IMU messages for gravity/bias initialisation)
GRAVITY = 9.81


def _phase(t, speed=1.0):
    tau = np.maximum(np.asarray(t, dtype=np.float64) - T_START, 0.0)
    return speed * 0.4 * (tau - np.tanh(tau))


def truth_position(t, speed=1.0):
    ph = _phase(t, speed)
    return np.stack([2.0 * np.cos(ph) - 2.0, 2.0 * np.sin(ph), 0.3 * np.sin(0.7 * ph)], axis=-1)


def truth_yaw(t, speed=1.0):
    return 0.5 * _phase(t, speed)


def _rot_z(a):
    c, s = np.cos(a), np.sin(a)
    R = np.zeros(np.shape(a) + (3, 3))
    R[..., 0, 0], R[..., 0, 1], R[..., 1, 0], R[..., 1, 1], R[..., 2, 2] = c, -s, s, c, 1.0
    return R


class Structure(object):
    """Noise-free stream."""
    pass


def make_structure(n_frames=200, feats_per_frame=300, structure_seed=0, mean_track_len=8.0,
                   n_landmarks=40000, config=None, speed=1.0):
    """``speed`` scales the trajectory's phase rate (1.0: about 0.8 m/s).  Above ~8 the camera moves more than the
    reference's key-frame distance (0.4 m, msckf.py:717) between consecutive frames, so pruning removes the OLDEST
    camera states instead of the ones next to the key frame and lost landmarks carry long tracks."""
    cfg = config or ConfigEuRoC()
    rng = np.random.default_rng(structure_seed)
    st = Structure()
    st.n_frames = n_frames
    st.speed = speed
    t_end = T_START + FRAME_DT * (n_frames + 1)
    n_imu = int(round(t_end / IMU_DT)) + 1
    ts = np.arange(n_imu) * IMU_DT
    h = 1e-4
    acc_w = (truth_position(ts + h, speed) - 2 * truth_position(ts, speed) + truth_position(ts - h, speed)) / (h * h)
    yaw_rate = (truth_yaw(ts + h, speed) - truth_yaw(ts - h, speed)) / (2 * h)
    R_wi = _rot_z(truth_yaw(ts, speed))                # IMU -> world
    g_w = np.array([0.0, 0.0, -GRAVITY])
    st.imu_ts = ts
    st.gyro_clean = np.stack([np.zeros(n_imu), np.zeros(n_imu), yaw_rate], axis=-1)
    st.acc_clean = np.einsum('kji,kj->ki', R_wi, acc_w - g_w)      # R_wi^T (a_w - g)
    st.gyro_bias = np.array([0.002, -0.001, 0.0015])
    st.acc_bias = np.array([0.02, -0.01, 0.015])

    # extrinsics: p_c0 = R_ci p_i + t_ci ; p_c1 = R_01 p_c0 + t_01
    R_ci, t_ci = cfg._vio_T_imu_cam0__[:3, :3], cfg._vio_T_imu_cam0__[:3, 3]
    R_01, t_01 = cfg._vio_T_cn_cnm1__[:3, :3], cfg._vio_T_cn_cnm1__[:3, 3]

    lm = np.stack([rng.uniform(-12, 8, n_landmarks), rng.uniform(-8, 12, n_landmarks),
                   rng.uniform(3, 8, n_landmarks)], axis=-1)
    st.landmarks = lm
    st.frame_ts = T_START + FRAME_DT * (np.arange(n_frames) + 1)
    st.frame_ids, st.frame_lm, st.frame_uv = [], [], []
    st.truth_p = truth_position(st.frame_ts, speed)
    st.truth_yaw = truth_yaw(st.frame_ts, speed)

    tracked_lm = np.zeros(0, dtype=np.int64)
    tracked_id = np.zeros(0, dtype=np.int64)
    next_id = 0
    survive = 1.0 - 1.0 / mean_track_len
    # a scalar, or one target per frame (e.g. a cheap lead-in that fills the camera window before the full-size frames)
    target = np.broadcast_to(np.asarray(feats_per_frame, dtype=np.int64), (n_frames,))
    for f in range(n_frames):
        t = st.frame_ts[f]
        R = _rot_z(truth_yaw(t, speed))
        p_i = (lm - truth_position(t, speed)) @ R             # R^T (X - p)
        p_c0 = p_i @ R_ci.T + t_ci
        p_c1 = p_c0 @ R_01.T + t_01
        uv = np.concatenate([p_c0[:, :2] / p_c0[:, 2:3], p_c1[:, :2] / p_c1[:, 2:3]], axis=1)
        vis = ((p_c0[:, 2] > 0.5) & (p_c1[:, 2] > 0.5)
               & (np.abs(uv[:, 0]) < 0.75) & (np.abs(uv[:, 1]) < 0.5)
               & (np.abs(uv[:, 2]) < 0.75) & (np.abs(uv[:, 3]) < 0.5))
        keep = vis[tracked_lm] & (rng.random(len(tracked_lm)) < survive)
        tracked_lm, tracked_id = tracked_lm[keep], tracked_id[keep]
        need = int(target[f]) - len(tracked_lm)
        if need > 0:
            cand = vis.copy()
            cand[tracked_lm] = False
            cand = np.flatnonzero(cand)
            new = rng.choice(cand, size=min(need, len(cand)), replace=False)
            tracked_lm = np.concatenate([tracked_lm, new])
            tracked_id = np.concatenate([tracked_id, next_id + np.arange(len(new))])
            next_id += len(new)
        st.frame_ids.append(tracked_id.copy())
        st.frame_lm.append(tracked_lm.copy())
        st.frame_uv.append(uv[tracked_lm].copy())
    return st


class Stream(object):
    """One noisy realisation: arrays plus reference-format message views."""

    def __init__(self, st, gyro, acc, frame_uv):
        self.structure = st
        self.imu_ts, self.gyro, self.acc = st.imu_ts, gyro, acc
        self.frame_ts, self.frame_ids, self.frame_uv = st.frame_ts, st.frame_ids, frame_uv
        self.n_frames = st.n_frames

    def imu_msgs(self):
        return [ImuMsg(float(t), self.gyro[k].copy(), self.acc[k].copy()) for k, t in enumerate(self.imu_ts)]

    def feature_msg(self, f):
        ids, uv = self.frame_ids[f], self.frame_uv[f]
        feats = [FeatureMeasurement(int(ids[i]), float(uv[i, 0]), float(uv[i, 1]), float(uv[i, 2]), float(uv[i, 3]))
                 for i in range(len(ids))]
        return FeatureMsg(float(self.frame_ts[f]), feats)

    def imu_range_for_frame(self, f):
        """Index range [a, b) of IMU samples with frame_ts[f-1] < ts <= frame_ts[f] (all earlier for f=0)."""
        b = int(np.searchsorted(self.imu_ts, self.frame_ts[f], side='right'))
        a = 0 if f == 0 else int(np.searchsorted(self.imu_ts, self.frame_ts[f - 1], side='right'))
        return a, b


def realize(st, noise_seed=0, sigma_z=0.001, outlier_frac=0.05, outlier_sigma=0.05, imu_noise_scale=1.0,
            config=None):
    cfg = config or ConfigEuRoC()
    rng = np.random.default_rng(noise_seed)
    sg = np.sqrt(cfg._vio_gyro_noise__ / IMU_DT) * imu_noise_scale
    sa = np.sqrt(cfg._vio_acc_noise__ / IMU_DT) * imu_noise_scale
    gyro = st.gyro_clean + st.gyro_bias + sg * rng.standard_normal(st.gyro_clean.shape)
    acc = st.acc_clean + st.acc_bias + sa * rng.standard_normal(st.acc_clean.shape)
    frame_uv = []
    for uv in st.frame_uv:
        n = sigma_z * rng.standard_normal(uv.shape)
        out = rng.random(len(uv)) < outlier_frac
        n[out] += outlier_sigma * rng.standard_normal((int(out.sum()), 4))
        frame_uv.append(uv + n)
    return Stream(st, gyro, acc, frame_uv)


def make_stream(n_frames=200, feats_per_frame=300, structure_seed=0, noise_seed=0, **kw):
    return realize(make_structure(n_frames, feats_per_frame, structure_seed), noise_seed, **kw)


def drive(filt, stream, n_frames=None, on_frame=None):
    """Single-threaded stepper (SURVEY.md section 8c hazard 2): push every IMU message with ts <= frame
    ts through imu_callback, then feature_callback; works with the reference, the oracle and
    the CUDA drop-in alike.  Returns the list of published results (None before gravity init)."""
    msgs = stream.imu_msgs()
    k = 0
    out = []
    nf = stream.n_frames if n_frames is None else n_frames
    for f in range(nf):
        t = stream.frame_ts[f]
        while k < len(msgs) and msgs[k].vio_timestamp__ <= t:
            filt.imu_callback(msgs[k])
            k += 1
        res = filt.feature_callback(stream.feature_msg(f))
        out.append(res)
        if on_frame is not None:
            on_frame(f, filt, res)
    return out


def stress_tracks(cam_q, cam_p, R01, t01, n_landmarks, rng, sigma_z=0.001, outlier_frac=0.05, outlier_sigma=0.05):
    """BASELINE.json configs[2] (feature-update stress frame): ``n_landmarks`` synthetic tracks over the camera states
    (cam_q [N][4] JPL world->cam0, cam_p [N][3]) except the newest one, so that every landmark is "lost" in the current
    frame.  World points lie in front of the middle camera and are projected into cam0 / cam1 of every observing state
    (msckf.py:443-457) plus pixel noise and gross outliers.  Track lengths M ~ U{3 .. N-1}; nine in ten tracks are
    contiguous in time (as real tracks are), the rest observe a random subset of the window.
    Returns (ids int64[n], masks uint64[n] with bit j = j-th camera state, obs [n][N][4])."""
    cam_q, cam_p = np.asarray(cam_q, dtype=np.float64), np.asarray(cam_p, dtype=np.float64)
    N = len(cam_q)
    Rs = [to_rotation(q) for q in cam_q]
    mid = (N - 1) // 2
    depth = rng.uniform(3.0, 8.0, n_landmarks)
    pc = np.stack([rng.uniform(-0.4, 0.4, n_landmarks) * depth, rng.uniform(-0.3, 0.3, n_landmarks) * depth, depth], axis=1)
    pw = pc @ Rs[mid] + cam_p[mid]                         # R^T p_c + t
    ids = 10 ** 6 + np.arange(n_landmarks, dtype=np.int64)
    masks = np.zeros(n_landmarks, dtype=np.uint64)
    obs = np.zeros((n_landmarks, N, 4))
    Ms = rng.integers(3, N, n_landmarks)                   # 3 .. N-1
    for j in range(n_landmarks):
        M = int(Ms[j])
        if rng.random() < 0.9:
            s = int(rng.integers(0, N - 1 - M + 1))
            sel = np.arange(s, s + M)
        else:
            sel = np.sort(rng.choice(N - 1, size=M, replace=False))
        m = 0
        for c in sel:
            p0 = Rs[c] @ (pw[j] - cam_p[c])
            p1 = R01 @ p0 + t01
            z = np.array([p0[0] / p0[2], p0[1] / p0[2], p1[0] / p1[2], p1[1] / p1[2]])
            z += sigma_z * rng.standard_normal(4)
            if rng.random() < outlier_frac:
                z += outlier_sigma * rng.standard_normal(4)
            obs[j, c] = z
            m |= 1 << int(c)
        masks[j] = m
    return ids, masks, obs
