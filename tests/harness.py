"""Shared parity harness: reads filter state through the reference's attribute names, so the
same code snapshots the reference (imported from /root/reference when present), the CPU oracle
and the CUDA drop-in."""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

REFERENCE_DIR = '/root/reference/MSCKF'


def reference_available():
    return os.path.isdir(REFERENCE_DIR)


def import_reference():
    """Import the unmodified reference (build container only).  Returns (msckf module, ConfigEuRoC)."""
    os.environ.setdefault('OPENBLAS_NUM_THREADS', '1')
    if REFERENCE_DIR not in sys.path:
        sys.path.insert(0, REFERENCE_DIR)
    import msckf as ref_msckf            # noqa
    from config import ConfigEuRoC as RefConfig   # noqa
    return ref_msckf, RefConfig


def reset_reference_globals(ref_msckf):
    """SURVEY.md 8c hazard 1: class-level ids / gravity are process-global in the reference."""
    ref_msckf.IMUState._vio_next_id__ = 0
    ref_msckf.IMUState._vio_gravity__ = np.array([0., 0., -9.81])


def snapshot(filt, with_P=True):
    ss = filt.state_server
    imu = ss._vio_imu_state__
    cams = ss._vio_cam_states__
    snap = {
        'imu': np.concatenate([np.asarray(imu.orientation, float), np.asarray(imu._vio_gyro_bias__, float),
                               np.asarray(imu.velocity, float), np.asarray(imu.acc_bias, float),
                               np.asarray(imu._vio_position__, float)]),
        'extr': np.concatenate([np.asarray(imu.R_imu_cam0, float).ravel(), np.asarray(imu.t_cam0_imu, float)]),
        'cam_ids': np.array(list(cams.keys()), dtype=np.int64),
        'cams': (np.array([np.concatenate([c.orientation, c._vio_position__]) for c in cams.values()])
                 if len(cams) else np.zeros((0, 7))),
        'map_ids': np.array(list(filt.map_server.keys()), dtype=np.int64),
    }
    if with_P:
        snap['P'] = np.array(ss._vio_state_cov__, dtype=float, copy=True)
    return snap


def rel_err(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    if a.shape != b.shape:
        return float('inf')
    d = np.linalg.norm(a - b)
    s = max(np.linalg.norm(b), 1e-300)
    return d / s


def compare_snapshots(a, b):
    """Returns dict of relative errors (state vector and P, norm-wise) + exact-match flags."""
    out = {
        'imu': rel_err(a['imu'], b['imu']),
        'extr': rel_err(a['extr'], b['extr']),
        'cams': rel_err(a['cams'], b['cams']) if a['cams'].size or b['cams'].size else 0.0,
        'cam_ids_equal': bool(np.array_equal(a['cam_ids'], b['cam_ids'])),
        'map_ids_equal': bool(np.array_equal(a['map_ids'], b['map_ids'])),
    }
    if 'P' in a and 'P' in b:
        out['P'] = rel_err(a['P'], b['P'])
    return out


def oracle_filter_from_units(u, max_cam=12):
    """Re-create the mid-stream filter state stored in units.npz inside an oracle MSCKF."""
    from oracle import msckf_oracle as orc
    cfg = orc.ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = orc.MSCKF(cfg)
    imu = filt.state_server._vio_imu_state__
    s = u['state_imu']
    imu.orientation, imu._vio_gyro_bias__, imu.velocity = s[0:4].copy(), s[4:7].copy(), s[7:10].copy()
    imu.acc_bias, imu._vio_position__ = s[10:13].copy(), s[13:16].copy()
    nl = u['state_imu_null']
    imu.orientation_null = nl[0:4].copy()
    # reference aliasing (msckf.py:366-368): after process_model the v/p null members ARE the live arrays
    assert np.array_equal(nl[4:7], imu.velocity) and np.array_equal(nl[7:10], imu._vio_position__)
    imu.velocity_null, imu.position_null = imu.velocity, imu._vio_position__
    imu.R_imu_cam0, imu.t_cam0_imu = u['state_extr'][:9].reshape(3, 3).copy(), u['state_extr'][9:].copy()
    imu.id = int(u['state_imu_id'])
    filt.gravity = u['j_gravity'].copy()
    for cid, row in zip(u['state_cam_ids'], u['state_cams']):
        c = orc.CAMState(int(cid))
        c.orientation, c._vio_position__ = row[0:4].copy(), row[4:7].copy()
        c.orientation_null = row[7:11].copy()
        c.position_null = c._vio_position__          # aliased in the reference (msckf.py:400)
        filt.state_server._vio_cam_states__[int(cid)] = c
    filt.state_server._vio_state_cov__ = u['state_P'].copy()
    filt.tracking_rate = float(u['state_tracking_rate'])
    off = u['f_off']
    for k, fid in enumerate(u['f_ids']):
        ft = orc.Feature(int(fid), filt.optimization_config, filt.R_cam0_cam1, filt.t_cam0_cam1)
        for j in range(off[k], off[k + 1]):
            ft.observations[int(u['f_obs_cam'][j])] = u['f_obs_z'][j].copy()
        filt.map_server[int(fid)] = ft
    return filt


def device_filter_from_units(u, max_cam=12, feature_capacity=256):
    """The same mid-stream state loaded into the CUDA drop-in (state upload through the C ABI's
    snapshot setters: IMU block, camera states, covariance, landmark map)."""
    from msckf_b200 import capi
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKF
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = MSCKF(cfg, feature_capacity=feature_capacity)
    b = filt.ctx.get_imu_state(0)
    s, nl, ex = u['state_imu'], u['state_imu_null'], u['state_extr']
    b[capi.IMU_Q:capi.IMU_Q + 4] = s[0:4]
    b[capi.IMU_BG:capi.IMU_BG + 3] = s[4:7]
    b[capi.IMU_V:capi.IMU_V + 3] = s[7:10]
    b[capi.IMU_BA:capi.IMU_BA + 3] = s[10:13]
    b[capi.IMU_P:capi.IMU_P + 3] = s[13:16]
    b[capi.IMU_Q_NULL:capi.IMU_Q_NULL + 4] = nl[0:4]
    b[capi.IMU_V_NULL:capi.IMU_V_NULL + 3] = nl[4:7]
    b[capi.IMU_P_NULL:capi.IMU_P_NULL + 3] = nl[7:10]
    b[capi.IMU_NULL_ALIASED] = 1.0
    b[capi.IMU_R_IC:capi.IMU_R_IC + 9] = ex[:9]
    b[capi.IMU_T_CI:capi.IMU_T_CI + 3] = ex[9:]
    b[capi.IMU_GRAVITY:capi.IMU_GRAVITY + 3] = u['j_gravity']
    b[capi.IMU_TIME] = 0.0
    b[capi.IMU_TRACKING_RATE] = float(u['state_tracking_rate'])
    filt.ctx.set_imu_state(0, b)
    cam_ids = [int(c) for c in u['state_cam_ids']]
    cams = np.zeros((len(cam_ids), capi.CAM_STRIDE))
    cams[:, :14] = u['state_cams']
    filt.ctx.set_cam_states(0, cam_ids, cams)
    filt.ctx.set_cov(0, u['state_P'])
    off = u['f_off']
    n = len(u['f_ids'])
    mask = np.zeros(n, dtype=np.uint64)
    obs = np.zeros((n, max_cam, 4))
    col = dict((c, i) for i, c in enumerate(cam_ids))
    for k in range(n):
        m = 0
        for j in range(off[k], off[k + 1]):
            c = col[int(u['f_obs_cam'][j])]
            m |= 1 << c
            obs[k, c] = u['f_obs_z'][j]
        mask[k] = m
    filt.ctx.set_features(0, u['f_ids'], mask, np.zeros(n, dtype=np.uint8), np.zeros((n, 3)), obs)
    filt._imu_id = int(u['state_imu_id'])
    filt._next_imu_id = filt._imu_id + 1
    filt._imu_time = 0.0
    filt.is_gravity_set = True
    filt.is_first_img = False
    filt.tracking_rate = float(u['state_tracking_rate'])
    return filt
