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
