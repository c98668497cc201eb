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


def block_rel_err(P, Pref):
    """Per-block error of a covariance (IMU core 0:15, extrinsics 15:21, every camera state 6 wide; diagonal and cross
    blocks).  Returns (scaled, strict):
      scaled  max over blocks of |dP_ij|_F / sqrt(|Pref_ii|_F |Pref_jj|_F)  -- every block on its natural scale
              (|P_ij| <= sqrt(P_ii P_jj)), so a small block next to a large one is not hidden by the norm of all of P;
      strict  max over the blocks that are not numerically zero (|Pref_ij| >= 1e-6 of that scale) of |dP_ij|_F / |Pref_ij|_F.
    """
    P, Pref = np.asarray(P, float), np.asarray(Pref, float)
    if P.shape != Pref.shape:
        return float('inf'), float('inf')
    n = P.shape[0]
    edges = [0, 15, 21] + list(range(27, n + 1, 6))
    nb = len(edges) - 1
    D = P - Pref

    def fro_blocks(M):
        sq = np.add.reduceat(np.add.reduceat(M * M, edges[:-1], axis=0), edges[:-1], axis=1)
        return np.sqrt(sq)

    nd, nr = fro_blocks(D), fro_blocks(Pref)
    diag = np.maximum(np.diag(nr), 1e-300)
    scale = np.sqrt(np.outer(diag, diag))
    scaled = float(np.max(nd / scale))
    sig = nr >= 1e-6 * scale
    strict = float(np.max(nd[sig] / nr[sig])) if sig.any() else 0.0
    assert nb >= 2
    return scaled, strict


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


def device_decisions(filt):
    """(gate decisions, gammas, removed cam ids or None, [update rows]) of the frame just processed."""
    from msckf_b200 import capi
    dec, gam, upd = [], [], []
    rm = None
    for stage in (0, 1):
        log = filt.stage_log(stage)
        for s, g in zip(log['status'], log['gamma']):
            if s in (capi.FEAT_GATE_PASS, capi.FEAT_GATE_FAIL):
                dec.append(s == capi.FEAT_GATE_PASS)
                gam.append(g)
        if log['update_rows'] > 0:
            upd.append(log['update_rows'])
        if stage == 1 and log['removed_cam_ids'][0] >= 0:
            rm = [int(x) for x in log['removed_cam_ids']]
    return dec, gam, rm, upd


def run_pair(stream, max_cam, n_frames, feature_capacity=1024, row_cap=None, before_frame=None, after_frame=None,
             check_P_every=1):
    """Steps the oracle and the device filter side by side through feature_callback; checks ids and decisions at every
    frame and returns the worst errors seen.

      worst['P']        |dP| / |P| over the whole covariance (north_star: 1e-9)
      worst['P_block']  the same per block on its natural scale, worst['P_block_strict'] relative to the block itself
                        (block_rel_err above)
      worst['gate_margin']  smallest |gamma - threshold| / threshold over all gates: how close any decision came to flipping
      counts['not_spd'] landmarks the device rejected because its innovation covariance was not SPD (must be 0: the
                        reference's LU would have produced some gamma there)
    before_frame(f, oracle_filter, device_filter) may perturb both filters before frame f (forced resets etc.);
    after_frame(f, oracle_filter, device_filter) may read further diagnostics."""
    from oracle import msckf_oracle as orc
    from msckf_b200 import capi
    from msckf_b200.config import ConfigEuRoC
    from msckf_b200.msckf import MSCKF
    ocfg = orc.ConfigEuRoC()
    ocfg._vio_max_cam_state_size__ = max_cam
    of = orc.MSCKF(ocfg)
    if row_cap is not None:
        of.ROW_CAP = row_cap
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    if row_cap is not None:
        dcfg.row_cap = row_cap if row_cap < 10 ** 8 else 0      # <= 0 lifts the cap in the library
    df = MSCKF(dcfg, feature_capacity=feature_capacity)
    worst = {'imu': 0.0, 'P': 0.0, 'P_block': 0.0, 'P_block_strict': 0.0, 'cams': 0.0, 'extr': 0.0, 'pos_m': 0.0,
             'gate_margin': float('inf')}
    counts = {'gates': 0, 'pass': 0, 'prunes': 0, 'updates': 0, 'resets': 0, 'not_spd': 0, 'max_update_rows': 0}
    msgs = stream.imu_msgs()
    k = 0
    try:
        for f in range(n_frames):
            if before_frame is not None:
                before_frame(f, of, df)
            t = stream.frame_ts[f]
            while k < len(msgs) and msgs[k].vio_timestamp__ <= t:
                of.imu_callback(msgs[k])
                df.imu_callback(msgs[k])
                k += 1
            of.log = {}
            ro = of.feature_callback(stream.feature_msg(f))
            rd = df.feature_callback(stream.feature_msg(f))
            assert (ro is None) == (rd is None)
            with_P = (f % check_P_every == 0) or f == n_frames - 1
            so, sd = snapshot(of, with_P), snapshot(df, with_P)
            assert np.array_equal(sd['cam_ids'], so['cam_ids']), f
            assert np.array_equal(sd['map_ids'], so['map_ids']), f
            for key in ('imu', 'cams', 'extr'):
                worst[key] = max(worst[key], rel_err(sd[key], so[key]))
            if with_P:
                worst['P'] = max(worst['P'], rel_err(sd['P'], so['P']))
                bs, bt = block_rel_err(sd['P'], so['P'])
                worst['P_block'] = max(worst['P_block'], bs)
                worst['P_block_strict'] = max(worst['P_block_strict'], bt)
            worst['pos_m'] = max(worst['pos_m'], float(np.linalg.norm(sd['imu'][13:16] - so['imu'][13:16])))
            dec, gam, rm, upd = device_decisions(df)
            ogates = of.log.get('gate', [])
            assert dec == [x[2] for x in ogates], f
            if ogates:
                og = np.array([x[0] for x in ogates])
                thr = np.array([x[1] for x in ogates])
                assert np.all(np.abs(np.array(gam) - og) <= 1e-7 * np.abs(og)), f
                worst['gate_margin'] = min(worst['gate_margin'], float(np.min(np.abs(og - thr) / thr)))
            orm = of.log.get('rm_cam_ids', [])
            assert (rm is None and not orm) or (orm and rm == [int(x) for x in orm[0]]), f
            did_reset = bool(of.log.get('online_reset'))
            assert bool(df.did_reset) == did_reset, f                     # online_reset decision (msckf.py:859-886)
            counts['gates'] += len(dec)
            counts['pass'] += sum(dec)
            counts['prunes'] += int(rm is not None)
            counts['updates'] += len(upd)
            counts['resets'] += int(did_reset)
            counts['max_update_rows'] = max([counts['max_update_rows']] + list(upd))
            if after_frame is not None:
                after_frame(f, of, df)
            if ro is not None:
                assert rel_err(rd.pose._vio_t__, ro.pose._vio_t__) < 1e-8 or np.linalg.norm(ro.pose._vio_t__) < 1e-6
                assert rel_err(rd.cam0_pose._vio_R__, ro.cam0_pose._vio_R__) < 1e-9
                assert rel_err(rd.velocity, ro.velocity) < 1e-8 or np.linalg.norm(ro.velocity) < 1e-6
        counts['not_spd'] = int(df.ctx.counters()[:, capi.CNT_GATE_NOT_SPD].sum())
    finally:
        df.close()
    assert counts['not_spd'] == 0, counts
    return worst, counts
