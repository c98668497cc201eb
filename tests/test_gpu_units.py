"""GPU parity, unit level: every piece of the hot path called through the C ABI and compared with
(a) the golden vectors recorded from the unmodified reference (tests/golden/units.npz) and
(b) the CPU oracle on the same inputs.

Tolerances (BASELINE.json north_star): FP64 state / covariance 1e-9 relative; decisions
(triangulation validity, gate accept / reject, camera states to prune) bit-exact.  The tests
use tighter bounds where the arithmetic allows, so drift is noticed early.
"""
import copy
import os

import numpy as np
import pytest

from harness import rel_err, snapshot, device_filter_from_units, oracle_filter_from_units
from oracle import msckf_oracle as orc

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TOL_STATE = 1e-9        # north_star tolerance, state vector and covariance
TOL_TIGHT = 1e-11       # what the kernels actually achieve on single operations


@pytest.fixture(scope='module')
def units():
    return dict(np.load(os.path.join(GOLD, 'units.npz')))


@pytest.fixture()
def dev(units):
    filt = device_filter_from_units(units)
    yield filt
    filt.close()


def _csr_for(u, ks, z_override=None):
    cam_row = dict((int(c), row) for c, row in zip(u['state_cam_ids'], u['state_cams']))
    off, uv, q, p = [0], [], [], []
    j = 0
    for k in ks:
        for i in range(u['f_off'][k], u['f_off'][k + 1]):
            row = cam_row[int(u['f_obs_cam'][i])]
            uv.append(u['f_obs_z'][i] if z_override is None else z_override[j])
            q.append(row[0:4])
            p.append(row[4:7])
            j += 1
        off.append(len(uv))
    return np.array(off, np.int32), np.array(uv), np.array(q), np.array(p)


def test_library_reports_sm100():
    from msckf_b200 import capi
    assert 'sm_100a' in capi.version()


def test_triangulation_golden(units, dev):
    """K1: initialize_position (feature.py:167-295) vs the reference's recorded results."""
    u = units
    n = len(u['f_ids'])
    off, uv, q, p = _csr_for(u, range(n))
    pos, valid, solves = dev.ctx.triangulate(off, uv, q, p)
    assert np.array_equal(valid, u['f_tri_ok'].astype(bool))            # bit-exact flags
    for k in range(n):
        assert rel_err(pos[k], u['f_tri_pos'][k]) < 1e-10, k
    assert solves.min() >= 1
    # behind-camera cases: flags bit-exact, position written anyway
    nb = len(u['f_bad_ok'])
    off, uv, q, p = _csr_for(u, range(nb), z_override=u['f_bad_z'])
    pos, valid, _ = dev.ctx.triangulate(off, uv, q, p)
    assert np.array_equal(valid, u['f_bad_ok'].astype(bool))
    for k in range(nb):
        assert rel_err(pos[k], u['f_bad_pos'][k]) < 1e-9, k


def test_triangulation_lm_schedule_vs_oracle(units, dev):
    """Number of damped solves (lambda schedule, single-accept quirk) equals the oracle's."""
    u = units
    of = oracle_filter_from_units(u)
    cams = of.state_server._vio_cam_states__
    n = len(u['f_ids'])
    off, uv, q, p = _csr_for(u, range(n))
    rng = np.random.default_rng(3)
    uv = uv + 0.02 * rng.standard_normal(uv.shape)      # noisy tracks force rejected steps
    pos, valid, solves = dev.ctx.triangulate(off, uv, q, p)
    for k, fid in enumerate(u['f_ids']):
        ft = orc.Feature(int(fid), of.optimization_config, of.R_cam0_cam1, of.t_cam0_cam1)
        for j, i in enumerate(range(u['f_off'][k], u['f_off'][k + 1])):
            ft.observations[int(u['f_obs_cam'][i])] = uv[off[k] + j]
        ok = ft.initialize_position(cams)
        assert ok == bool(valid[k]), k
        assert ft.lm_solves == int(solves[k]), k
        assert rel_err(pos[k], ft._vio_position__) < 1e-8, k


def test_check_motion_and_initial_guess(units, dev):
    from msckf_b200 import capi
    from msckf_b200.config import ConfigEuRoC
    u = units
    of = oracle_filter_from_units(u)
    cams = of.state_server._vio_cam_states__
    n = 20
    off, uv, q, p = _csr_for(u, range(n))
    guess = dev.ctx.initial_guess(off, uv, q, p)
    for k in range(n):
        ft = of.map_server[int(u['f_ids'][k])]
        ids = list(ft.observations.keys())
        c0 = orc.Isometry3d(orc.to_rotation(cams[ids[0]].orientation).T, cams[ids[0]]._vio_position__)
        cl = orc.Isometry3d(orc.to_rotation(cams[ids[-1]].orientation).T, cams[ids[-1]]._vio_position__)
        g = ft.generate_initial_guess(cl.inverse() * c0, ft.observations[ids[0]][:2], ft.observations[ids[-1]][:2])
        assert rel_err(guess[k], g) < 1e-11, k
    # check_motion with a positive threshold: some pass, some fail -- decisions equal the oracle's
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = 12
    for thr in (0.01, 0.05, 0.2):
        cfg._vio_optimization_config__._vio_translation_threshold__ = thr
        ctx = capi.Context(capi.make_cconfig(cfg, n_filters=1, feature_capacity=8))
        moved = ctx.check_motion(off, uv, q, p)
        oc = orc.OptimizationConfigEuRoC()
        oc._vio_translation_threshold__ = thr
        exp = []
        for k in range(n):
            ft = copy.copy(of.map_server[int(u['f_ids'][k])])
            ft.optimization_config = oc
            exp.append(ft.check_motion(cams))
        assert list(moved) == exp
        ctx.close()


def test_feature_class_dropin(units, dev):
    """Feature.check_motion / generate_initial_guess / initialize_position with the reference's signatures."""
    from msckf_b200.feature import Feature, triangulate_batch
    from msckf_b200.config import OptimizationConfigEuRoC
    u = units
    cams = dev.state_server._vio_cam_states__
    feats = []
    for k in range(12):
        ft = Feature(int(u['f_ids'][k]), OptimizationConfigEuRoC())
        for i in range(u['f_off'][k], u['f_off'][k + 1]):
            ft.observations[int(u['f_obs_cam'][i])] = u['f_obs_z'][i].copy()
        feats.append(ft)
    for k, ft in enumerate(feats[:4]):
        assert ft.check_motion(cams) is True
        assert ft.initialize_position(cams) == bool(u['f_tri_ok'][k])
        assert rel_err(ft._vio_position__, u['f_tri_pos'][k]) < 1e-10
    valid = triangulate_batch(feats, cams)
    assert np.array_equal(valid, u['f_tri_ok'][:12].astype(bool))
    T = orc.Isometry3d(orc.to_rotation(np.array([0.01, -0.02, 0.03, 1.0])), np.array([0.1, -0.05, 0.02]))
    z1, z2 = np.array([0.1, -0.2]), np.array([0.08, -0.21])
    g = feats[0].generate_initial_guess(T, z1, z2)
    assert rel_err(g, orc.Feature(0, OptimizationConfigEuRoC()).generate_initial_guess(T, z1, z2)) < 1e-10


def test_measurement_and_feature_jacobian_golden(units, dev):
    """K2: measurement_jacobian element-wise; feature_jacobian through its invariants H^T H, H^T r, gamma
    (the null-space basis is not unique: SVD in the reference, Householder here)."""
    u = units
    for k, fid in enumerate(u['f_ids']):
        dev.ctx.set_feature_position(0, int(fid), u['f_tri_pos'][k], True)
    ms = dev.map_server
    for k in range(len(u['m_jac'])):
        fid = int(u['f_ids'][k])
        cid = list(ms[fid].observations.keys())[1]
        Hx, Hf, r = dev.measurement_jacobian(cid, fid)
        assert rel_err(np.concatenate([Hx.ravel(), Hf.ravel(), r]), u['m_jac'][k]) < 1e-12, k
    P = u['state_P']
    for k in range(len(u['fj_rows'])):
        fid = int(u['f_ids'][k])
        H, r = dev.feature_jacobian(fid, list(ms[fid].observations.keys()))
        assert len(r) == int(u['fj_rows'][k])
        assert np.all(H[:, :21] == 0.0)
        assert rel_err(H.T @ H, u['fj_HtH'][k]) < 1e-11, k
        assert rel_err(H.T @ r, u['fj_Htr'][k]) < 1e-10, k
        gamma, ok = dev.ctx.gating_test(0, H, r, len(r) // 4)
        assert abs(gamma - u['fj_gamma'][k]) <= 1e-9 * abs(u['fj_gamma'][k]), k
        S = H @ P @ H.T + 0.035 ** 2 * np.identity(len(H))
        assert ok == bool(r @ np.linalg.solve(S, r) < dev.chi_squared_test_table[len(r) // 4])


@pytest.mark.parametrize('tag', ['wide', 'tall'])
def test_measurement_update_golden(units, tag):
    """K3 + K4: measurement_update(H, r) with the reference's own H (wide: no QR; tall: QR)."""
    u = units
    dev = device_filter_from_units(u)
    try:
        dev.measurement_update(u['mu_%s_H' % tag], u['mu_%s_r' % tag])
        s = snapshot(dev)
        assert rel_err(s['imu'], u['mu_%s_imu' % tag]) < TOL_TIGHT
        assert rel_err(s['extr'], u['mu_%s_extr' % tag]) < TOL_TIGHT
        assert rel_err(s['cams'], u['mu_%s_cams' % tag]) < TOL_TIGHT
        assert rel_err(s['P'], u['mu_%s_P' % tag]) < TOL_STATE
        assert np.array_equal(s['P'], s['P'].T)
        i2 = dev.state_server._vio_imu_state__
        nul = np.concatenate([i2.orientation_null, i2.velocity_null, i2.position_null])
        assert rel_err(nul, u['mu_%s_null' % tag]) < TOL_TIGHT      # aliasing quirk (msckf.py:366-368)
    finally:
        dev.close()


def test_propagation_vs_oracle(units, dev):
    """K5: process_model / predict_new_state incl. the stale-skew RK4 quirk, several samples."""
    u = units
    of = oracle_filter_from_units(u)
    t0 = 100.0
    of.state_server._vio_imu_state__.timestamp = t0
    b = dev.ctx.get_imu_state(0)
    from msckf_b200.capi import IMU_TIME
    b[IMU_TIME] = t0
    dev.ctx.set_imu_state(0, b)
    dev._imu_time = t0
    rng = np.random.default_rng(5)
    ts = t0 + 0.005 * (1 + np.arange(10))
    gy = 0.2 * rng.standard_normal((10, 3))
    gy[3] = [1e-7, -2e-7, 1e-7]                     # small-rate branch (jit_utils.py:66-76)
    ac = np.array([0.1, -0.2, 9.8]) + 0.5 * rng.standard_normal((10, 3))
    for k in range(10):
        of.process_model(ts[k], gy[k], ac[k])
        of.state_server._vio_imu_state__.timestamp = ts[k]
    dev.ctx.propagate([0, 10], ts, gy, ac)
    so, sd = snapshot(of), snapshot(dev)
    assert rel_err(sd['imu'], so['imu']) < TOL_TIGHT
    assert rel_err(sd['P'], so['P']) < 1e-10
    assert np.array_equal(sd['P'], sd['P'].T)
    io, idv = of.state_server._vio_imu_state__, dev.state_server._vio_imu_state__
    assert rel_err(np.concatenate([idv.orientation_null, idv.velocity_null, idv.position_null]),
                   np.concatenate([io.orientation_null, io.velocity_null, io.position_null])) < TOL_TIGHT
    # jit_utils golden vectors (reference's own outputs): Phi / predict through a single step from rest
    for i in range(3):
        of2 = oracle_filter_from_units(u)
        d2 = device_filter_from_units(u)
        try:
            of2.state_server._vio_imu_state__.timestamp = 0.0
            b = d2.ctx.get_imu_state(0)
            b[IMU_TIME] = 0.0
            d2.ctx.set_imu_state(0, b)
            # zero biases so that gyro/acc equal the golden inputs
            dt = float(u['j_dt'][i])
            imu_o = of2.state_server._vio_imu_state__
            gyro_m = u['j_gyro'][i] + imu_o._vio_gyro_bias__
            acc_m = u['j_acc'][i] + imu_o.acc_bias
            d2.ctx.propagate([0, 1], [dt], gyro_m[None], acc_m[None])
            got = d2.ctx.get_imu_state(0)
            from msckf_b200.capi import IMU_Q, IMU_V, IMU_P
            pred = np.concatenate([got[IMU_Q:IMU_Q + 4], got[IMU_V:IMU_V + 3], got[IMU_P:IMU_P + 3]])
            assert rel_err(pred, u['j_pred'][i]) < 1e-12, i
        finally:
            d2.close()


def test_augmentation_vs_golden(units, dev):
    """K6: state_augmentation (jit_utils.py:137-167 golden) + the new camera state."""
    u = units
    of = oracle_filter_from_units(u)
    of.state_server._vio_imu_state__.id = 777
    of.state_augmentation(1.0)
    dev._imu_id = 777
    dev.state_augmentation(1.0)
    so, sd = snapshot(of), snapshot(dev)
    assert np.array_equal(sd['cam_ids'], so['cam_ids'])
    assert rel_err(sd['cams'], so['cams']) < 1e-13
    assert rel_err(sd['P'], so['P']) < 1e-12
    aug = u['j_aug']
    assert rel_err(sd['P'], (aug + aug.T) / 2.) < 1e-12
    assert np.array_equal(sd['P'], sd['P'].T)


def test_find_redundant_cam_states(units, dev):
    u = units
    of = oracle_filter_from_units(u)
    assert dev.find_redundant_cam_states() == of.find_redundant_cam_states()
    # force the "close to key frame" branch: make the two frames after the key frame coincide with it
    ids, cams = dev.ctx.get_cam_states(0)
    key = len(ids) - 4
    cams[key + 1, :7] = cams[key, :7]
    cams[key + 2, :7] = cams[key, :7]
    dev.ctx.set_cam_states(0, ids, cams)
    oc = of.state_server._vio_cam_states__
    keys = list(oc.keys())
    for j in (1, 2):
        oc[keys[key + j]].orientation = oc[keys[key]].orientation.copy()
        oc[keys[key + j]]._vio_position__ = oc[keys[key]]._vio_position__.copy()
    for rate in (0.9, 0.3):
        of.tracking_rate = rate
        b = dev.ctx.get_imu_state(0)
        from msckf_b200.capi import IMU_TRACKING_RATE
        b[IMU_TRACKING_RATE] = rate
        dev.ctx.set_imu_state(0, b)
        assert dev.find_redundant_cam_states() == of.find_redundant_cam_states()


def test_map_roundtrip(units, dev):
    """The device map reproduces ids, order, observation sets and values it was fed."""
    u = units
    ms = dev.map_server
    assert list(ms.keys()) == [int(x) for x in u['f_ids']]
    for k, fid in enumerate(u['f_ids']):
        ft = ms[int(fid)]
        sl = slice(u['f_off'][k], u['f_off'][k + 1])
        assert list(ft.observations.keys()) == [int(c) for c in u['f_obs_cam'][sl]]
        assert np.array_equal(np.array(list(ft.observations.values())), u['f_obs_z'][sl])


def test_remove_lost_features_vs_oracle(units, dev):
    """a21 end to end on the mid-stream state: invalid / processed id lists and gate decisions bit-exact,
    state and covariance after the update within tolerance."""
    u = units
    of = oracle_filter_from_units(u)
    of.log = {}
    of.remove_lost_features()
    dev.remove_lost_features()
    from msckf_b200 import capi
    log = dev.stage_log(0)
    st = log['status']
    invalid = [int(i) for i, s in zip(log['ids'], st) if s in (capi.FEAT_FEW_OBS, capi.FEAT_TRI_FAILED)]
    processed = [int(i) for i, s in zip(log['ids'], st) if s in (capi.FEAT_GATE_PASS, capi.FEAT_GATE_FAIL, capi.FEAT_AFTER_CAP)]
    assert invalid == of.log.get('lost_invalid', [[]])[0]
    assert processed == of.log.get('lost_processed', [[]])[0]
    gates = of.log.get('gate', [])
    dec = [s == capi.FEAT_GATE_PASS for s in st if s in (capi.FEAT_GATE_PASS, capi.FEAT_GATE_FAIL)]
    assert dec == [g[2] for g in gates]
    gam = np.array([g for g, s in zip(log['gamma'], st) if s in (capi.FEAT_GATE_PASS, capi.FEAT_GATE_FAIL)])
    ref = np.array([g[0] for g in gates])
    assert np.all(np.abs(gam - ref) <= 1e-8 * np.abs(ref))
    so, sd = snapshot(of), snapshot(dev)
    assert np.array_equal(sd['map_ids'], so['map_ids'])
    assert rel_err(sd['imu'], so['imu']) < TOL_STATE
    assert rel_err(sd['cams'], so['cams']) < TOL_STATE
    assert rel_err(sd['P'], so['P']) < TOL_STATE


def test_gate_with_indefinite_innovation_covariance_uses_lu(units, dev):
    """gating_test as the reference writes it (msckf.py:606-614) solves with LU, which also answers when H P H^T + sigma^2 I
    is not positive definite.  With P replaced by -1e4 P on both sides every gate of the mid-stream state is indefinite: the
    device's Cholesky gate must hand those landmarks to its LU kernel and arrive at the reference's gamma and decision.
    (The update that follows is a different matter: its indefinite S is an error by design, like a LinAlgError.)"""
    from msckf_b200 import capi
    u = units
    of = oracle_filter_from_units(u)
    Pneg = -1e4 * np.array(of.state_server._vio_state_cov__)      # H P H^T far below -sigma^2 in some directions
    of.state_server._vio_state_cov__ = Pneg.copy()
    dev.ctx.set_cov(0, Pneg)
    of.log = {}
    of.remove_lost_features()
    try:
        dev.remove_lost_features()          # gates first; the update then meets an indefinite S of its own and reports it
    except np.linalg.LinAlgError:
        pass
    log = dev.stage_log(0)
    st = log['status']
    gates = of.log.get('gate', [])
    sel = [i for i, s in enumerate(st) if s in (capi.FEAT_GATE_PASS, capi.FEAT_GATE_FAIL)]
    assert len(sel) == len(gates) and len(gates) >= 5
    gam = log['gamma'][sel]
    ref = np.array([g[0] for g in gates])
    assert np.all(np.isfinite(gam))
    assert np.all(np.abs(gam - ref) <= 1e-7 * np.abs(ref)), np.max(np.abs(gam - ref) / np.abs(ref))
    assert [st[i] == capi.FEAT_GATE_PASS for i in sel] == [g[2] for g in gates]
    assert (ref < 0).any()                                    # indefinite: some gammas are negative (and pass)
    assert int(dev.ctx.counters()[0, capi.CNT_GATE_NOT_SPD]) >= 1


def test_duplicate_feature_ids_are_rejected(units):
    """add_feature_observations requires unique ids inside one message (the front end guarantees it, image.py:425-438):
    a repeated id is reported as an argument error instead of silently creating two landmarks."""
    from msckf_b200 import capi
    dev = device_filter_from_units(units)
    try:
        ids = np.array([900001, 900002, 900001, 900003], dtype=np.int64)
        uv = np.zeros((4, 4))
        dev.ctx.add_observations([0, 4], ids, uv)
        with pytest.raises(capi.MsckfCudaError) as ei:
            dev.ctx.sync()
        assert ei.value.code == capi.ERR_ARG and 'duplicate' in str(ei.value)
    finally:
        dev.close()


def test_errors_are_loud(units):
    from msckf_b200 import capi
    from msckf_b200.config import ConfigEuRoC
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = 3            # below the supported window
    with pytest.raises(capi.MsckfCudaError):
        capi.Context(capi.make_cconfig(cfg))
    dev = device_filter_from_units(units)
    try:
        with pytest.raises(capi.MsckfCudaError):
            dev.ctx.get_cov(0, 7)                # wrong size
        with pytest.raises(capi.MsckfCudaError):
            dev.measurement_jacobian(10 ** 9, int(units['f_ids'][0]))   # unknown camera state
        # a covariance that makes S indefinite -> numeric error, like the reference's LinAlgError
        n = dev.ctx.n_state(0)
        dev.ctx.set_cov(0, -1e3 * np.identity(n))
        H = np.zeros((3, n))
        H[0, 30], H[1, 31], H[2, 32] = 1, 1, 1
        with pytest.raises(np.linalg.LinAlgError):
            dev.measurement_update(H, np.ones(3))
    finally:
        dev.close()


@pytest.mark.gpu
def test_undistort_points_vs_golden_and_oracle():
    """k_undistort through the C ABI (msckf_undistort_points / msckf_undistort_stereo) against the reference's
    recorded outputs and the oracle: FP64, 1e-12 relative (tan / division orders differ in the last bit)."""
    from oracle import image_oracle as io
    from msckf_b200 import image as gimg
    from msckf_b200.config import ConfigEuRoC
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'undistort.npz'))
    for name in ['radtan_cam0', 'radtan_cam1', 'radtan_rect', 'equi_a', 'equi_rect']:
        model = 'equidistant' if name.startswith('equi') else 'radtan'
        out = gimg.undistort_points(g[name + '_in'], g[name + '_intr'], model, g[name + '_dist'], g[name + '_R'], g[name + '_newK'])
        ref = g[name + '_out']
        assert np.max(np.abs(out - ref) / (1.0 + np.abs(ref))) < 1e-12, name
    assert gimg.undistort_points([], g['radtan_cam0_intr'], 'radtan', g['radtan_cam0_dist']) == []
    cfg = ConfigEuRoC()
    rng = np.random.default_rng(3)
    n = 100003                                           # ragged tail, more points than one grid pass
    p0 = np.stack([rng.uniform(0, 752, n), rng.uniform(0, 480, n)], axis=1)
    p1 = p0 + rng.normal(0, 3.0, p0.shape)
    uv = gimg.undistort_stereo(p0, p1, cfg)
    sel = rng.choice(n, 500, replace=False)
    ref = io.publish_measurements(p0[sel], p1[sel], cfg)
    assert np.max(np.abs(uv[sel] - ref)) < 1e-12
    msg = gimg.publish(1.5, np.arange(5), p0[:5], p1[:5], cfg)
    assert msg.timestamp == 1.5 and [f.id for f in msg.vio_features] == [0, 1, 2, 3, 4]
    assert np.allclose([[f.u0, f.v0, f.u1, f.v1] for f in msg.vio_features], uv[:5], rtol=0, atol=0)


def test_synth_frame_generator():
    """msckf_synth_frame: layout of msckf_feature_callback_dev, noise that is a pure function of
    (seed, sample / frame, element) -- filter f of seed0 equals filter 0 of seed0 + f, a frame regenerated alone is
    identical -- and the noise model of SURVEY 8d (sigmas, outlier fraction)."""
    import torch
    from msckf_b200 import capi
    F, K, n = 6, 10, 4000
    rng = np.random.default_rng(0)
    ts = 1.5 + 0.005 * np.arange(K)
    gyro_c, acc_c = rng.standard_normal((K, 3)) * 0.1, rng.standard_normal((K, 3))
    gb, ab = np.array([0.002, -0.001, 0.0015]), np.array([0.02, -0.01, 0.015])
    ids = np.arange(1000, 1000 + n, dtype=np.int64)
    uv_c = rng.uniform(-0.5, 0.5, (n, 4))
    sg, sa, sz, of, so = 0.07, 0.7, 0.001, 0.05, 0.05
    dev = torch.device('cuda', 0)

    def gen(nf, seed0, frame, sample0):
        t = {'imu_off': torch.empty(nf + 1, dtype=torch.int32, device=dev), 'ts': torch.empty(nf * K, dtype=torch.float64, device=dev),
             'gyro': torch.empty((nf, K, 3), dtype=torch.float64, device=dev), 'acc': torch.empty((nf, K, 3), dtype=torch.float64, device=dev),
             'cam_ids': torch.empty(nf, dtype=torch.int64, device=dev), 'feat_off': torch.empty(nf + 1, dtype=torch.int32, device=dev),
             'ids': torch.empty(nf * n, dtype=torch.int64, device=dev), 'uv': torch.empty((nf, n, 4), dtype=torch.float64, device=dev)}
        keys = ['imu_off', 'ts', 'gyro', 'acc', 'cam_ids', 'feat_off', 'ids', 'uv']
        capi.synth_frame(0, nf, seed0, frame, sample0, ts, gyro_c, acc_c, gb, ab, sg, sa, ids, uv_c, sz, of, so, 77, [t[k].data_ptr() for k in keys])
        return dict((k, v.cpu().numpy()) for k, v in t.items())

    a = gen(F, 10, 5, 300)
    b = gen(F, 10, 5, 300)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a['imu_off'], np.arange(F + 1) * K) and np.array_equal(a['feat_off'], np.arange(F + 1) * n)
    assert np.array_equal(a['cam_ids'], np.full(F, 77)) and np.array_equal(a['ids'], np.tile(ids, F))
    assert np.array_equal(a['ts'], np.tile(ts, F))
    one = gen(1, 13, 5, 300)                             # filter 3 of the batch, alone
    assert np.array_equal(one['uv'][0], a['uv'][3]) and np.array_equal(one['gyro'][0], a['gyro'][3])
    other = gen(F, 10, 6, 310)                           # another frame: new pixel noise, new IMU samples
    assert not np.array_equal(other['uv'], a['uv']) and not np.array_equal(other['gyro'], a['gyro'])
    shifted = gen(F, 10, 5, 295)                         # IMU noise is keyed by the global sample index
    assert np.allclose(shifted['gyro'][:, 5:] - gyro_c[5:], a['gyro'][:, :5] - gyro_c[:5], rtol=0, atol=1e-14)
    # noise model
    dg = (a['gyro'] - gyro_c - gb).ravel()
    da = (a['acc'] - acc_c - ab).ravel()
    assert abs(dg.std() / sg - 1) < 0.25 and abs(da.std() / sa - 1) < 0.25 and abs(dg.mean()) < 4 * sg / np.sqrt(dg.size)
    duv = a['uv'] - uv_c
    big = np.abs(duv).max(axis=2) > 6 * sz
    assert 0.03 < big.mean() < 0.07                      # gross outliers (sigma 0.05) on ~5 % of the observations
    inl = duv[~big].ravel()
    assert abs(inl.std() / sz - 1) < 0.03 and abs(inl.mean()) < 5 * sz / np.sqrt(inl.size)
    assert abs(duv[big].ravel().std() / so - 1) < 0.15
    for f1 in range(F):                                  # independent streams per seed
        for f2 in range(f1):
            assert abs(np.corrcoef(duv[f1].ravel(), duv[f2].ravel())[0, 1]) < 0.05
