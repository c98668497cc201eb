"""CPU: pins oracle/msckf_oracle.py to the golden vectors that tests/golden/make_golden.py
recorded from the unmodified reference (the reference has no tests of its own)."""
import copy
import os

import numpy as np
import pytest

from harness import snapshot, rel_err, oracle_filter_from_units as build_filter_from_units
from oracle import msckf_oracle as orc
from msckf_b200 import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TOL = 1e-11       # oracle vs reference: same algorithm, different BLAS call order only


def probe_vector(n):
    return np.cos(1.0 + 0.37 * np.arange(n) ** 1.5)


@pytest.fixture(scope='module')
def units():
    return dict(np.load(os.path.join(GOLD, 'units.npz')))


@pytest.fixture(scope='module')
def stream_gold():
    return dict(np.load(os.path.join(GOLD, 'stream_small.npz')))


def test_utils_golden(units):
    u = units
    for q, R in zip(u['u_q'], u['u_to_rotation']):
        assert rel_err(orc.to_rotation(q), R) < 1e-14
    for R, q in zip(u['u_R'], u['u_to_quaternion']):
        assert rel_err(orc.to_quaternion(R), q) < 1e-14
    for i in range(8):
        assert rel_err(orc.quaternion_multiplication(u['u_q'][i], u['u_q'][(i + 1) % 8]), u['u_qmul'][i]) < 1e-14
    for d, q in zip(u['u_dtheta'], u['u_small_angle']):
        assert rel_err(orc.small_angle_quaternion(d), q) < 1e-14
    for a, b, q in zip(u['u_v0'], u['u_v1'], u['u_from_two']):
        assert rel_err(orc.from_two_vectors(a, b), q) < 1e-14


def test_jit_kernels_golden(units):
    u = units
    q, v, p, g = u['j_q'], u['j_v'], u['j_p'], u['j_gravity']
    Rwi = orc.to_rotation(q)
    for i in range(3):
        G, Phi = orc.process_model_matrices(u['j_gyro'][i], Rwi, u['j_acc'][i], u['j_dt'][i])
        assert rel_err(G, u['j_G'][i]) < 1e-14
        assert rel_err(Phi, u['j_Phi'][i]) < 1e-14
        qn, vn, pn = orc.predict_new_state(u['j_dt'][i], u['j_gyro'][i], u['j_acc'][i], q, v, p, g)
        assert rel_err(np.concatenate([qn, vn, pn]), u['j_pred'][i]) < 1e-14
    P = u['state_P']
    ext = u['state_extr']
    aug = orc.augment_covariance(ext[:9].reshape(3, 3), Rwi, ext[9:], P.copy())
    assert rel_err(aug, u['j_aug']) < 1e-13
    cfg = orc.ConfigEuRoC()
    filt = orc.MSCKF(cfg)
    G, Phi = orc.process_model_matrices(u['j_gyro'][0], Rwi, u['j_acc'][0], u['j_dt'][0])
    Pp = orc.propagate_imu_covariance(filt.state_server._vio_continuous_noise_cov__, P.copy(), G, Phi, u['j_dt'][0])
    assert rel_err(Pp[:21, :21], u['j_prop']) < 1e-13


def test_triangulation_golden(units):
    u = units
    filt = build_filter_from_units(u)
    cams = filt.state_server._vio_cam_states__
    for k, fid in enumerate(u['f_ids']):
        ft = filt.map_server[int(fid)]
        ok = ft.initialize_position(cams)
        assert ok == bool(u['f_tri_ok'][k])
        assert rel_err(ft._vio_position__, u['f_tri_pos'][k]) < 1e-10
    # degenerate (behind-camera) cases: flags bit-exact
    j = 0
    for k in range(len(u['f_bad_ok'])):
        src = filt.map_server[int(u['f_ids'][k])]
        probe = orc.Feature(0, filt.optimization_config, filt.R_cam0_cam1, filt.t_cam0_cam1)
        for c in src.observations:
            probe.observations[c] = u['f_bad_z'][j]
            j += 1
        assert probe.initialize_position(cams) == bool(u['f_bad_ok'][k])
        assert rel_err(probe._vio_position__, u['f_bad_pos'][k]) < 1e-9


def test_jacobian_gate_update_golden(units):
    u = units
    filt = build_filter_from_units(u)
    for k, fid in enumerate(u['f_ids']):
        ft = filt.map_server[int(fid)]
        ft._vio_position__ = u['f_tri_pos'][k].copy()
    P = filt.state_server._vio_state_cov__
    for k in range(len(u['m_jac'])):
        ft = filt.map_server[int(u['f_ids'][k])]
        cid = list(ft.observations.keys())[1]
        Hx, Hf, r = filt.measurement_jacobian(cid, ft.id)
        assert rel_err(np.concatenate([Hx.ravel(), Hf.ravel(), r]), u['m_jac'][k]) < 1e-13
    Hs, rs = [], []
    for k in range(len(u['fj_rows'])):
        ft = filt.map_server[int(u['f_ids'][k])]
        H, r = filt.feature_jacobian(ft.id, list(ft.observations.keys()))
        assert len(r) == int(u['fj_rows'][k])
        assert rel_err(H.T @ H, u['fj_HtH'][k]) < 1e-12
        assert rel_err(H.T @ r, u['fj_Htr'][k]) < 1e-11
        S = H @ P @ H.T + filt.config._vio_observation_noise__ * np.identity(len(H))
        assert abs(r @ np.linalg.solve(S, r) - u['fj_gamma'][k]) <= 1e-10 * abs(u['fj_gamma'][k])
        Hs.append(H)
        rs.append(r)
    for tag in ('wide', 'tall'):
        f2 = copy.deepcopy(filt)
        f2.measurement_update(u['mu_%s_H' % tag], u['mu_%s_r' % tag])
        s2 = snapshot(f2)
        assert rel_err(s2['imu'], u['mu_%s_imu' % tag]) < TOL
        assert rel_err(s2['extr'], u['mu_%s_extr' % tag]) < TOL
        assert rel_err(s2['cams'], u['mu_%s_cams' % tag]) < TOL
        assert rel_err(s2['P'], u['mu_%s_P' % tag]) < 1e-10
        i2 = f2.state_server._vio_imu_state__
        nul = np.concatenate([i2.orientation_null, i2.velocity_null, i2.position_null])
        assert rel_err(nul, u['mu_%s_null' % tag]) < TOL


def test_stream_golden(stream_gold):
    g = stream_gold
    nf, nfeat, sseed, nseed = [int(x) for x in g['stream_kw']]
    stream = synth.make_stream(nf, nfeat, sseed, nseed)
    cfg = orc.ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = int(g['max_cam'])
    filt = orc.MSCKF(cfg)
    filt.log = {}
    worst = {'imu': 0.0, 'Pv': 0.0, 'cams': 0.0}

    def seg(key, f):
        o = g[key + '_off']
        return g[key][o[f]:o[f + 1]]

    def check(f, flt, res):
        s = snapshot(flt)
        assert np.array_equal(s['cam_ids'], seg('cam_ids', f)), f
        assert np.array_equal(s['map_ids'], seg('map_ids', f)), f
        n = s['P'].shape[0]
        assert n == int(g['n'][f])
        worst['imu'] = max(worst['imu'], rel_err(s['imu'], g['imu'][f]))
        worst['cams'] = max(worst['cams'], rel_err(s['cams'].ravel(), seg('cams', f)))
        worst['Pv'] = max(worst['Pv'], rel_err(s['P'] @ probe_vector(n), seg('Pv', f)))
        if 'P_%d' % f in g:
            assert rel_err(s['P'], g['P_%d' % f]) < 1e-10

    synth.drive(filt, stream, on_frame=check)
    assert worst['imu'] < TOL and worst['cams'] < TOL and worst['Pv'] < 1e-10, worst
    # decisions: gating accept/reject sequence and removed cam ids are bit-exact
    gates = filt.log['gate']
    assert len(gates) == len(g['gate'])
    assert [bool(x[2]) for x in gates] == [bool(x) for x in g['gate'][:, 3]]
    gam = np.array([x[0] for x in gates])
    assert np.all(np.abs(gam - g['gate'][:, 2]) <= 1e-8 * np.abs(g['gate'][:, 2]))
    rm = np.array(filt.log['rm_cam_ids'], dtype=np.int64)
    assert np.array_equal(rm, g['rm'][:, 1:])


def test_image_oracle_undistort_vs_reference_golden():
    """oracle/image_oracle.py against the reference's ImageProcessor.undistort_points (cv2 4.13 underneath), recorded by
    tests/golden/make_golden_image.py: radtan (both EuRoC cameras, with and without rectification / new camera matrix)
    and equidistant (including the reference's positional-argument quirk).  FP64 in, FP64 out: 1e-12 relative."""
    from oracle import image_oracle as io
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'undistort.npz'))
    for name in ['radtan_cam0', 'radtan_cam1', 'radtan_rect', 'equi_a', 'equi_rect']:
        model = 'equidistant' if name.startswith('equi') else 'radtan'
        out = io.undistort_points(g[name + '_in'], g[name + '_intr'], model, g[name + '_dist'], g[name + '_R'], g[name + '_newK'])
        ref = g[name + '_out']
        assert np.max(np.abs(out - ref) / (1.0 + np.abs(ref))) < 1e-12, name
    assert io.undistort_points([], g['radtan_cam0_intr'], 'radtan', g['radtan_cam0_dist']) == []
