"""Generates the golden vectors that pin ``oracle/msckf_oracle.py`` to the reference.

Runs ONLY in the build container: it imports the unmodified reference from
/root/reference/MSCKF (which does not travel to the GPU box) and writes small ``.npz``
fixtures next to this file.  The reference ships no tests or fixtures of its own (SURVEY.md
section 4), so "the reference itself, run here on the synthetic stream" is the pin.

    python tests/golden/make_golden.py

Fixtures:
  stream_small.npz  per-frame state / ids / decisions of the reference stepping a
                    45-frame, 120-features/frame stream with max_cam_state_size = 12
  units.npz         input -> output vectors of the reference's own functions
                    (utils.py, jit_utils.py, Feature.*, measurement_jacobian, feature_jacobian,
                    gating_test, measurement_update) taken at a mid-stream state
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from harness import REPO, import_reference, reset_reference_globals, snapshot   # noqa: E402

sys.path.insert(0, REPO)
from msckf_b200 import synth   # noqa: E402

STREAM_KW = dict(n_frames=45, feats_per_frame=120, structure_seed=0, noise_seed=0)
MAX_CAM = 12


def probe_vector(n):
    """Fixed pseudo-random probe; P @ v is stored per frame as a fingerprint of P."""
    return np.cos(1.0 + 0.37 * np.arange(n) ** 1.5)


def pack_ragged(list_of_arrays, dtype):
    off = np.zeros(len(list_of_arrays) + 1, dtype=np.int64)
    for i, a in enumerate(list_of_arrays):
        off[i + 1] = off[i] + len(a)
    flat = np.concatenate([np.asarray(a, dtype=dtype).ravel() for a in list_of_arrays]) if list_of_arrays else np.zeros(0, dtype)
    return flat, off


def main():
    ref, RefConfig = import_reference()
    import utils as ref_utils
    import jit_utils as ref_jit
    from feature import Feature as RefFeature
    reset_reference_globals(ref)

    log = {'gate': [], 'tri': [], 'rm': [], 'upd_rows': [], 'frame_of_gate': []}
    cur = {'f': -1}

    class Wrapped(ref.MSCKF):
        def gating_test(self, H, r, dof):
            ok = super().gating_test(H, r, dof)
            P = self.state_server._vio_state_cov__
            gamma = r @ np.linalg.solve(H @ P @ H.T + self.config._vio_observation_noise__ * np.identity(len(H)), r)
            log['gate'].append((cur['f'], dof, gamma, float(ok)))
            return ok

        def find_redundant_cam_states(self):
            rm = super().find_redundant_cam_states()
            log['rm'].append((cur['f'], rm[0], rm[1]))
            return rm

        def measurement_update(self, H, r):
            log['upd_rows'].append((cur['f'], H.shape[0], H.shape[1]))
            return super().measurement_update(H, r)

    orig_init = RefFeature.initialize_position

    def logged_init(self, cam_states):
        ok = orig_init(self, cam_states)
        log['tri'].append((cur['f'], self.id, float(ok), *self._vio_position__))
        return ok
    RefFeature.initialize_position = logged_init

    cfg = RefConfig()
    cfg._vio_max_cam_state_size__ = MAX_CAM
    filt = Wrapped(cfg)
    stream = synth.make_stream(**STREAM_KW)

    frames = {'imu': [], 'extr': [], 'cam_ids': [], 'cams': [], 'map_ids': [], 'Pv': [], 'Pdiag': [], 'Pfro': [], 'n': []}
    fullP = {}
    mid = {}

    msgs = stream.imu_msgs()
    k = 0
    for f in range(stream.n_frames):
        cur['f'] = f
        t = stream.frame_ts[f]
        while k < len(msgs) and msgs[k].vio_timestamp__ <= t:
            filt.imu_callback(msgs[k])
            k += 1
        if f == 30:
            # ---- mid-stream unit vectors, taken after propagation+augmentation+observations
            filt.is_first_img = False
            fm = stream.feature_msg(f)
            filt.batch_imu_processing(fm.timestamp)
            filt.state_augmentation(fm.timestamp)
            filt.add_feature_observations(fm)
            mid = collect_units(filt, ref, ref_utils, ref_jit, RefFeature, orig_init)
            filt.remove_lost_features()
            filt.prune_cam_state_buffer()
            filt.publish(fm.timestamp)
            filt.online_reset()
        else:
            filt.feature_callback(stream.feature_msg(f))
        s = snapshot(filt)
        n = s['P'].shape[0]
        frames['imu'].append(s['imu'])
        frames['extr'].append(s['extr'])
        frames['cam_ids'].append(s['cam_ids'])
        frames['cams'].append(s['cams'].ravel())
        frames['map_ids'].append(s['map_ids'])
        frames['Pv'].append(s['P'] @ probe_vector(n))
        frames['Pdiag'].append(np.diag(s['P']).copy())
        frames['Pfro'].append(np.linalg.norm(s['P']))
        frames['n'].append(n)
        if f in (5, 20, 44):
            fullP['P_%d' % f] = s['P']

    out = {'imu': np.array(frames['imu']), 'extr': np.array(frames['extr']), 'Pfro': np.array(frames['Pfro']),
           'n': np.array(frames['n'])}
    for key, dt in (('cam_ids', np.int64), ('cams', np.float64), ('map_ids', np.int64), ('Pv', np.float64), ('Pdiag', np.float64)):
        out[key], out[key + '_off'] = pack_ragged(frames[key], dt)
    out['gate'] = np.array(log['gate'])
    out['tri'] = np.array(log['tri'])
    out['rm'] = np.array(log['rm'], dtype=np.int64)
    out['upd_rows'] = np.array(log['upd_rows'], dtype=np.int64)
    out['max_cam'] = np.array(MAX_CAM)
    out['stream_kw'] = np.array([STREAM_KW['n_frames'], STREAM_KW['feats_per_frame'], STREAM_KW['structure_seed'], STREAM_KW['noise_seed']])
    out.update(fullP)
    np.savez_compressed(os.path.join(HERE, 'stream_small.npz'), **out)
    np.savez_compressed(os.path.join(HERE, 'units.npz'), **mid)
    print('gates', len(log['gate']), 'pass', int(sum(g[3] for g in log['gate'])), 'tri', len(log['tri']),
          'prunes', len(log['rm']), 'updates', len(log['upd_rows']))
    for name in ('stream_small.npz', 'units.npz'):
        print(name, os.path.getsize(os.path.join(HERE, name)) // 1024, 'KiB')


def collect_units(filt, ref, U, J, RefFeature, raw_init):
    """Input->output vectors of individual reference functions at the current filter state."""
    rng = np.random.default_rng(7)
    out = {}
    ss = filt.state_server
    # ---------------- utils.py
    qs = rng.standard_normal((8, 4))
    out['u_q'] = qs
    out['u_to_rotation'] = np.array([U.to_rotation(q) for q in qs])
    Rs = list(out['u_to_rotation'])
    # force all four to_quaternion branches
    for q in ([0.9, 0.1, 0.2, 0.1], [0.1, 0.9, 0.2, 0.1], [0.1, 0.2, 0.9, 0.1], [0.1, 0.2, 0.1, 0.9]):
        Rs.append(U.to_rotation(np.array(q)))
    out['u_R'] = np.array(Rs)
    out['u_to_quaternion'] = np.array([U.to_quaternion(R) for R in Rs])
    out['u_qmul'] = np.array([U.quaternion_multiplication(qs[i], qs[(i + 1) % 8]) for i in range(8)])
    dth = np.concatenate([0.05 * rng.standard_normal((4, 3)), 3.0 * rng.standard_normal((4, 3))])
    out['u_dtheta'] = dth
    out['u_small_angle'] = np.array([U.small_angle_quaternion(d) for d in dth])
    v0 = np.concatenate([rng.standard_normal((5, 3)), [[0, 0, 1.0]], [[0, 0, 1.0]], [[1.0, 0, 0]]])
    v1 = np.concatenate([rng.standard_normal((5, 3)), [[0, 0, 1.0]], [[0, 0, -1.0]], [[-1.0, 0, 0]]])
    out['u_v0'], out['u_v1'] = v0, v1
    out['u_from_two'] = np.array([U.from_two_vectors(a, b) for a, b in zip(v0, v1)])

    # ---------------- jit_utils.py
    imu = ss._vio_imu_state__
    gy = np.array([[0.1, -0.2, 0.3], [1e-7, 2e-7, -1e-7], [0.5, 0.4, -0.9]])
    ac = np.array([[0.3, -0.1, 9.7], [0.0, 0.0, 9.81], [2.0, -1.0, 8.0]])
    dts = np.array([0.005, 0.005, 0.01])
    out['j_gyro'], out['j_acc'], out['j_dt'] = gy, ac, dts
    out['j_q'], out['j_v'], out['j_p'] = imu.orientation.copy(), imu.velocity.copy(), imu._vio_position__.copy()
    out['j_gravity'] = ref.IMUState._vio_gravity__.copy()
    Rwi = U.to_rotation(imu.orientation)
    pm = [J._process_model(gy[i], Rwi, ac[i], dts[i]) for i in range(3)]
    out['j_G'] = np.array([p[1] for p in pm])
    out['j_Phi'] = np.array([p[5] for p in pm])
    pn = [J._predict_new_state(dts[i], gy[i], ac[i], imu.orientation, imu.velocity, imu._vio_position__,
                               ref.IMUState._vio_gravity__) for i in range(3)]
    out['j_pred'] = np.array([np.concatenate(p) for p in pn])
    P = ss._vio_state_cov__.copy()
    out['state_P'] = P
    out['j_aug'] = J._state_augmentation(imu.R_imu_cam0, Rwi, imu.t_cam0_imu, P.copy(), P.shape[0])
    Pc = P.copy()
    out['j_prop'] = J._propaget_state_Covariance(ss._vio_continuous_noise_cov__, Pc, pm[0][1], pm[0][5], dts[0])[:21, :21].copy()

    # ---------------- filter state at this point
    out['state_imu'] = np.concatenate([imu.orientation, imu._vio_gyro_bias__, imu.velocity, imu.acc_bias, imu._vio_position__])
    out['state_imu_null'] = np.concatenate([imu.orientation_null, imu.velocity_null, imu.position_null])
    out['state_extr'] = np.concatenate([imu.R_imu_cam0.ravel(), imu.t_cam0_imu])
    cams = ss._vio_cam_states__
    out['state_cam_ids'] = np.array(list(cams.keys()), dtype=np.int64)
    out['state_cams'] = np.array([np.concatenate([c.orientation, c._vio_position__, c.orientation_null, c.position_null])
                                  for c in cams.values()])
    out['state_imu_id'] = np.array(imu.id)
    out['state_tracking_rate'] = np.array(filt.tracking_rate)

    # ---------------- features: observations + triangulation results
    feats = [ft for ft in filt.map_server.values() if len(ft.observations) >= 3][:60]
    obs_cam, obs_z, off = [], [], [0]
    tri_pos, tri_ok, was_init, prev_pos = [], [], [], []
    for ft in feats:
        for cid, z in ft.observations.items():
            obs_cam.append(cid)
            obs_z.append(z)
        off.append(len(obs_cam))
        was_init.append(ft.is_initialized)
        prev_pos.append(ft._vio_position__.copy())
        probe = RefFeature(ft.id, ft.optimization_config)
        probe.observations = dict(ft.observations)
        ok = raw_init(probe, cams)
        tri_ok.append(ok)
        tri_pos.append(probe._vio_position__.copy())
    out['f_ids'] = np.array([ft.id for ft in feats], dtype=np.int64)
    out['f_obs_cam'] = np.array(obs_cam, dtype=np.int64)
    out['f_obs_z'] = np.array(obs_z)
    out['f_off'] = np.array(off, dtype=np.int64)
    out['f_tri_pos'] = np.array(tri_pos)
    out['f_tri_ok'] = np.array(tri_ok)
    out['f_was_init'] = np.array(was_init)
    out['f_prev_pos'] = np.array(prev_pos)

    # degenerate triangulations: mirrored observations put the point behind the cameras
    bad_pos, bad_ok, bad_z = [], [], []
    for ft in feats[:6]:
        probe = RefFeature(ft.id, ft.optimization_config)
        probe.observations = {c: np.array([-z[0] + 0.3, z[1], -z[2] - 0.3, z[3]]) for c, z in ft.observations.items()}
        bad_z += list(probe.observations.values())
        bad_ok.append(raw_init(probe, cams))
        bad_pos.append(probe._vio_position__.copy())
    out['f_bad_z'] = np.array(bad_z)
    out['f_bad_ok'] = np.array(bad_ok)
    out['f_bad_pos'] = np.array(bad_pos)

    # ---------------- Jacobians, gate, update (positions = fresh triangulation above)
    saved = [(ft._vio_position__, ft.is_initialized) for ft in feats]
    for ft, p in zip(feats, tri_pos):
        ft._vio_position__ = p.copy()
    mj = []
    for ft in feats[:10]:
        cid = list(ft.observations.keys())[1]
        Hx, Hf, r = filt.measurement_jacobian(cid, ft.id)
        mj.append(np.concatenate([Hx.ravel(), Hf.ravel(), r]))
    out['m_jac'] = np.array(mj)
    HtH, Htr, gam, rows = [], [], [], []
    Hs, rs = [], []
    for ft in feats[:24]:
        cids = list(ft.observations.keys())
        H, r = filt.feature_jacobian(ft.id, cids)
        HtH.append(H.T @ H)
        Htr.append(H.T @ r)
        S = H @ P @ H.T + filt.config._vio_observation_noise__ * np.identity(len(H))
        gam.append(r @ np.linalg.solve(S, r))
        rows.append(len(r))
        Hs.append(H)
        rs.append(r)
    out['fj_HtH'] = np.array(HtH)
    out['fj_Htr'] = np.array(Htr)
    out['fj_gamma'] = np.array(gam)
    out['fj_rows'] = np.array(rows)
    for ft, (p, ini) in zip(feats, saved):
        ft._vio_position__, ft.is_initialized = p, ini

    # measurement_update on a deep copy of the filter: tall (QR) and wide cases
    import copy
    n_state = P.shape[0]
    wide_take = 1
    while sum(rows[:wide_take + 1]) <= n_state:
        wide_take += 1
    for tag, take in (('wide', wide_take), ('tall', 24)):
        H = np.vstack(Hs[:take])
        r = np.concatenate(rs[:take])
        if tag == 'tall':
            assert H.shape[0] > H.shape[1], H.shape
        else:
            assert H.shape[0] <= H.shape[1], H.shape
        f2 = copy.deepcopy(filt)
        f2.measurement_update(H, r)
        s2 = snapshot(f2)
        i2 = f2.state_server._vio_imu_state__
        out['mu_%s_H' % tag], out['mu_%s_r' % tag] = H, r
        out['mu_%s_imu' % tag], out['mu_%s_extr' % tag] = s2['imu'], s2['extr']
        out['mu_%s_cams' % tag], out['mu_%s_P' % tag] = s2['cams'], s2['P']
        out['mu_%s_null' % tag] = np.concatenate([i2.orientation_null, i2.velocity_null, i2.position_null])
    return out


if __name__ == '__main__':
    main()
