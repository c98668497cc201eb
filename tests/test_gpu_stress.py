"""GPU parity at the STATED sizes of BASELINE.json configs[2] and configs[4] (the CUDA drop-in through the C ABI against
the CPU oracle on the same inputs):

  configs[2]  feature-update stress frame: one filter, 30 camera states (n = 201), 2000 lost landmarks with up to 29
              stereo observations each, row cap lifted: batched triangulation + Jacobian + null-space projection + gate,
              and the compression of a ~1.2e5 x 201 stacked Jacobian (grid-wide TSQR, k_qr_wide) followed by the wide
              tensor-pipe update (k_update_wide).  A lost landmark is by definition not seen by the newest camera state,
              so with 30 states in the window its track has at most 29 observations.
  configs[4]  large sliding window: max_cam_state_size = 64 (n = 405), 4096 features per frame, 40 frames after the
              window has filled: the covariance update on the full 405 x 405 P every frame and the pruning update
              over >= 4096 live landmarks every other frame.

Criteria as everywhere (north_star): state and covariance within 1e-9 relative, trajectory within 1 mm, decisions
(gate accept / reject, ids, removed camera states) bit-exact.
"""
import numpy as np
import pytest

from harness import block_rel_err, device_decisions, rel_err, run_pair, snapshot
from oracle import msckf_oracle as orc
from msckf_b200 import capi, synth
from msckf_b200.config import ConfigEuRoC
from msckf_b200.msckf import MSCKF

pytestmark = pytest.mark.gpu

TOL = 1e-9


def stress_landmarks(of, n_landmarks, rng, **kw):
    """synth.stress_tracks over the oracle filter's camera states (all but the newest one observe)."""
    cams = list(of.state_server._vio_cam_states__.values())
    return synth.stress_tracks(np.array([c.orientation for c in cams]), np.array([c._vio_position__ for c in cams]),
                               of.R_cam0_cam1, of.t_cam0_cam1, n_landmarks, rng, **kw)


def test_c3_stress_frame_2000_lost_landmarks():
    """configs[2] at its stated size.  Both filters are stepped through 29 ordinary frames, then frame 30 is taken stage
    by stage through the reference's public methods (SURVEY 8c hazard 3): batch_imu_processing, state_augmentation
    (30 camera states, n = 201), the map replaced on both sides by the same 2000 synthetic tracks, remove_lost_features
    with the 1500-row cap lifted, prune_cam_state_buffer."""
    max_cam, n_lead, n_lm = 30, 29, 2000
    stream = synth.make_stream(n_lead + 1, 40, structure_seed=8, noise_seed=3)
    ocfg = orc.ConfigEuRoC()
    ocfg._vio_max_cam_state_size__ = max_cam
    of = orc.MSCKF(ocfg)
    of.ROW_CAP = 10 ** 9
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    dcfg.row_cap = 0                                        # <= 0 lifts the cap in the library
    df = MSCKF(dcfg, feature_capacity=2304)
    msgs = stream.imu_msgs()
    k = 0
    try:
        for f in range(n_lead):
            while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
                of.imu_callback(msgs[k])
                df.imu_callback(msgs[k])
                k += 1
            of.feature_callback(stream.feature_msg(f))
            df.feature_callback(stream.feature_msg(f))
        t = stream.frame_ts[n_lead]
        while k < len(msgs) and msgs[k].vio_timestamp__ <= t:
            of.imu_callback(msgs[k])
            df.imu_callback(msgs[k])
            k += 1
        for flt in (of, df):
            flt.batch_imu_processing(t)
            flt.state_augmentation(t)
        so, sd = snapshot(of), snapshot(df)
        assert len(so['cam_ids']) == 30 and so['P'].shape == (201, 201)
        assert np.array_equal(sd['cam_ids'], so['cam_ids']) and rel_err(sd['P'], so['P']) < TOL

        ids, masks, obs = stress_landmarks(of, n_lm, np.random.default_rng(42), outlier_frac=0.005)
        cam_ids = [int(c) for c in so['cam_ids']]
        of.map_server.clear()
        for j in range(n_lm):
            ft = orc.Feature(int(ids[j]), of.optimization_config, of.R_cam0_cam1, of.t_cam0_cam1)
            for c in range(30):
                if (int(masks[j]) >> c) & 1:
                    ft.observations[cam_ids[c]] = obs[j, c].copy()
            of.map_server[int(ids[j])] = ft
        df.ctx.set_features(0, ids, masks, None, None, obs)

        of.log = {}
        of.remove_lost_features()
        df.remove_lost_features()
        so, sd = snapshot(of), snapshot(df)
        log = df.stage_log(0)
        dec, gam, _, upd = device_decisions(df)
        ogates = of.log['gate']
        assert len(ogates) > 1900                          # nearly every track triangulates
        assert dec == [x[2] for x in ogates]               # bit-exact gate decisions, in map order
        og = np.array([x[0] for x in ogates])
        assert np.all(np.abs(np.array(gam) - og) <= 1e-7 * np.abs(og))
        assert 0 < sum(dec) < len(dec)
        # lost ids: processed vs invalid (triangulation failed) are the same sets in the same order
        dev_invalid = [int(i) for i, s in zip(log['ids'], log['status']) if s == capi.FEAT_TRI_FAILED]
        assert dev_invalid == [int(x) for x in of.log['lost_invalid'][0]]
        assert len(upd) == 1 and upd[0] > 100000, upd      # one update with > 1e5 stacked rows
        m, na, kr, path = df.ctx.update_shapes()[0]
        assert m == upd[0] and na == 174 and kr == 174 and path == 2, (m, na, kr, path)   # TSQR + wide update
        assert len(sd['map_ids']) == 0 and len(so['map_ids']) == 0
        errs = {key: rel_err(sd[key], so[key]) for key in ('imu', 'P', 'cams', 'extr')}
        errs['P_block'], errs['P_block_strict'] = block_rel_err(sd['P'], so['P'])
        errs['pos_m'] = float(np.linalg.norm(sd['imu'][13:16] - so['imu'][13:16]))
        print('c3 parity', errs, 'rows', upd[0], 'gates', len(dec), 'pass', sum(dec))
        assert max(errs['imu'], errs['P'], errs['cams'], errs['extr'], errs['P_block']) < TOL, errs
        assert errs['pos_m'] < 1e-3
        assert int(df.ctx.counters()[0, capi.CNT_GATE_NOT_SPD]) == 0

        of.prune_cam_state_buffer()
        df.prune_cam_state_buffer()
        so, sd = snapshot(of), snapshot(df)
        assert np.array_equal(sd['cam_ids'], so['cam_ids']) and len(so['cam_ids']) == 28
        assert rel_err(sd['P'], so['P']) < TOL and rel_err(sd['cams'], so['cams']) < TOL
    finally:
        df.close()


def test_c5_window64_4096_features_per_frame():
    """configs[4] at its stated size, update dominated: 64 cheap lead-in frames (300 features) fill the 64-state window,
    then 40 frames with 4096 features per frame, every one compared with the oracle.  The platform moves fast
    (speed = 10: more than the key-frame distance of msckf.py:717 per frame), so pruning removes the oldest camera states
    and every frame's lost landmarks carry long tracks: a 1500-row (capped, msckf.py:678) x 405 stacked Jacobian over
    up to 384 active columns is compressed and applied to the full 405 x 405 covariance on every frame."""
    n_lead, n_full = 64, 40
    n_frames = n_lead + n_full
    target = np.array([300] * n_lead + [4096] * n_full)
    st = synth.make_structure(n_frames, target, structure_seed=12, n_landmarks=150000, speed=10.0)
    assert min(len(x) for x in st.frame_ids[n_lead + 1:]) == 4096
    stream = synth.realize(st, noise_seed=5)
    wide = []

    def after(f, of, df):
        log = df.stage_log(0)
        if log['update_rows'] > 1500 and log['update_cols'] >= 399:
            wide.append(f)

    worst, counts = run_pair(stream, 64, n_frames, feature_capacity=8192, after_frame=after)
    print('c5 parity', worst, counts, 'wide updates', len(wide))
    assert counts['prunes'] >= 19 and counts['gates'] > 4000 and 0 < counts['pass'] < counts['gates'], counts
    assert len(wide) >= 30, wide                           # capped 1500-row updates on the (almost) full window
    assert max(worst['imu'], worst['P'], worst['cams'], worst['extr'], worst['P_block']) < TOL, worst
    assert worst['pos_m'] < 1e-3, worst


def test_c5_window64_pruning_dominated():
    """configs[4] on the ordinary (slow) trajectory: the camera states next to the key frame are pruned, so almost every
    lost track is short, and the work is the pruning update over all 4096 live landmarks every other frame: 4096
    two-view Jacobians and a 20480-row x 12-column stack, which a one-filter batch spreads over the grid (k_qr_wide)."""
    n_lead, n_full = 64, 16
    n_frames = n_lead + n_full
    target = np.array([300] * n_lead + [4096] * n_full)
    st = synth.make_structure(n_frames, target, structure_seed=12, n_landmarks=150000)
    stream = synth.realize(st, noise_seed=5)
    worst, counts = run_pair(stream, 64, n_frames, feature_capacity=8192)
    print('c5 pruning parity', worst, counts)
    assert counts['prunes'] >= 8 and counts['gates'] > 15000 and 0 < counts['pass'] < counts['gates'], counts
    assert counts['max_update_rows'] >= 5 * 2500, counts    # 5 rows per live landmark that passes the gate
    assert max(worst['imu'], worst['P'], worst['cams'], worst['extr'], worst['P_block']) < TOL, worst
    assert worst['pos_m'] < 1e-3, worst
