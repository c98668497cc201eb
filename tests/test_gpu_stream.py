"""GPU parity, stream level: the CUDA drop-in stepped frame by frame through feature_callback
(the reference's public API) against

  * the golden per-frame record of the unmodified reference (tests/golden/stream_small.npz), and
  * the CPU oracle stepped beside it on a longer / wider stream (BASELINE.json configs[1] shape:
    max_cam_state_size = 30), and as a batch of independent filters (configs[3] shape).

Criteria (north_star): state and covariance within 1e-9 relative at every frame, trajectory within
1 mm, and bit-exact decisions: gate accept / reject sequence, lost / processed / pruned landmark
ids (map contents), removed camera-state ids.
"""
import os

import numpy as np
import pytest

from harness import device_decisions, rel_err, run_pair, snapshot
from oracle import msckf_oracle as orc
from msckf_b200 import capi, synth
from msckf_b200.config import ConfigEuRoC
from msckf_b200.msckf import MSCKF, MSCKFBatch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TOL = 1e-9


def probe_vector(n):
    return np.cos(1.0 + 0.37 * np.arange(n) ** 1.5)


def test_stream_vs_reference_golden():
    g = dict(np.load(os.path.join(GOLD, 'stream_small.npz')))
    nf, nfeat, sseed, nseed = [int(x) for x in g['stream_kw']]
    stream = synth.make_stream(nf, nfeat, sseed, nseed)
    cfg = ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = int(g['max_cam'])
    filt = MSCKF(cfg, feature_capacity=512)
    worst = {'imu': 0.0, 'Pv': 0.0, 'cams': 0.0, 'extr': 0.0, 'Pdiag': 0.0}
    dec_all, gam_all, rm_all, upd_all = [], [], [], []

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
        worst['extr'] = max(worst['extr'], rel_err(s['extr'], g['extr'][f]))
        worst['cams'] = max(worst['cams'], rel_err(s['cams'].ravel(), seg('cams', f)))
        worst['Pv'] = max(worst['Pv'], rel_err(s['P'] @ probe_vector(n), seg('Pv', f)))
        worst['Pdiag'] = max(worst['Pdiag'], rel_err(np.diag(s['P']), seg('Pdiag', f)))
        assert abs(np.linalg.norm(s['P']) - g['Pfro'][f]) <= TOL * g['Pfro'][f]
        if 'P_%d' % f in g:
            assert rel_err(s['P'], g['P_%d' % f]) < TOL
        dec, gam, rm, upd = device_decisions(flt)
        dec_all.extend(dec)
        gam_all.extend(gam)
        upd_all.extend(upd)
        if rm is not None:
            rm_all.append([f] + rm)

    try:
        synth.drive(filt, stream, on_frame=check)
    finally:
        filt.close()
    assert max(worst.values()) < TOL, worst
    assert dec_all == [bool(x) for x in g['gate'][:, 3]]                 # bit-exact gate decisions
    gam = np.array(gam_all)
    rel = np.abs(gam - g['gate'][:, 2]) / np.abs(g['gate'][:, 2])
    # gamma of a grossly rejected outlier (gamma ~ 1e6) amplifies the <= 1e-9 state differences; decisions stay bit-exact
    assert rel.max() <= 1e-7, (rel.max(), int(rel.argmax()), gam[rel.argmax()], g['gate'][rel.argmax()])
    assert np.array_equal(np.array(rm_all, dtype=np.int64), g['rm'])      # removed camera states
    # rows of every update; the generator's two probe calls of measurement_update on deep copies at
    # frame 30 (make_golden.py, collect_units) are not part of the stream
    gold_upd = [(int(fr), int(r)) for fr, r, _ in g['upd_rows']]
    probe = [i for i, (fr, _) in enumerate(gold_upd) if fr == 30][:2]
    gold_upd = [r for i, (fr, r) in enumerate(gold_upd) if i not in probe and r > 0]
    assert upd_all == gold_upd


def test_stream_vs_oracle_window30():
    """configs[1] in full: one filter, max_cam_state_size = 30, ~300 features per frame, all 200 frames of the sequence;
    state and P (whole and per block) within 1e-9 at every frame, trajectory within 1 mm over the sequence."""
    n_frames = 200
    stream = synth.make_stream(n_frames, 300, structure_seed=1, noise_seed=11)
    worst, counts = run_pair(stream, 30, n_frames, feature_capacity=2048)
    print('c2 parity', worst, counts)
    assert counts['gates'] > 15000 and counts['prunes'] >= 80 and 0 < counts['pass'] < counts['gates'], counts
    assert worst['imu'] < TOL and worst['P'] < TOL and worst['cams'] < TOL and worst['extr'] < TOL, worst
    assert worst['P_block'] < TOL, worst
    assert worst['pos_m'] < 1e-3, worst


def test_online_reset_forced():
    """online_reset (msckf.py:859-886): the position variance of both filters is inflated past the 8 m threshold before
    two frames of the stream; at the end of those frames both must clear the camera window and the map and reset P --
    the same decision, the same state afterwards, and the same evolution over the frames that follow (the camera-state
    ids continue, the map refills)."""
    n_frames = 36
    stream = synth.make_stream(n_frames, 100, structure_seed=4, noise_seed=2)
    forced = (12, 25)

    def inflate(f, of, df):
        if f not in forced:
            return
        Po = of.state_server._vio_state_cov__
        Po[12 + (f % 3), 12 + (f % 3)] = 70.0               # sigma = 8.4 m >= position_std_threshold (config.py:64)
        ss = df.state_server
        Pd = ss._vio_state_cov__.copy()
        Pd[12 + (f % 3), 12 + (f % 3)] = 70.0
        ss._vio_state_cov__ = Pd                            # uploads through msckf_set_cov

    after_reset = {}

    def after(f, of, df):
        if f in forced:
            so, sd = snapshot(of), snapshot(df)
            after_reset[f] = (len(sd['cam_ids']), len(sd['map_ids']), sd['P'].shape, bool(df.did_reset),
                              np.array_equal(sd['P'], so['P']))

    worst, counts = run_pair(stream, 8, n_frames, feature_capacity=512, before_frame=inflate, after_frame=after)
    assert counts['resets'] == 2, counts
    for f in forced:
        assert after_reset[f] == (0, 0, (21, 21), True, True), after_reset
    assert counts['prunes'] >= 2 and counts['gates'] > 50, counts     # the window refills and prunes again after a reset
    assert max(worst['imu'], worst['P'], worst['cams'], worst['extr'], worst['P_block']) < TOL, worst


def test_reset_method():
    """reset() (msckf.py:832-857): IMU state back to its initial values (extrinsics and id kept), empty window and map,
    initial covariance, gravity / bias initialisation armed again; the stream then restarts identically on both sides."""
    n_frames = 30
    stream = synth.make_stream(n_frames, 80, structure_seed=7, noise_seed=4)
    ocfg = orc.ConfigEuRoC()
    ocfg._vio_max_cam_state_size__ = 6
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = 6
    of, df = orc.MSCKF(ocfg), MSCKF(dcfg, feature_capacity=512)
    try:
        synth.drive(of, stream, 14)
        synth.drive(df, stream, 14)
        assert len(snapshot(df)['cam_ids']) >= 4
        of.reset()
        df.reset()
        so, sd = snapshot(of), snapshot(df)
        assert len(sd['cam_ids']) == 0 and len(sd['map_ids']) == 0 and sd['P'].shape == (21, 21)
        assert np.array_equal(sd['P'], so['P']) and np.array_equal(sd['imu'], so['imu'])
        assert rel_err(sd['extr'], so['extr']) < TOL                 # the calibrated extrinsics survive a reset
        assert not df.is_gravity_set and df.is_first_img and len(df.imu_msg_buffer) == 0
        # restart: a static lead-in for the gravity initialisation, then frames again (a second realisation of the stream)
        again = synth.realize(stream.structure, noise_seed=9)
        worst = 0.0
        msgs = again.imu_msgs()
        k = 0
        for f in range(n_frames):
            while k < len(msgs) and msgs[k].vio_timestamp__ <= again.frame_ts[f]:
                of.imu_callback(msgs[k])
                df.imu_callback(msgs[k])
                k += 1
            ro, rd = of.feature_callback(again.feature_msg(f)), df.feature_callback(again.feature_msg(f))
            assert (ro is None) == (rd is None) and ro is not None
            so, sd = snapshot(of), snapshot(df)
            assert np.array_equal(sd['cam_ids'], so['cam_ids']) and np.array_equal(sd['map_ids'], so['map_ids']), f
            worst = max(worst, rel_err(sd['imu'], so['imu']), rel_err(sd['P'], so['P']), rel_err(sd['cams'], so['cams']))
        assert worst < TOL, worst
    finally:
        df.close()


def test_stress_long_tracks_cap_lifted():
    """configs[2] in small: a 30-state window, long tracks, the row cap lifted, and a final frame in which EVERY live
    landmark is lost at once (its ids are replaced): long-track Jacobians (one CTA per landmark), the general
    stacking + QR kernel with up to 174 active columns and the tiled update."""
    n_frames = 44
    st = synth.make_structure(n_frames, 120, structure_seed=3, mean_track_len=40.0)
    st.frame_ids[-1] = st.frame_ids[-1] + 10 ** 6
    stream = synth.realize(st, noise_seed=7)
    worst, counts = run_pair(stream, 30, n_frames, feature_capacity=1024, row_cap=10 ** 9)
    assert counts['gates'] > 500 and counts['updates'] > 20, counts
    assert max(worst['imu'], worst['P'], worst['cams'], worst['extr'], worst['P_block']) < TOL, worst
    assert worst['pos_m'] < 1e-3, worst


def test_large_window_64():
    """configs[4] in small: max_cam_state_size = 64 (state dimension 405): Jacobian scratch in global memory, the
    QR fallback buffers and the tiled update on the full-size covariance."""
    n_frames = 76
    st = synth.make_structure(n_frames, 100, structure_seed=5, mean_track_len=30.0)
    stream = synth.realize(st, noise_seed=9)
    worst, counts = run_pair(stream, 64, n_frames, feature_capacity=1024)
    assert counts['gates'] > 500 and counts['prunes'] >= 3, counts
    assert max(worst['imu'], worst['P'], worst['cams'], worst['extr'], worst['P_block']) < TOL, worst
    assert worst['pos_m'] < 1e-3, worst


def test_stream_row_cap_break():
    """The row-cap break of remove_lost_features (msckf.py:676-679) with a small cap so that it triggers."""
    n_frames = 40
    stream = synth.realize(synth.make_structure(n_frames, 200, structure_seed=2, mean_track_len=4.0), noise_seed=5)
    worst, counts = run_pair(stream, 10, n_frames, feature_capacity=1024, row_cap=60)
    assert counts['gates'] > 200, counts
    assert max(worst['imu'], worst['P'], worst['cams']) < TOL, worst


def test_batch_of_filters_vs_oracle():
    """configs[3] shape: independent filters (noise seeds) in one context, each equal to its own oracle run."""
    F, n_frames, max_cam = 6, 30, 8
    st = synth.make_structure(n_frames, 80, structure_seed=3)
    streams = [synth.realize(st, noise_seed=100 + k) for k in range(F)]
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    batch = MSCKFBatch(dcfg, F, feature_capacity=512)
    try:
        s0 = streams[0]
        batch.initialize_gravity_and_bias(np.stack([s.gyro[:200] for s in streams]), np.stack([s.acc[:200] for s in streams]))
        traj = np.zeros((F, n_frames, 3))
        for f in range(n_frames):
            a, b = s0.imu_range_for_frame(f)
            imu, rst = batch.feature_callback(float(s0.frame_ts[f]), s0.imu_ts[a:b], np.stack([s.gyro[a:b] for s in streams]),
                                              np.stack([s.acc[a:b] for s in streams]), s0.frame_ids[f],
                                              np.stack([s.frame_uv[f] for s in streams]))
            assert not rst.any()
            traj[:, f] = imu[:, capi.IMU_P:capi.IMU_P + 3]
        Ps = [batch.ctx.get_cov(k) for k in range(F)]
        cams = [batch.ctx.get_cam_states(k) for k in range(F)]
        feats = [batch.ctx.get_features(k, with_obs=False)[0] for k in range(F)]
    finally:
        batch.close()
    for k in range(F):
        ocfg = orc.ConfigEuRoC()
        ocfg._vio_max_cam_state_size__ = max_cam
        of = orc.MSCKF(ocfg)
        otraj = []
        synth.drive(of, streams[k], n_frames, on_frame=lambda f, flt, res: otraj.append(flt.state_server._vio_imu_state__._vio_position__.copy()))
        so = snapshot(of)
        assert np.max(np.linalg.norm(traj[k] - np.array(otraj), axis=1)) < 1e-3
        assert rel_err(traj[k], np.array(otraj)) < TOL
        assert rel_err(Ps[k], so['P']) < TOL
        assert np.array_equal(cams[k][0], so['cam_ids'])
        assert np.array_equal(feats[k], so['map_ids'])


def test_two_frames_in_flight_equal_sequential():
    """msckf_feature_callback_async with two frames in flight (inputs in page-locked memory, copied on the copy stream)
    returns, frame by frame, exactly what the blocking call returns."""
    F, n_frames, max_cam = 4, 24, 6
    st = synth.make_structure(n_frames, 60, structure_seed=5)
    streams = [synth.realize(st, noise_seed=200 + k) for k in range(F)]
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    s0 = streams[0]

    def frame_args(batch, f, pinned):
        a, b = s0.imu_range_for_frame(f)
        take = batch.select_imu(float(s0.frame_ts[f]), s0.imu_ts[a:b])
        ts = s0.imu_ts[a:b][take]
        K, n = len(ts), len(s0.frame_ids[f])
        cam_id = batch.next_imu_id
        batch.next_imu_id += 1
        args = [(np.arange(F + 1) * K).astype(np.int32), np.tile(ts, F),
                np.ascontiguousarray(np.stack([s.gyro[a:b][take] for s in streams])),
                np.ascontiguousarray(np.stack([s.acc[a:b][take] for s in streams])),
                np.full(F, cam_id, dtype=np.int64), (np.arange(F + 1) * n).astype(np.int32),
                np.tile(s0.frame_ids[f].astype(np.int64), F), np.ascontiguousarray(np.stack([s.frame_uv[f] for s in streams]))]
        return [capi.pinned_copy(x) for x in args] if pinned else args

    outs = {}
    for mode in ('blocking', 'pipelined'):
        batch = MSCKFBatch(dcfg, F, feature_capacity=256)
        try:
            batch.initialize_gravity_and_bias(np.stack([s.gyro[:200] for s in streams]), np.stack([s.acc[:200] for s in streams]))
            res = []
            if mode == 'blocking':
                for f in range(n_frames):
                    res.append(batch.ctx.feature_callback(*frame_args(batch, f, False))[0])
            else:
                args = [frame_args(batch, f, True) for f in range(n_frames)]
                batch.ctx.feature_callback_async(*args[0])
                for f in range(1, n_frames):
                    batch.ctx.feature_callback_async(*args[f])
                    res.append(batch.ctx.collect()[0])
                res.append(batch.ctx.collect()[0])
            outs[mode] = (np.array(res), [batch.ctx.get_cov(k) for k in range(F)])
        finally:
            batch.close()
    assert np.array_equal(outs['blocking'][0], outs['pipelined'][0])
    for k in range(F):
        assert np.array_equal(outs['blocking'][1][k], outs['pipelined'][1][k])


def test_filter_groups_equal_one_context():
    """A batch spread over two contexts (MSCKFBatch(groups=2), uneven split 3 + 2) returns bit for bit what one context
    returns: through the blocking host call, through device-resident messages (device_frame_parts) and through the staged
    calls, with ragged per-filter message lengths."""
    import torch
    F, n_frames, max_cam = 5, 26, 6
    st = synth.make_structure(n_frames, 70, structure_seed=9)
    streams = [synth.realize(st, noise_seed=300 + k) for k in range(F)]
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    s0 = streams[0]
    keys = ['imu_off', 'ts', 'gyro', 'acc', 'cam_ids', 'feat_off', 'ids', 'uv']

    def frame_args(batch, f):
        """ragged: filter k drops its last k features of the frame"""
        a, b = s0.imu_range_for_frame(f)
        take = batch.select_imu(float(s0.frame_ts[f]), s0.imu_ts[a:b])
        ts = s0.imu_ts[a:b][take]
        K = len(ts)
        cam_id = batch.next_imu_id
        batch.next_imu_id += 1
        ns = [max(0, len(s0.frame_ids[f]) - k) for k in range(F)]
        return [(np.arange(F + 1) * K).astype(np.int32), np.tile(ts, F),
                np.ascontiguousarray(np.stack([s.gyro[a:b][take] for s in streams])),
                np.ascontiguousarray(np.stack([s.acc[a:b][take] for s in streams])),
                np.full(F, cam_id, dtype=np.int64), np.concatenate([[0], np.cumsum(ns)]).astype(np.int32),
                np.concatenate([s0.frame_ids[f][:n].astype(np.int64) for n in ns]),
                np.concatenate([s.frame_uv[f][:n].reshape(-1, 4) for s, n in zip(streams, ns)])]

    outs = {}
    for mode, groups in (('one', 1), ('host', 2), ('device', 2), ('staged', 2)):
        batch = MSCKFBatch(dcfg, F, feature_capacity=256, groups=groups)
        try:
            assert batch.groups == groups and batch.ctx.F == F
            batch.initialize_gravity_and_bias(np.stack([s.gyro[:200] for s in streams]), np.stack([s.acc[:200] for s in streams]))
            res = []
            for f in range(n_frames):
                args = frame_args(batch, f)
                if mode in ('one', 'host'):
                    imu, rst = batch.ctx.feature_callback(*args)
                elif mode == 'device':
                    d = dict((k, torch.from_numpy(np.ascontiguousarray(v)).cuda()) for k, v in zip(keys, args))
                    out = torch.zeros((F, capi.IMU_STRIDE), dtype=torch.float64, device='cuda')
                    batch.feature_callback_dev(batch.device_frame_parts(d, keys), out)
                    batch.ctx.sync()
                    imu, rst = out.cpu().numpy(), np.zeros(F, dtype=bool)
                else:
                    c = batch.ctx
                    c.propagate(*args[:4])
                    c.augment(args[4])
                    c.add_observations(*args[5:])
                    c.remove_lost_features()
                    c.prune_cam_states()
                    rst = c.online_reset()
                    imu = c.get_imu_states()
                assert not rst.any()
                res.append(imu)
            cnt = batch.ctx.counters()
            assert cnt.shape[0] == F and batch.ctx.update_shapes().shape[0] == F and batch.ctx.kernel_launches() > 0
            outs[mode] = (np.array(res), [batch.ctx.get_cov(k) for k in range(F)], [batch.ctx.get_cam_states(k)[0] for k in range(F)],
                          [batch.ctx.get_features(k, with_obs=False)[0] for k in range(F)], cnt[:, capi.CNT_FEATURE_UPDATES])
        finally:
            batch.close()
    assert outs['one'][4].sum() > 100
    for mode in ('host', 'device', 'staged'):
        assert np.array_equal(outs['one'][0], outs[mode][0]), mode
        assert np.array_equal(outs['one'][4], outs[mode][4]), mode
        for k in range(F):
            assert np.array_equal(outs['one'][1][k], outs[mode][1][k]), (mode, k)
            assert np.array_equal(outs['one'][2][k], outs[mode][2][k]), (mode, k)
            assert np.array_equal(outs['one'][3][k], outs[mode][3][k]), (mode, k)


def test_euroc_replay_through_publishers_vs_oracle(tmp_path):
    """a sequence on disk in the EuRoC layout, replayed by two DataPublisher threads into queues and consumed
    the way vio.py does (IMU thread -> imu_callback, image thread -> feature_callback), drives the CUDA drop-in; the oracle
    fed with the same messages in the same order agrees (1e-9, ids exact) and the ATE against the ground-truth file is small."""
    from queue import Queue
    import time
    from msckf_b200 import dataset
    stream = synth.make_stream(50, 90, structure_seed=6, noise_seed=12)
    dataset.write_euroc(stream, str(tmp_path), t0_ns=0)
    ds = dataset.EuRoCDataset(str(tmp_path))
    q_imu, q_feat = Queue(), Queue()
    pubs = [dataset.DataPublisher(ds.imu, q_imu, ratio=50.0), dataset.DataPublisher(ds.features, q_feat, ratio=50.0)]
    now = time.time()
    for p in pubs:
        p.start(now)
    imu_msgs, feat_msgs = [], []
    for q, out in ((q_imu, imu_msgs), (q_feat, feat_msgs)):
        while True:
            m = q.get(timeout=60)
            if m is None:
                break
            out.append(m)
    for p in pubs:
        p.stop()
    assert len(imu_msgs) == len(stream.imu_ts) and len(feat_msgs) == stream.n_frames
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = 10
    ocfg = orc.ConfigEuRoC()
    ocfg._vio_max_cam_state_size__ = 10
    df, of = MSCKF(dcfg, feature_capacity=512), orc.MSCKF(ocfg)
    worst, k = 0.0, 0
    est_t, est_p = [], []
    try:
        for msg in feat_msgs:
            while k < len(imu_msgs) and imu_msgs[k].vio_timestamp__ <= msg.timestamp:
                df.imu_callback(imu_msgs[k])
                of.imu_callback(imu_msgs[k])
                k += 1
            rd, ro = df.feature_callback(msg), of.feature_callback(msg)
            assert (rd is None) == (ro is None)
            sd, so = snapshot(df), snapshot(of)
            assert np.array_equal(sd['cam_ids'], so['cam_ids']) and np.array_equal(sd['map_ids'], so['map_ids'])
            worst = max(worst, rel_err(sd['imu'], so['imu']), rel_err(sd['P'], so['P']), rel_err(sd['cams'], so['cams']))
            est_t.append(msg.timestamp)
            est_p.append(sd['imu'][13:16].copy())
    finally:
        df.close()
    assert worst < TOL, worst
    assert dataset.absolute_trajectory_error(est_t, est_p, ds.groundtruth) < 0.25


def test_batch_of_different_sequences_ragged():
    """MSCKFBatch.feature_callback_ragged: the filters of one context follow DIFFERENT sequences (own landmark fields, own
    feature counts and ids per frame); each equals the single-filter drop-in run on its own stream."""
    F, n_frames, max_cam = 3, 28, 7
    streams = [synth.make_stream(n_frames, 60 + 25 * k, structure_seed=20 + k, noise_seed=40 + k) for k in range(F)]
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    batch = MSCKFBatch(dcfg, F, feature_capacity=512)
    try:
        batch.initialize_gravity_and_bias(np.stack([s.gyro[:200] for s in streams]), np.stack([s.acc[:200] for s in streams]))
        for f in range(n_frames):
            rng = [s.imu_range_for_frame(f) for s in streams]
            batch.feature_callback_ragged(float(streams[0].frame_ts[f]), [s.imu_ts[a:b] for s, (a, b) in zip(streams, rng)],
                                          [s.gyro[a:b] for s, (a, b) in zip(streams, rng)], [s.acc[a:b] for s, (a, b) in zip(streams, rng)],
                                          [s.frame_ids[f] for s in streams], [s.frame_uv[f] for s in streams])
        got = [(batch.ctx.get_imu_state(k), batch.ctx.get_cov(k), batch.ctx.get_cam_states(k)[0],
                batch.ctx.get_features(k, with_obs=False)[0]) for k in range(F)]
    finally:
        batch.close()
    assert len(set(len(g[3]) for g in got)) > 1               # the maps really differ in size
    for k in range(F):
        single = MSCKF(dcfg, feature_capacity=512)
        try:
            synth.drive(single, streams[k])
            s1 = snapshot(single)
        finally:
            single.close()
        b = got[k][0]
        imu = np.concatenate([b[capi.IMU_Q:capi.IMU_Q + 4], b[capi.IMU_BG:capi.IMU_BG + 3], b[capi.IMU_V:capi.IMU_V + 3],
                              b[capi.IMU_BA:capi.IMU_BA + 3], b[capi.IMU_P:capi.IMU_P + 3]])
        assert rel_err(imu, s1['imu']) < 1e-12 and rel_err(got[k][1], s1['P']) < 1e-12
        assert np.array_equal(got[k][2], s1['cam_ids']) and np.array_equal(got[k][3], s1['map_ids'])


def test_two_devices_in_one_process():
    """Kernel attributes (dynamic shared-memory opt-in) and the SM count are per device: two contexts on two GPUs in one
    process must both run the wide kernels.  Needs a box with at least two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('one GPU on this box')
    n_frames = 24
    stream = synth.make_stream(n_frames, 100, structure_seed=2, noise_seed=6)
    outs = []
    for dev in (1, 0):                      # the second device first: nothing may rely on device 0 having run before
        cfg = ConfigEuRoC()
        cfg._vio_max_cam_state_size__ = 16
        filt = MSCKF(cfg, device=dev, feature_capacity=512)
        try:
            synth.drive(filt, stream)
            outs.append(snapshot(filt))
        finally:
            filt.close()
    assert np.array_equal(outs[0]['imu'], outs[1]['imu']) and np.array_equal(outs[0]['P'], outs[1]['P'])
    assert np.array_equal(outs[0]['map_ids'], outs[1]['map_ids'])


def test_graph_replay_equals_eager_on_ragged_frames(monkeypatch):
    """The captured frame graph is replayed with new message pointers every frame; its launch dimensions must not depend
    on the message.  Filters of one batch receive different numbers of features and IMU samples per frame (ragged
    offsets of the C ABI) and the counts change from frame to frame: the replayed graph returns, bit for bit, what
    eager launches return."""
    F, n_frames, max_cam = 3, 26, 6
    st = synth.make_structure(n_frames, 70, structure_seed=9)
    streams = [synth.realize(st, noise_seed=300 + k) for k in range(F)]
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    s0 = streams[0]

    def run():
        batch = MSCKFBatch(dcfg, F, feature_capacity=256)
        try:
            batch.initialize_gravity_and_bias(np.stack([s.gyro[:200] for s in streams]), np.stack([s.acc[:200] for s in streams]))
            out = []
            for f in range(n_frames):
                a, b = s0.imu_range_for_frame(f)
                take = np.flatnonzero(batch.select_imu(float(s0.frame_ts[f]), s0.imu_ts[a:b])) + a
                ids = s0.frame_ids[f].astype(np.int64)
                imu_off, feat_off, ts, gy, ac, fid, uv = [0], [0], [], [], [], [], []
                for k in range(F):
                    tk = take[(k + f) % 2:] if len(take) > 3 else take          # some filters skip their first sample
                    nk = len(ids) - ((3 * k + f) % 11)                          # and see fewer features
                    ts.append(s0.imu_ts[tk]); gy.append(streams[k].gyro[tk]); ac.append(streams[k].acc[tk])
                    fid.append(ids[:nk]); uv.append(streams[k].frame_uv[f][:nk])
                    imu_off.append(imu_off[-1] + len(tk)); feat_off.append(feat_off[-1] + nk)
                cam_ids = np.full(F, batch.next_imu_id, dtype=np.int64)
                batch.next_imu_id += 1
                imu, rst = batch.ctx.feature_callback(imu_off, np.concatenate(ts), np.concatenate(gy), np.concatenate(ac), cam_ids,
                                                      feat_off, np.concatenate(fid), np.concatenate(uv))
                out.append(imu)
            return np.array(out), [batch.ctx.get_cov(k) for k in range(F)], [batch.ctx.get_features(k, with_obs=False)[0] for k in range(F)]
        finally:
            batch.close()

    g_imu, g_P, g_ids = run()
    monkeypatch.setenv('MSCKF_NO_GRAPH', '1')
    e_imu, e_P, e_ids = run()
    assert np.array_equal(g_imu, e_imu)
    for k in range(F):
        assert np.array_equal(g_P[k], e_P[k]) and np.array_equal(g_ids[k], e_ids[k])
    assert not np.array_equal(g_P[0], g_P[1])                  # the filters really did see different messages


@pytest.mark.parametrize('knob', ['MSCKF_NO_DMMA', 'MSCKF_NO_LANES', 'MSCKF_NO_JAC_ROWS', 'MSCKF_NO_QR_TALL', 'MSCKF_NO_GRAPH'])
def test_alternative_kernel_paths(knob, monkeypatch):
    """The library's debugging knobs route the same work through its other kernels: the tiled update kernels instead of
    the one-CTA DMMA update (the path wide windows take), in-order launches instead of concurrent size classes, the
    group-per-landmark Jacobian kernel instead of the lane-per-row one, the blocked tensor-pipe QR instead of the
    tall-skinny TSQR for the pruning stacks, eager launches instead of the captured frame graph.  Each must pass the same stepped comparison."""
    monkeypatch.setenv(knob, '1')
    n_frames = 40
    stream = synth.make_stream(n_frames, 120, structure_seed=6, noise_seed=21)
    worst, counts = run_pair(stream, 10, n_frames, feature_capacity=512)
    assert counts['gates'] > 300 and counts['prunes'] >= 5 and counts['updates'] > 20, counts
    assert worst['imu'] < TOL and worst['P'] < TOL and worst['cams'] < TOL and worst['extr'] < TOL, worst
    assert worst['pos_m'] < 1e-3, worst


def test_update_shapes_of_the_pruning_stage():
    """msckf_get_update_shapes after feature_callback describes the pruning update (the last stage of a frame): every
    landmark sits on the two camera states being removed, so n_a = 12, k = min(m, 12), the one-CTA class-0 update runs,
    and m is the row count of the decision log.  These are the stacks k_qr_tall compresses."""
    n_frames, max_cam = 26, 6
    stream = synth.make_stream(n_frames, 100, structure_seed=5, noise_seed=3)
    dcfg = ConfigEuRoC()
    dcfg._vio_max_cam_state_size__ = max_cam
    df = MSCKF(dcfg, feature_capacity=512)
    msgs = stream.imu_msgs()
    k = 0
    pruned = tall = 0
    try:
        for f in range(n_frames):
            while k < len(msgs) and msgs[k].vio_timestamp__ <= stream.frame_ts[f]:
                df.imu_callback(msgs[k])
                k += 1
            df.feature_callback(stream.feature_msg(f))
            m, na, kr, path = df.ctx.update_shapes()[0]
            log = df.ctx.stage_log(0, 1)
            assert m == log['update_rows']
            if m == 0:
                assert path == 0 and kr == 0
                continue
            pruned += 1
            assert na == 12 and kr == min(m, 12) and path == 5, (m, na, kr, path)
            assert m % 5 == 0                       # 4 M - 3 rows per landmark with M = 2
            tall += int(m > 12)
    finally:
        df.close()
    assert pruned >= 5 and tall >= 5, (pruned, tall)
