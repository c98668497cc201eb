"""EuRoC on-disk reader and online replay (msckf_b200/dataset.py, mirror of the reference's dataset.py) on a
synthetic sequence written in the EuRoC directory layout; the replayed messages drive the CPU oracle here and the CUDA
drop-in in tests/test_gpu_stream.py."""
import os
import sys
import time
from queue import Queue

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

from msckf_b200 import dataset, synth       # noqa: E402
from oracle import msckf_oracle as orc      # noqa: E402


def make_sequence(tmp_path, n_frames=40, feats=60, t0_ns=0):
    stream = synth.make_stream(n_frames, feats, structure_seed=3, noise_seed=8)
    dataset.write_euroc(stream, str(tmp_path), t0_ns=t0_ns)
    return stream, dataset.EuRoCDataset(str(tmp_path))


def test_readers_round_trip(tmp_path):
    stream, ds = make_sequence(tmp_path)
    imu = list(ds.imu)
    assert len(imu) == len(stream.imu_ts) and imu[0]._fields == ('vio_timestamp__', 'angular_velocity', 'linear_acceleration')
    assert np.allclose([m.vio_timestamp__ for m in imu], stream.imu_ts, rtol=0, atol=1e-9)
    assert np.array_equal(np.array([m.angular_velocity for m in imu]), stream.gyro)       # repr() round-trips doubles exactly
    assert np.array_equal(np.array([m.linear_acceleration for m in imu]), stream.acc)
    gt = list(ds.groundtruth)
    assert len(gt) == len(stream.imu_ts) and gt[0]._fields == ('vio_timestamp__', 'vio_p__', 'vio_q__', 'vio_v__', 'bw_', 'ba_')
    assert np.allclose(gt[-1].vio_p__, synth.truth_position(stream.imu_ts[-1]), atol=1e-12)
    assert np.allclose(gt[-1].bw_, stream.structure.gyro_bias) and abs(np.linalg.norm(gt[-1].vio_q__) - 1) < 1e-12
    feats = list(ds.features)
    assert len(feats) == stream.n_frames
    for f, msg in enumerate(feats):
        assert abs(msg.timestamp - stream.frame_ts[f]) < 1e-9
        assert [m.id for m in msg.vio_features] == [int(i) for i in stream.frame_ids[f]]
        assert np.array_equal(np.array([(m.u0, m.v0, m.u1, m.v1) for m in msg.vio_features]), stream.frame_uv[f])
    # start-time offsets skip the beginning of every stream (dataset.py:218-223)
    ds.set_starttime(1.0)
    t_first = ds.starttime + 1.0
    assert next(iter(ds.imu)).vio_timestamp__ >= t_first and next(iter(ds.features)).timestamp >= t_first
    assert next(iter(ds.groundtruth)).vio_timestamp__ >= t_first


def test_dataset_clock_and_missing_images(tmp_path):
    """Timestamps in the dataset's own epoch (1.4e18 ns) survive as float seconds to ~2.4e-7 s; a track-only sequence has no
    image folders and still opens."""
    stream, ds = make_sequence(tmp_path, n_frames=5, t0_ns=1403636579000000000)
    assert ds.stereo is None and ds.cam0 is None and ds.features is not None
    t = np.array([m.vio_timestamp__ for m in ds.imu])
    assert np.allclose(np.diff(t), synth.IMU_DT, atol=1e-6) and abs(t[0] - 1403636579.0) < 1e-6


def test_publisher_replays_in_order_and_stops(tmp_path):
    stream, ds = make_sequence(tmp_path, n_frames=20)
    q_imu, q_feat = Queue(), Queue()
    duration = 1.8                         # seconds of data after the dataset start
    p_imu = dataset.DataPublisher(ds.imu, q_imu, duration, ratio=200.0)
    p_feat = dataset.DataPublisher(ds.features, q_feat, duration, ratio=200.0)
    now = time.time()
    p_imu.start(now)
    p_feat.start(now)
    imu, feat = [], []
    for q, out in ((q_imu, imu), (q_feat, feat)):
        while True:
            m = q.get(timeout=20)
            if m is None:
                break
            out.append(m)
    p_imu.stop()
    p_feat.stop()
    ti = np.array([m.vio_timestamp__ for m in imu])
    assert len(ti) > 100 and np.all(np.diff(ti) > 0) and ti[-1] <= ds.starttime + duration + 2e-3
    assert np.sum(stream.imu_ts <= ds.starttime + duration + 1e-3) == len(ti)
    assert [abs(m.timestamp - t) < 1e-9 for m, t in zip(feat, stream.frame_ts)] == [True] * len(feat)
    assert len(feat) == int(np.sum(stream.frame_ts <= ds.starttime + duration + 1e-3))


def test_replay_drives_the_filter_and_ate(tmp_path):
    """The replayed messages through imu_callback / feature_callback (what vio.py does with the publisher queues) run the
    filter; its ATE against the ground-truth FILE equals the error against the generator's own truth.  (A frame-by-frame
    comparison with synth.drive is not meaningful: frame and IMU times coincide, and after the trip through integer
    nanoseconds the boundary sample of a frame may fall on the other side of the `<=`.)"""
    stream, ds = make_sequence(tmp_path, n_frames=60, feats=80)
    replay = orc.MSCKF(orc.ConfigEuRoC())
    imu_iter = iter(ds.imu)
    pending = next(imu_iter)
    est_t, est_p = [], []
    for msg in ds.features:
        while pending is not None and pending.vio_timestamp__ <= msg.timestamp:
            replay.imu_callback(pending)
            pending = next(imu_iter, None)
        out = replay.feature_callback(msg)
        if out is not None:
            est_t.append(msg.timestamp)
            est_p.append(replay.state_server._vio_imu_state__._vio_position__.copy())
    assert len(est_p) == stream.n_frames
    ate = dataset.absolute_trajectory_error(est_t, est_p, ds.groundtruth)
    truth = synth.truth_position(np.array(est_t))
    direct_rmse = float(np.sqrt(np.mean(np.sum(((np.array(est_p) - est_p[0]) - (truth - truth[0])) ** 2, axis=1))))
    assert abs(ate - direct_rmse) < 1e-6 and ate < 0.25, (ate, direct_rmse)
