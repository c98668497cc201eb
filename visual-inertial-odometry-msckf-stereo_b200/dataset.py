"""EuRoC MAV on-disk reader and online replay with the reference's interface (dataset.py:12-271), SURVEY.md section 8f-4.

Same class names, constructor arguments, message namedtuples and attribute names as the reference (``EuRoCDataset``,
``GroundTruthReader``, ``IMUDataReader``, ``ImageReader``, ``Stereo``, ``DataPublisher``), so ``vio.py`` can import this
module in place of the reference's.  Host I/O only: nothing here is on the filter's hot path.

Two additions, because the image front end (image.py, cv2 KLT / FAST) is out of scope of this repo:

* ``FeatureTrackReader`` reads pre-computed stereo feature tracks (``mav0/features0/data.csv``: timestamp [ns], id, u0, v0,
  u1, v1 in normalised coordinates) and yields the ``vio_feature_msg__`` messages the back end consumes
  (image.py:425-438), and ``EuRoCDataset.features`` exposes it when that file exists;
* ``write_euroc`` writes a synthetic stream (synth.py) in the EuRoC directory layout, so that the readers, the replay
  and the trajectory evaluation (``absolute_trajectory_error``) can be exercised without the dataset, which is not
  available offline.

The ground-truth message carries six fields (timestamp, p, q, v, b_w, b_a); the reference declares five and passes six
(dataset.py:16,33), which raises on the first line read -- the field list is completed here.
"""
import os
import time
from collections import namedtuple
from threading import Thread

import numpy as np

GtMsg = namedtuple('gt_msg', ['vio_timestamp__', 'vio_p__', 'vio_q__', 'vio_v__', 'bw_', 'ba_'])
ImuMsg = namedtuple('imu_msg', ['vio_timestamp__', 'angular_velocity', 'linear_acceleration'])
ImgMsg = namedtuple('img_msg', ['vio_timestamp__', 'image'])
StereoMsg = namedtuple('stereo_msg', ['vio_timestamp__', 'cam0_image', 'cam1_image', 'cam0_msg', 'cam1_msg'])
FeatureMsg = namedtuple('vio_feature_msg__', ['timestamp', 'vio_features'])


class FeatureMeasurement(object):
    __slots__ = ('id', 'u0', 'v0', 'u1', 'v1')

    def __init__(self, fid, u0, v0, u1, v1):
        self.id, self.u0, self.v0, self.u1, self.v1 = fid, u0, v0, u1, v1


class _CsvReader(object):
    """A csv file with one header line, replayed from ``starttime`` on."""

    def __init__(self, _vio_path__, scaler, starttime=-float('inf')):
        self.scaler = scaler                 # converts the timestamp column (ns) to seconds
        self._vio_path__ = _vio_path__
        self.starttime = starttime

    def _rows(self):
        with open(self._vio_path__, 'r') as f:
            next(f)
            for line in f:
                line = line.strip()
                if line:
                    yield [float(x) for x in line.split(',')]

    def set_starttime(self, starttime):
        self.starttime = starttime

    def start_time(self):
        for row in self._rows():
            return row[0] * self.scaler
        return None

    def __iter__(self):
        for row in self._rows():
            msg = self.parse_row(row)
            if msg[0] < self.starttime:
                continue
            yield msg


class GroundTruthReader(_CsvReader):
    """dataset.py:12-47: timestamp, p_RS_R (3), q_RS (w x y z), v_RS_R (3), b_w (3), b_a (3)."""
    field = GtMsg

    def parse(self, _vio_line__):
        return self.parse_row([float(x) for x in _vio_line__.strip().split(',')])

    def parse_row(self, v):
        return GtMsg(v[0] * self.scaler, np.array(v[1:4]), np.array(v[4:8]), np.array(v[8:11]), np.array(v[11:14]),
                     np.array(v[14:17]))


class IMUDataReader(_CsvReader):
    """dataset.py:51-92: timestamp, w_RS_S (3) [rad/s], a_RS_S (3) [m/s^2]."""
    field = ImuMsg

    def parse(self, _vio_line__):
        return self.parse_row([float(x) for x in _vio_line__.strip().split(',')])

    def parse_row(self, v):
        return ImuMsg(v[0] * self.scaler, np.array(v[1:4]), np.array(v[4:7]))


class FeatureTrackReader(_CsvReader):
    """Pre-computed stereo tracks: rows (timestamp [ns], id, u0, v0, u1, v1), grouped by timestamp into feature messages."""
    field = FeatureMsg

    def start_time(self):
        for row in self._rows():
            return row[0] * self.scaler
        return None

    def __iter__(self):
        cur_t, feats = None, []
        for row in self._rows():
            t = row[0] * self.scaler
            if cur_t is not None and t != cur_t:
                if cur_t >= self.starttime:
                    yield FeatureMsg(cur_t, feats)
                feats = []
            cur_t = t
            feats.append(FeatureMeasurement(int(row[1]), row[2], row[3], row[4], row[5]))
        if cur_t is not None and cur_t >= self.starttime:
            yield FeatureMsg(cur_t, feats)


class ImageReader(object):
    """dataset.py:96-167: lazily decoded image files with their timestamps (decoding needs cv2, as in the reference)."""

    def __init__(self, ids, _vio_timestamps__, starttime=-float('inf')):
        self.ids = ids
        self._vio_timestamps__ = _vio_timestamps__
        self.starttime = starttime
        self.cache = dict()
        self._vio_idx__ = 0
        self.field = ImgMsg

    def read(self, _vio_path__):
        import cv2                      # only the front end needs pixels; the back end never calls this
        return cv2.imread(_vio_path__, -1)

    def __len__(self):
        return len(self.ids)

    def __getitem__(self, _vio_idx__):
        self._vio_idx__ = _vio_idx__
        if _vio_idx__ in self.cache:
            return self.cache.pop(_vio_idx__)
        return self.read(self.ids[_vio_idx__])

    def __iter__(self):
        for i, t in enumerate(self._vio_timestamps__):
            if t < self.starttime:
                continue
            yield ImgMsg(t, self[i])

    def start_time(self):
        return self._vio_timestamps__[0]

    def set_starttime(self, starttime):
        self.starttime = starttime


class Stereo(object):
    """dataset.py:171-196"""

    def __init__(self, cam0, cam1):
        assert len(cam0) == len(cam1)
        self.cam0 = cam0
        self.cam1 = cam1
        self._vio_timestamps__ = cam0._vio_timestamps__
        self.starttime = cam0.starttime
        self.field = StereoMsg

    def __iter__(self):
        for left, right in zip(self.cam0, self.cam1):
            assert abs(left.vio_timestamp__ - right.vio_timestamp__) < 0.01, 'unsynced stereo pair'
            yield StereoMsg(left.vio_timestamp__, left.image, right.image, left, right)

    def __len__(self):
        return len(self.cam0)

    def start_time(self):
        return self.cam0.start_time()

    def set_starttime(self, starttime):
        self.starttime = starttime
        self.cam0.set_starttime(starttime)
        self.cam1.set_starttime(starttime)


class EuRoCDataset(object):
    """dataset.py:200-232 -- ``_vio_path__`` is e.g. '.../EuRoC MAV Dataset/MH_01_easy'.  The image folders are optional
    (a track-only export has none); ``features`` is set when mav0/features0/data.csv exists."""

    def __init__(self, _vio_path__):
        mav = os.path.join(_vio_path__, 'mav0')
        self.groundtruth = GroundTruthReader(os.path.join(mav, 'state_groundtruth_estimate0', 'data.csv'), 1e-9)
        self.imu = IMUDataReader(os.path.join(mav, 'imu0', 'data.csv'), 1e-9)
        starts = [self.imu.start_time()]
        self.cam0 = self.cam1 = self.stereo = None
        if os.path.isdir(os.path.join(mav, 'cam0', 'data')) and os.path.isdir(os.path.join(mav, 'cam1', 'data')):
            self.cam0 = ImageReader(*self.list_imgs(os.path.join(mav, 'cam0', 'data')))
            self.cam1 = ImageReader(*self.list_imgs(os.path.join(mav, 'cam1', 'data')))
            if len(self.cam0):
                self.stereo = Stereo(self.cam0, self.cam1)
                self._vio_timestamps__ = self.cam0._vio_timestamps__
                starts.append(self.stereo.start_time())
        self.features = None
        fpath = os.path.join(mav, 'features0', 'data.csv')
        if os.path.exists(fpath):      # (tracks do not move the start: a static IMU lead-in before the first frame is kept)
            self.features = FeatureTrackReader(fpath, 1e-9)
        self.starttime = max(s for s in starts if s is not None)
        self.set_starttime(0)

    def set_starttime(self, offset):
        for r in (self.groundtruth, self.imu, self.cam0, self.cam1, self.stereo, self.features):
            if r is not None:
                r.set_starttime(self.starttime + offset)

    def list_imgs(self, dir):
        xs = sorted((x for x in os.listdir(dir) if x.endswith('.png')), key=lambda x: float(x[:-4]))
        return [os.path.join(dir, x) for x in xs], [float(x[:-4]) * 1e-9 for x in xs]


class DataPublisher(object):
    """dataset.py:236-271: replays a reader into a queue in (scaled) real time; ``None`` marks the end."""

    def __init__(self, _vio_dataset__, out_queue, _vio_duration__=float('inf'), ratio=1.):
        self._vio_dataset__ = _vio_dataset__
        self.dataset_starttime = _vio_dataset__.starttime
        self.out_queue = out_queue
        self._vio_duration__ = _vio_duration__
        self.ratio = ratio
        self.starttime = None
        self.started = False
        self.stopped = False
        self.publish_thread = Thread(target=self.publish)

    def start(self, starttime):
        self.started = True
        self.starttime = starttime
        self.publish_thread.start()

    def stop(self):
        self.stopped = True
        if self.started:
            self.publish_thread.join()
        self.out_queue.put(None)

    def publish(self):
        for data in self._vio_dataset__:
            if self.stopped:
                return
            interval = data[0] - self.dataset_starttime
            if interval < 0:
                continue
            while (time.time() - self.starttime) * self.ratio < interval + 1e-3:
                time.sleep(1e-3)        # assumption: data frequency < 1000 Hz
                if self.stopped:
                    return
            if interval <= self._vio_duration__ + 1e-3:
                self.out_queue.put(data)
            else:
                break
        self.out_queue.put(None)


# ------------------------------------------------------------------------------------------ synthetic export + evaluation
def _quat_wxyz_from_yaw(yaw):
    return np.array([np.cos(yaw / 2.0), 0.0, 0.0, np.sin(yaw / 2.0)])


def write_euroc(stream, path, t0_ns=1403636579000000000):
    """Writes a synth.Stream as an EuRoC sequence directory: mav0/imu0/data.csv, mav0/state_groundtruth_estimate0/data.csv
    (truth of synth.truth_position / truth_yaw at the IMU rate) and mav0/features0/data.csv (the stereo tracks).
    Timestamps are integer nanoseconds from ``t0_ns`` like the dataset's; values are written with repr precision."""
    from . import synth
    st = stream.structure
    mav = os.path.join(path, 'mav0')
    for sub in ('imu0', 'state_groundtruth_estimate0', 'features0'):
        os.makedirs(os.path.join(mav, sub), exist_ok=True)

    def ns(t):
        return t0_ns + int(round(t * 1e9))

    with open(os.path.join(mav, 'imu0', 'data.csv'), 'w') as f:
        f.write('#timestamp [ns],w_RS_S_x [rad s^-1],w_RS_S_y [rad s^-1],w_RS_S_z [rad s^-1],a_RS_S_x [m s^-2],a_RS_S_y [m s^-2],a_RS_S_z [m s^-2]\n')
        for k, t in enumerate(stream.imu_ts):
            f.write('%d,%s\n' % (ns(t), ','.join(repr(float(x)) for x in list(stream.gyro[k]) + list(stream.acc[k]))))
    speed = getattr(st, 'speed', 1.0)
    with open(os.path.join(mav, 'state_groundtruth_estimate0', 'data.csv'), 'w') as f:
        f.write('#timestamp,p_RS_R_x [m],p_RS_R_y [m],p_RS_R_z [m],q_RS_w [],q_RS_x [],q_RS_y [],q_RS_z [],v_RS_R_x [m s^-1],'
                'v_RS_R_y [m s^-1],v_RS_R_z [m s^-1],b_w_RS_S_x [rad s^-1],b_w_RS_S_y [rad s^-1],b_w_RS_S_z [rad s^-1],'
                'b_a_RS_S_x [m s^-2],b_a_RS_S_y [m s^-2],b_a_RS_S_z [m s^-2]\n')
        h = 1e-4
        for t in stream.imu_ts:
            p = synth.truth_position(t, speed)
            v = (synth.truth_position(t + h, speed) - synth.truth_position(t - h, speed)) / (2 * h)
            q = _quat_wxyz_from_yaw(float(synth.truth_yaw(t, speed)))
            row = list(p) + list(q) + list(v) + list(st.gyro_bias) + list(st.acc_bias)
            f.write('%d,%s\n' % (ns(t), ','.join(repr(float(x)) for x in row)))
    with open(os.path.join(mav, 'features0', 'data.csv'), 'w') as f:
        f.write('#timestamp [ns],id,u0,v0,u1,v1\n')
        for fr in range(stream.n_frames):
            t = ns(stream.frame_ts[fr])
            for i, uv in zip(stream.frame_ids[fr], stream.frame_uv[fr]):
                f.write('%d,%d,%s\n' % (t, int(i), ','.join(repr(float(x)) for x in uv)))
    return t0_ns


def absolute_trajectory_error(est_t, est_p, groundtruth, t_offset=0.0):
    """RMSE [m] between estimated positions ``est_p`` [n][3] at times ``est_t`` [n] (dataset clock, seconds) and the ground
    truth interpolated at those times, after removing the ground-truth position at the first estimate (the filter's world
    frame is the IMU pose at the first image).  ``groundtruth``: a GroundTruthReader or any iterable of gt messages."""
    gt = [(m.vio_timestamp__, m.vio_p__) for m in groundtruth]
    gt_t = np.array([g[0] for g in gt]) + t_offset
    gt_p = np.array([g[1] for g in gt])
    est_t, est_p = np.asarray(est_t, dtype=np.float64), np.asarray(est_p, dtype=np.float64)
    ref = np.stack([np.interp(est_t, gt_t, gt_p[:, k]) for k in range(3)], axis=1)
    err = (est_p - est_p[0]) - (ref - ref[0])
    return float(np.sqrt(np.mean(np.sum(err * err, axis=1))))
