"""Scratch: how the difference between the CUDA filters and the CPU oracle evolves over a long run.  Runs the headline stream
(device generator, noise seed = filter index) for F filters, downloads state and covariance of a few filters every STEP frames
and compares them with the oracle fed with the same messages.
usage: python tools/drift_probe.py F n_frames step f0,f1,..."""
import multiprocessing as mp
import os
import sys

import numpy as np

from _stream import make_batch, KEYS
import bench
from msckf_b200 import capi


def oracle_run(job):
    os.environ.setdefault('OPENBLAS_NUM_THREADS', '1')
    from collections import namedtuple
    from oracle import msckf_oracle as orc
    from msckf_b200 import synth
    max_cam, init_ts, init_g, init_a, frames, step = job
    cfg = orc.ConfigEuRoC()
    cfg._vio_max_cam_state_size__ = max_cam
    filt = orc.MSCKF(cfg)
    for t, g, a in zip(init_ts, init_g, init_a):
        filt.imu_callback(synth.ImuMsg(float(t), g.copy(), a.copy()))
    PackedMsg = namedtuple('PackedMsg', ['timestamp', 'vio_features'])
    out = []
    for fi, (t_frame, ts, g, a, ids, uv) in enumerate(frames):
        for k in range(len(ts)):
            filt.imu_callback(synth.ImuMsg(float(ts[k]), g[k].copy(), a[k].copy()))
        feats = [synth.FeatureMeasurement(int(ids[i]), float(uv[i, 0]), float(uv[i, 1]), float(uv[i, 2]), float(uv[i, 3])) for i in range(len(ids))]
        filt.feature_callback(PackedMsg(float(t_frame), feats))
        if fi % step == step - 1:
            ss = filt.state_server
            imu = ss._vio_imu_state__
            out.append((fi, np.concatenate([imu.orientation, imu._vio_gyro_bias__, imu.velocity, imu.acc_bias, imu._vio_position__]),
                        np.array(ss._vio_state_cov__)))
    return out


if __name__ == '__main__':
    F, NF, STEP = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    watch = [int(x) for x in sys.argv[4].split(',')]
    b, st, frames = make_batch(F, NF)
    ctx = b.ctx
    g0, a0 = bench.DeviceStream(st, b, F, 0, 0, __import__('torch')).static_samples()
    per = dict((s, []) for s in watch)
    gpu = dict((s, []) for s in watch)
    for f in range(NF):
        d = frames[f]
        ts = d['ts'].reshape(F, -1)[0].cpu().numpy()
        ids = d['ids'].reshape(F, -1)[0].cpu().numpy()
        gy, ac, uv = d['gyro'][watch].cpu().numpy(), d['acc'][watch].cpu().numpy(), d['uv'][watch].cpu().numpy()
        for j, s in enumerate(watch):
            per[s].append((float(st.frame_ts[f]), ts, gy[j], ac[j], ids, uv[j]))
        ctx.feature_callback_dev(*[d[k].data_ptr() for k in KEYS])
        if f % STEP == STEP - 1:
            ctx.sync()
            for s in watch:
                blk = ctx.get_imu_state(s)
                gpu[s].append((f, np.concatenate([blk[capi.IMU_Q:capi.IMU_Q + 4], blk[capi.IMU_BG:capi.IMU_BG + 3], blk[capi.IMU_V:capi.IMU_V + 3],
                                                  blk[capi.IMU_BA:capi.IMU_BA + 3], blk[capi.IMU_P:capi.IMU_P + 3]]), ctx.get_cov(s)))
    print('not_spd per watched filter', ctx.counters()[watch, capi.CNT_GATE_NOT_SPD], flush=True)
    b.close()
    jobs = [(20, st.imu_ts[:200], g0[s], a0[s], per[s], STEP) for s in watch]
    with mp.get_context('spawn').Pool(len(jobs)) as pool:
        outs = pool.map(oracle_run, jobs)
    for s, o in zip(watch, outs):
        line = []
        for (f, gi, gP), (fo, oi, oP) in zip(gpu[s], o):
            e = max(np.linalg.norm(gi - oi) / np.linalg.norm(oi), np.linalg.norm(gP - oP) / np.linalg.norm(oP)) if gP.shape == oP.shape else float('inf')
            line.append('%d:%.1e' % (f + 1, e))
        print('filter %d  ' % s + ' '.join(line), flush=True)
