"""Scratch: run the same batch with the medium fused update on / off and report the first divergence."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
from msckf_b200 import capi, synth
from msckf_b200.config import ConfigEuRoC
from msckf_b200.msckf import MSCKFBatch

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
NF = int(sys.argv[2]) if len(sys.argv) > 2 else 90
cfg = ConfigEuRoC()
cfg._vio_max_cam_state_size__ = 20
st = synth.make_structure(NF, 300, structure_seed=0)
rng = np.random.default_rng(1000)
gyro, acc, uvs = bench.build_frames(st, rng, F, NF)

def make(env):
    for k in ('MSCKF_NO_MED', 'MSCKF_NO_SMALL'):
        os.environ.pop(k, None)
    if env:
        os.environ[env] = '1'
    b = MSCKFBatch(cfg, F, device=0, feature_capacity=1024)
    b.initialize_gravity_and_bias(gyro[:, :200], acc[:, :200])
    return b

ba, bb = make(None), make('MSCKF_NO_MED')
STOP = int(sys.argv[3]) if len(sys.argv) > 3 else 75
WF = int(sys.argv[4]) if len(sys.argv) > 4 else 589
def cmpP(tag):
    Pa, Pb = ba.ctx.get_cov(WF), bb.ctx.get_cov(WF)
    ev = np.linalg.eigvalsh((Pa + Pa.T) / 2)
    print(tag, 'n', Pa.shape[0], 'relP', np.linalg.norm(Pa - Pb) / np.linalg.norm(Pb), 'min eig A', ev[0], 'B', np.linalg.eigvalsh(Pb)[0],
          'asym A', np.abs(Pa - Pa.T).max())
for f in range(STOP + 1):
    fa = bench.frame_arrays(ba, st, gyro, acc, uvs, f)
    fb = bench.frame_arrays(bb, st, gyro, acc, uvs, f)
    if f < STOP:
        ba.ctx.feature_callback(*[fa[k] for k in bench.KEYS])
        bb.ctx.feature_callback(*[fb[k] for k in bench.KEYS])
        continue
    cmpP('before frame')
    for b, fr in ((ba, fa), (bb, fb)):
        b.ctx.propagate(fr['imu_off'], fr['ts'], fr['gyro'], fr['acc'])
        b.ctx.augment(fr['cam_ids'])
        b.ctx.add_observations(fr['feat_off'], fr['ids'], fr['uv'])
    cmpP('after observe')
    for b in (ba, bb):
        b.ctx.remove_lost_features()
        b.ctx.sync()
    cmpP('after lost')
    for name, b in (('A', ba), ('B', bb)):
        l = b.ctx.stage_log(WF, 0)
        st_ = np.array(l['status'])
        print(name, 'lost stage rows/cols', l['update_rows'], l['update_cols'], 'n', len(st_), 'pass', int((st_ == 3).sum()), 'dof', [int(x) for x, s_ in zip(l['dof'], st_) if s_ == 3])
    # P of all filters
    worst = 0
    for w in range(0, F, 1):
        Pa, Pb = ba.ctx.get_cov(w), bb.ctx.get_cov(w)
        e = np.linalg.norm(Pa - Pb) / np.linalg.norm(Pb)
        if e > 1e-9:
            la = ba.ctx.stage_log(w, 0)
            print('filter', w, 'relP', e, 'rows', la['update_rows'], 'dofmax', max([int(x) for x, s_ in zip(la['dof'], la['status']) if s_ == 3] + [0]))

    for w in range(F):
        ia, ib = ba.ctx.get_imu_state(w), bb.ctx.get_imu_state(w)
        ca, cb = ba.ctx.get_cam_states(w), bb.ctx.get_cam_states(w)
        e1 = np.abs(ia - ib).max()
        e2 = np.abs(np.asarray(ca[1]) - np.asarray(cb[1])).max()
        if e1 > 1e-9 or e2 > 1e-9 or not np.isfinite(e1 + e2):
            print('filter', w, 'imu diff', e1, 'cam diff', e2)
    for name, b in (('A', ba), ('B', bb)):
        try:
            b.ctx.prune_cam_states()
            b.ctx.sync()
            print(name, 'prune ok')
        except Exception as e:
            print(name, 'prune failed', e)
        l = b.ctx.stage_log(WF, 1)
        st_ = np.array(l['status'])
        print(name, 'prune rows/cols', l['update_rows'], l['update_cols'], 'n', len(st_), 'pass', int((st_ == 3).sum()), 'gamma[:8]', np.array(l['gamma'])[:8])
    cmpP('after prune')
