"""Scratch: per-stage device times (msckf_last_timing) of steady-state frames, host-buffer path, F filters."""
import os, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench
from msckf_b200 import capi, synth
from msckf_b200.config import ConfigEuRoC
from msckf_b200.msckf import MSCKFBatch
F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
NF = 130
cfg = ConfigEuRoC(); cfg._vio_max_cam_state_size__ = 20
st = synth.make_structure(NF, 300, structure_seed=0)
gyro, acc, uvs = bench.build_frames(st, np.random.default_rng(1000), F, NF)
b = MSCKFBatch(cfg, F, device=0, feature_capacity=1024)
b.initialize_gravity_and_bias(gyro[:, :200], acc[:, :200])
frames = [dict((k, capi.pinned_copy(v)) for k, v in bench.frame_arrays(b, st, gyro, acc, uvs, f).items()) for f in range(NF)]
ctx = b.ctx
for f in range(90):
    ctx.feature_callback(*[frames[f][k] for k in bench.KEYS])
ctx.set_timing(True)
rows = []
for f in range(90, NF):
    ctx.feature_callback(*[frames[f][k] for k in bench.KEYS])
    rows.append(ctx.last_timing())
rows = np.array(rows)
names = ['propagate+augment', 'observe', 'remove_lost', 'prune(+finish)', 'total']
pr = rows[:, 3] > 0.2
for tag, sel in (('pruning frames', pr), ('other frames', ~pr), ('all', np.ones(len(rows), bool))):
    print(tag, int(sel.sum()), ' '.join('%s=%.3f' % (n, v) for n, v in zip(names, rows[sel].mean(axis=0))))
b.close()
