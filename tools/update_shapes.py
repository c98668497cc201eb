"""Scratch: histogram of the measurement-update shapes (m, n_a, k, kernel path) per stage at the headline workload,
plus per-stage device times.  usage: python tools/update_shapes.py [F]"""
import sys

import numpy as np

from _stream import make_batch, host_copy, KEYS

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
NF = 130
b, st, frames = make_batch(F, NF)
ctx = b.ctx
hist = {0: [], 1: []}
for f in range(NF):
    a = dict(zip(KEYS, host_copy(frames[f])))
    if f < 90:
        ctx.feature_callback(*[a[k] for k in KEYS])
        continue
    ctx.propagate(a['imu_off'], a['ts'], a['gyro'], a['acc'])
    ctx.augment(a['cam_ids'])
    ctx.add_observations(a['feat_off'], a['ids'], a['uv'])
    ctx.remove_lost_features()
    hist[0].append(ctx.update_shapes())
    ctx.prune_cam_states()
    hist[1].append(ctx.update_shapes())
for stage in (0, 1):
    h = np.concatenate(hist[stage])
    h = h[h[:, 3] > 0]
    print('stage', stage, 'updates per frame per filter %.3f' % (len(h) / (len(hist[stage]) * F)))
    for path in sorted(set(h[:, 3])):
        s = h[h[:, 3] == path]
        print('  path %d: share %.3f  m mean %.1f p50 %d p99 %d max %d | n_a mean %.1f p50 %d max %d | k mean %.1f max %d | compressed %.2f' % (
            path, len(s) / len(h), s[:, 0].mean(), np.median(s[:, 0]), np.percentile(s[:, 0], 99), s[:, 0].max(),
            s[:, 1].mean(), np.median(s[:, 1]), s[:, 1].max(), s[:, 2].mean(), s[:, 2].max(), (s[:, 0] > s[:, 1]).mean()))
    big = h[h[:, 3] == 2]
    if len(big):
        print('  tiled n_a histogram', np.bincount(big[:, 1] // 6)[10:])
b.close()
