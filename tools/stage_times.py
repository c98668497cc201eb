"""Scratch: per-stage device times (msckf_last_timing) of steady-state frames, host-buffer path, F filters."""
import sys

import numpy as np

from _stream import make_batch, host_copy

F = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
NF = 130
b, st, frames = make_batch(F, NF)
ctx = b.ctx
for f in range(90):
    ctx.feature_callback(*host_copy(frames[f]))
ctx.set_timing(True)
rows = []
for f in range(90, NF):
    ctx.feature_callback(*host_copy(frames[f]))
    rows.append(ctx.last_timing())
rows = np.array(rows)
names = ['propagate+augment(+observe)', '-', 'remove_lost', 'prune(+finish)', 'total']
pr = rows[:, 3] > 0.2
for tag, sel in (('pruning frames', pr), ('other frames', ~pr), ('all', np.ones(len(rows), bool))):
    print(tag, int(sel.sum()), ' '.join('%s=%.3f' % (n, v) for n, v in zip(names, rows[sel].mean(axis=0))))
b.close()
