#!/usr/bin/env python
"""A/B of library knobs on one box: runs the c3 (and optionally c5) bench legs several times in one process, switching
environment variables the launchers read at every launch.  usage: python tools/ab_env.py "VAR=val,VAR2=val" "..." [--c5]"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402

import bench  # noqa: E402
from msckf_b200 import capi  # noqa: E402

combos = [a for a in sys.argv[1:] if not a.startswith('--')]
with_c5 = '--c5' in sys.argv
peaks = capi.fp64_peak(0)
for rep in range(2):
    for combo in combos:
        keys = []
        for kv in [x for x in combo.split(',') if x and x != '-']:
            k, v = kv.split('=')
            os.environ[k] = v
            keys.append(k)
        c3 = bench.c3_leg(0, False, peaks)
        line = 'rep %d  %-40s c3 %.3f ms (qr %.3f)' % (rep, combo, c3['ms_per_frame'], c3['families']['stacked_qr']['ms'])
        if with_c5:
            c5 = bench.c5_leg(0, 32, False, peaks, torch)
            line += '   c5 %.3f ms/frame (qr %.3f)' % (c5['ms_per_frame'], c5['families']['stacked_qr']['ms'] / 10)
        print(line, flush=True)
        for k in keys:
            del os.environ[k]
