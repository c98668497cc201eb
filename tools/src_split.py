#!/usr/bin/env python
"""Extracts the source-page tables of the launches of one kernel from a (gzipped) `ncu --page source --csv` export that
holds many kernels (tools/ncu_box.sh).  usage: python tools/src_split.py <export.csv[.gz]> <kernel-name-substring> <out.csv>"""
import gzip
import sys

path, pat, out = sys.argv[1:4]
op = gzip.open if path.endswith('.gz') else open
keep = False
n = 0
with op(path, 'rt') as f, open(out, 'w') as o:
    for line in f:
        if line.startswith('"Kernel Name"'):
            keep = pat in line
            n += keep
        if keep:
            o.write(line)
print('%d launches of %s' % (n, pat))
