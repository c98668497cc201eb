"""Attribute the warp-stall samples of an ncu SASS source page to CUDA source lines.

`ncu -i rep --page source --csv -k regex:<kernel>` lists one row per SASS instruction (samples, executed count) but no
source line; `nvdisasm -g` of the cubin of the SAME build lists the instructions with `//## File ..., line N` markers.
Both walk the function in address order, so instruction k of one is instruction k of the other; the opcodes are compared
and the tool stops if the two listings are not the same binary.

usage: python tools/sass_hotspots.py <source_page.csv> <object.o> <mangled-name-substring> [top N]
e.g.   python tools/sass_hotspots.py gpurun_out/src_qr.csv visual-inertial-odometry-msckf-stereo_b200/csrc/k_qr_dmma.o k_qr_dmma 25
"""
import csv
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict


def opcode(text):
    t = text.replace('{', ' ').replace('}', ' ').split()
    if not t:
        return ''
    o = t[1] if t[0].startswith('@') and len(t) > 1 else t[0]
    return o.rstrip(';')


def disasm_lines(obj, name):
    """[(opcode, file, line)] of the function whose mangled name contains `name`."""
    tmp = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(obj)], cwd=tmp, check=True, capture_output=True)
    out = []
    for cubin in sorted(os.listdir(tmp)):
        txt = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
        inside, cur = False, (None, 0)
        for ln in txt.splitlines():
            m = re.match(r'\s*\.section\s+\.text\.(\S+?),', ln)
            if m:
                inside = name in m.group(1)
                continue
            if not inside:
                continue
            m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
            if m:
                cur = (m.group(1), int(m.group(2)))
                continue
            m = re.match(r'\s*/\*[0-9a-f]{4,}\*/\s+(.*?);', ln)
            if m:
                out.append((opcode(m.group(1)), cur[0], cur[1]))
        if out:
            break
    return out


def main():
    page, obj, name = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
    rows = list(csv.reader(open(page)))
    # the page holds one table per profiled launch of the kernel: take the launch with the most samples
    heads = [i for i, r in enumerate(rows) if 'Source' in r and '# Samples' in r]
    best = []
    for n, hi in enumerate(heads):
        h = rows[hi]
        si, ci, ei = h.index('Source'), h.index('# Samples'), h.index('Instructions Executed')
        end = heads[n + 1] if n + 1 < len(heads) else len(rows)
        inst = [(opcode(r[si]), int(r[ci] or 0), int(r[ei] or 0)) for r in rows[hi + 1:end]
                if len(r) > max(si, ci, ei) and re.match(r'^[0-9a-fx]+$', r[0] or '')]
        if sum(x[1] for x in inst) > sum(x[1] for x in best):
            best = inst
    sass = best
    dis = disasm_lines(obj, name)
    if len(dis) < len(sass):
        sys.exit('nvdisasm lists %d instructions, the page %d: not the same build' % (len(dis), len(sass)))
    bad = sum(1 for (a, _, _), (b, _, _) in zip(sass, dis) if a.split('.')[0] != b.split('.')[0])
    if bad > len(sass) // 100:
        sys.exit('%d of %d opcodes differ: not the same build' % (bad, len(sass)))
    per = defaultdict(lambda: [0, 0, 0])
    for (op, smp, ex), (_, f, l) in zip(sass, dis):
        p = per[(f, l)]
        p[0] += smp; p[1] += ex; p[2] += 1
    tot = sum(v[0] for v in per.values()) or 1
    src = {}
    print('%d SASS instructions, %d samples' % (len(sass), tot))
    for (f, l), v in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
        if f and f not in src and os.path.exists(f):
            src[f] = open(f).read().splitlines()
        text = src.get(f, [])[l - 1].strip() if f in src and 0 < l <= len(src[f]) else ''
        print('%5.1f %%  %9d inst  %3d sass  %s:%d  %s' % (100.0 * v[0] / tot, v[1], v[2], os.path.basename(f or '?'), l, text[:110]))


if __name__ == '__main__':
    main()
