"""Turn the two ncu exports of a round into the committed artefacts under profiles/:
   launch list (gpu__time_duration per launch)  -> per-kernel table (markdown)
   --set full raw page (csv)                    -> condensed per-launch lines + dram bytes per launch per kernel family (json)
usage: python tools/ncu_profiles.py <launches.csv> <raw.csv> <tag> [traffic json name, e.g. traffic_r2.json]"""
import csv, json, os, re, subprocess, sys
from collections import defaultdict
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
launches, raw, tag = sys.argv[1:4]
traffic_name = sys.argv[4] if len(sys.argv) > 4 else None

FAMILY = [('triangulate', r'k_triangulate'), ('jacobian_nullspace_gate', r'k_jacobian'), ('stacked_qr', r'k_qr'),
          ('gain_update', r'k_upd|k_wupd'), ('propagate', r'k_propagate')]


def family(name):
    for fam, pat in FAMILY:
        if re.search(pat, name):
            return fam
    return None


def short(name):
    name = re.sub(r'^void ', '', name)
    name = re.sub(r'\(.*$', '', name)
    return re.sub(r'\(int\)', '', name.replace('msckf::', ''))


rows = [r for r in csv.reader(l for l in open(launches) if l.startswith('"'))]
hdr, rows = rows[0], rows[1:]
ci = hdr.index('Kernel Name'); vi = hdr.index('Metric Value')
agg = defaultdict(lambda: [0, 0.0])
for r in rows:
    k = short(r[ci])
    agg[k][0] += 1
    agg[k][1] += float(r[vi].replace(',', '')) / 1e3
tot = sum(v[1] for k, v in agg.items() if k.startswith('k_'))
out = ['| kernel | launches | total us | avg us | share of our kernels |', '|---|---|---|---|---|']
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if not k.startswith('k_'):
        continue
    out.append('| %s | %d | %.1f | %.1f | %.3f |' % (k, v[0], v[1], v[1] / v[0], v[1] / tot))
open(os.path.join(REPO, 'profiles', tag + '_launches.md'), 'w').write('\n'.join(out) + '\n')

summ = subprocess.run([sys.executable, os.path.join(REPO, 'tools', 'ncu_summary.py'), raw], capture_output=True, text=True).stdout
open(os.path.join(REPO, 'profiles', tag + '_ncu_full_summary.txt'), 'w').write(summ)

rr = list(csv.reader(open(raw)))
h, data = rr[0], rr[2:]
kn, rd, wr, du = h.index('Kernel Name'), h.index('dram__bytes_read.sum'), h.index('dram__bytes_write.sum'), h.index('gpu__time_duration.sum')
units = rr[1]
def to_bytes(v, unit):
    v = float(v.replace(',', ''))
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(unit, 1)
fam = defaultdict(lambda: [0, 0.0, 0.0])
for r in data:
    f = family(r[kn])
    if f is None:
        continue
    fam[f][0] += 1
    fam[f][1] += to_bytes(r[rd], units[rd]) + to_bytes(r[wr], units[wr])
    fam[f][2] += float(r[du].replace(',', ''))
js = {'source': 'ncu --set full --clock-control none, %s (bench.py default workload, 2 frames)' % os.path.basename(raw),
      'dram_bytes_per_launch': dict((f, v[1] / v[0]) for f, v in fam.items()),
      'profiled_launches': dict((f, v[0]) for f, v in fam.items())}
if traffic_name:
    json.dump(js, open(os.path.join(REPO, 'profiles', traffic_name), 'w'), indent=1)
print('\n'.join(out[:14]))
print(json.dumps(js, indent=1))
