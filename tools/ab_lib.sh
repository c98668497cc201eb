#!/bin/bash
# A/B of two builds of the library on one box: tools/ab_lib.sh csrc/libA.so csrc/libB.so [rounds]  (alternates A B A B ...)
PKG=visual-inertial-odometry-msckf-stereo_b200
R=${3:-2}
for r in $(seq 1 $R); do
  for v in "$1" "$2"; do
    cp "$PKG/$v" $PKG/libmsckf_b200.so
    python bench.py --steps 20 --warmup 5 --no-latency --no-cpu-baseline --no-c3 --no-c5 --no-parity-spot --no-strong $BENCH_ARGS > /tmp/ab.json 2>/tmp/ab.err || tail -3 /tmp/ab.err
    python - "$v" <<'PY'
import json, sys
d = json.load(open('/tmp/ab.json'))
print('%-14s value %.1f M/s (mean %.1f, min %.1f, max %.1f)  e2e %.1f' % (sys.argv[1], d['value'] / 1e6, d['blocks']['value_mean'] / 1e6, d['blocks']['value_min'] / 1e6, d['blocks']['value_max'] / 1e6, d['e2e']['value'] / 1e6))
PY
  done
done
