#!/bin/bash
# Headline bench under different environment knobs of the library, alternating, on one box:
#   tools/ab_knobs.sh rounds "A=1 B=2" "A=3" ...     ("-" = no knob)
R=$1; shift
for r in $(seq 1 $R); do
  for combo in "$@"; do
    if [ "$combo" = "-" ]; then envs=""; else envs="$combo"; fi
    env $envs python bench.py --steps 20 --warmup 5 --no-latency --no-cpu-baseline --no-c3 --no-c5 --no-parity-spot --no-strong $BENCH_ARGS > /tmp/abk.json 2>/tmp/abk.err || tail -3 /tmp/abk.err
    python - "$combo" <<'PY'
import json, sys
d = json.load(open('/tmp/abk.json'))
print('%-44s value %.1f M/s (mean %.1f, min %.1f, max %.1f)  e2e %.1f' % (sys.argv[1], d['value'] / 1e6, d['blocks']['value_mean'] / 1e6, d['blocks']['value_min'] / 1e6, d['blocks']['value_max'] / 1e6, d['e2e']['value'] / 1e6))
PY
  done
done
