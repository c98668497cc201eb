#!/bin/bash
# headline numbers only (no CPU legs, no configs[2] / configs[4] legs), printed in readable form
python bench.py --steps 20 --warmup 5 --no-latency --no-cpu-baseline --no-c3 --no-c5 --no-parity-spot > /tmp/bench_short.json && python tools/show_legs.py /tmp/bench_short.json
