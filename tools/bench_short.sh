python bench.py --steps 60 --no-latency --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for x in sys.stdin:
    if x.startswith('{'):
        j=json.loads(x); print(round(j['value']/1e6,2), round(j['ms_per_step'],4), round(j['e2e']['value']/1e6,2), j['clocks']['sm_mhz']); print({k: round(v['ms']/20,3) for k,v in j['roofline']['families'].items()}, 'prop', round(j['roofline']['all_kernels']['k_propagate']['ms']/20,3))"
