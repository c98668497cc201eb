python bench.py --steps 60 --no-latency --no-cpu-baseline 2>/dev/null | python -c "
import sys,json
for x in sys.stdin:
    if x.startswith('{'):
        j=json.loads(x); print(round(j['value']/1e6,2), round(j['ms_per_step'],4), round(j['e2e']['value']/1e6,2), j['clocks']['sm_mhz']); print({k: round(v['ms']/20,3) for k,v in j['roofline']['families'].items()})
        print(' '.join('%s=%.0f' % (k.replace('k_',''), v['ms']/20*1e3) for k,v in sorted(j['roofline']['all_kernels'].items(), key=lambda kv:-kv[1]['ms'])))"
