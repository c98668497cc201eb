#!/usr/bin/env python
"""Prints the c3 / c5 blocks (and the headline numbers, when present) of a bench.py JSON line in readable form."""
import json
import sys

d = json.load(open(sys.argv[1]))
if 'value' in d:
    print('value %.1f M/s  ms/step %.4f  e2e %.1f M/s  launches %d  clocks %s' % (d['value'] / 1e6, d['ms_per_step'], d['e2e']['value'] / 1e6,
                                                                                d['gpu_launches'], d.get('clocks')))
    print('blocks', d.get('blocks'))
    print('parity_spot', d.get('parity_spot'), 'not_spd', d.get('gate_not_spd'), 'ate', d.get('ate', {}).get('rmse_m'))
    r = d['roofline']
    print('dominant', r['kernel'], 'frac %.4f' % r['frac'])
    for k, v in r['families'].items():
        print('  %-24s ms %.3f share %.3f tflops %.3f frac %.4f' % (k, v['ms'], v['share'], v['tflops'], v['frac']),
              ('exec_frac %.4f' % v['executed_frac']) if 'executed_frac' in v else '')
    for k, v in sorted(r['all_kernels'].items(), key=lambda x: -x[1]['ms']):
        print('     %-18s %.3f ms  %d launches  %.1f us/launch' % (k, v['ms'], v['launches'], 1e3 * v['ms'] / max(v['launches'], 1)))
for key in ('c3', 'c5'):
    if key not in d:
        continue
    c = d[key]
    print('=====', key)
    for k, v in c.items():
        if k == 'families':
            for kk, vv in v.items():
                print('   %-24s ms %.3f tflops %.3f frac %.4f' % (kk, vv['ms'], vv['tflops'], vv['frac']),
                      ('exec_tflops %.3f exec_frac %.4f' % (vv['executed_tflops'], vv['executed_frac'])) if 'executed_frac' in vv else '')
        elif k != 'workload':
            print('  ', k, v)
