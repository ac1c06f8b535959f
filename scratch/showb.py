import json, sys
for f in sys.argv[1:]:
    for line in open(f):
        if not line.startswith('{'): continue
        d = json.loads(line)
        if d.get('impl') == 'reference':
            print(f, 'REF', round(d['value']), d['ms_per_step']); continue
        print('==', f, d['config']['workload'], 'ms', round(d['ms_per_step'], 3), 'e2e', round(d['e2e']['ms_per_step'], 3),
              'bit_exact', d.get('bit_exact'), 'ver', d.get('verified'), 'launches', d.get('gpu_launches'))
        print('   H acc ms', round(d['roofline']['ms_per_launch'], 3), 'frac', round(d['roofline']['frac'], 3), 'step frac', round(d['roofline_step']['frac'], 3),
              'ntt alone', round(d['roofline_ntt']['ms_alone'], 3), 'in-step', round(d['roofline_ntt']['ms_in_step'], 3), 'ntt frac', round(d['roofline_ntt']['frac'], 3))
        print('   msm', {k: (v['additions'], v['c'], v['W'], round(v['accumulate_ms'], 2)) for k, v in d['msm'].items()})
        for k, v in d['trace_last_step_ms'].items():
            print('   ', k, v)
        print('   phases', {k: round(v, 2) for k, v in d['phases_ms'].items()})
        for pc in d.get('per_config', []) or []:
            print('   PC', pc['config']['workload'], pc.get('ms_per_step'), pc.get('e2e', {}).get('ms_per_step'), pc.get('bit_exact'), pc.get('roofline_step', {}).get('frac'))
        if 'strong' in d: print('   STRONG', json.dumps(d['strong'])[:600])
