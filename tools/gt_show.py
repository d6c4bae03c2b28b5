import json, sys
d = json.loads(open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/gt_bench.json').read().strip().splitlines()[-1])
print({k: d.get(k) for k in ('value', 'ms_per_step')}, 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
print('roofline', d.get('roofline'))
print(' '.join('%s=%.1f' % (k['name'], 1e3 * k['ms_per_step']) for k in d['kernels']))
