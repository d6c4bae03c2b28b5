import json, sys
for l in open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/tk_knn.jsonl'):
    d = json.loads(l)
    print(d['B'], d['V'], d['P2'], 'mode', d['mode'], 'kq', d['kq'], 'ms %.3f' % d['ms_median'], 'exec/alg %.4f' % (d['pairs_exec'] / d['pairs_alg']),
          'slow', d['slow_queries'], 'max', d.get('max_evals'), 'groups', d.get('groups'), {k: round(v, 4) for k, v in d['kernels_ms'].items()}, d['loss'], d['idx_sum'])
