timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pdl_tests.log 2>&1; echo tests=$?; tail -3 gpurun_out/pdl_tests.log
for v in 1 0 1 0; do FINDB200_PDL=$v timeout 200 python bench.py --skip-cpu-baseline --steps 100 > gpurun_out/pdl_$v.json 2>gpurun_out/pdl_$v.err; echo -n "pdl=$v rc=$? "; python -c "
import json;d=json.loads(open('gpurun_out/pdl_$v.json').read().strip().splitlines()[-1]);print('step %.4f e2e %.4f shipped %.4f'%(d['ms_per_step'],d['e2e']['ms_per_step'],d['shipped_3d_step']['ms_per_iteration_median']))"; done
