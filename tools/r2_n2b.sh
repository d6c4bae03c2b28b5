# round 2, final build (dW split + presort): 2-rank -m gpu test, dp_check and the bench at N = 2
timeout 300 python -m pytest tests/test_parallel.py -m gpu -x -q 2>&1 | tail -3
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 tools/dp_check.py > gpurun_out/dp_check.json 2> gpurun_out/dp_check.err; echo dp=$?; cat gpurun_out/dp_check.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/dp_check.err | tail -5
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 5 --skip-cpu-baseline --skip-eager > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo rc2=$?; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/n2_bench.err | tail -5
python - <<PY
import json
for ln in open('gpurun_out/n2_bench.json'):
    if ln.startswith('{'):
        d=json.loads(ln); print(d["n_gpus"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["config"]["collective"], {k:(v.get("ms_per_step"), v.get("value")) for k,v in d.get("other_workloads",{}).items()})
        print([(k['name'], k['ms_per_step']) for k in d['kernels']])
PY
