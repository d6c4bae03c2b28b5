# round 2: 8-GPU sanity — the 2-rank -m gpu test, then the bench at N = 8 and N = 4
timeout 300 python -m pytest tests/test_parallel.py -m gpu -x -q 2>&1 | tail -3
for n in 8 4; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 30 --warmup 3 --skip-cpu-baseline --skip-eager > gpurun_out/n${n}_bench.json 2> gpurun_out/n${n}_bench.err; echo rc$n=$?
python - <<PY
import json
for ln in open('gpurun_out/n${n}_bench.json'):
    if ln.startswith('{'):
        d=json.loads(ln); print(d["n_gpus"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["config"]["collective"], {k:(v.get("ms_per_step"), v.get("value")) for k,v in d.get("other_workloads",{}).items()})
PY
done
