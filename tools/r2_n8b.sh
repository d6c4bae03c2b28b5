# round 2, final build: bench at N = 8 and N = 4
for n in 8 4; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 100 --warmup 5 --skip-cpu-baseline --skip-eager > gpurun_out/n${n}_bench.json 2> gpurun_out/n${n}_bench.err; echo rc$n=$?
python - <<PY
import json
for ln in open('gpurun_out/n${n}_bench.json'):
    if ln.startswith('{'):
        d=json.loads(ln); print(d["n_gpus"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["config"]["collective"], {k:(v.get("ms_per_step"), v.get("value")) for k,v in d.get("other_workloads",{}).items()})
        print([(k['name'], k['ms_per_step']) for k in d['kernels'] if k['name'] in ('knn_main','allreduce_p2p','head_fwd','dw_gemms')])
PY
done
