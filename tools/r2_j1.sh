# NVLS allreduce: isolated timing against the peer-memory kernel and NCCL, correctness, bench at this N
N=${1:-2}
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29731 tools/p2p_bench.py 2>gpurun_out/j1_p2p.err | grep "^{" | tee gpurun_out/j1_p2p_n$N.json; tail -3 gpurun_out/j1_p2p.err | grep -v "OMP_NUM\|^\*\*\*"
FINDB200_COLLECTIVE=nvls timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29732 tools/dp_check.py 2>gpurun_out/j1_dp.err | grep "^{" | tee gpurun_out/j1_dp_nvls_n$N.json; tail -3 gpurun_out/j1_dp.err | grep -v "OMP_NUM\|^\*\*\*"
for c in p2p nvls; do
FINDB200_COLLECTIVE=$c timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29733 bench.py --gpus $N --steps 200 --warmup 5 --skip-cpu-baseline --skip-eager --skip-other-workloads > gpurun_out/j1_bench_${c}_n$N.json 2> gpurun_out/j1_bench.err; echo "$c rc=$?"
python - <<PY
import json
for ln in open('gpurun_out/j1_bench_${c}_n$N.json'):
    if ln.startswith('{'):
        d=json.loads(ln); print(' ', d["n_gpus"], d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"], d["config"]["collective"], [(k['name'], k['ms_per_step']) for k in d['kernels'] if k['name'].startswith('allreduce')])
PY
done
