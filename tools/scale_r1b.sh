# final weak-scaling verification with the peer-memory allreduce as the default collective (N = 4, 8; cfg5 at 8; dp_check at 8)
run() { n=$1; tag=$2; shift 2; timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+RANDOM%400)) "$@" > gpurun_out/$tag.json 2> gpurun_out/$tag.err; echo $tag=$?; }
run 8 dp_check_n8 tools/dp_check.py; cat gpurun_out/dp_check_n8.json
run 8 sc_n8 bench.py --gpus 8 --steps 100 --warmup 5
run 4 sc_n4 bench.py --gpus 4 --steps 100 --warmup 5
run 8 sc_cfg5_n8 bench.py --gpus 8 --workload cfg5 --steps 30 --warmup 5
run 8 sc_cfg3_n8 bench.py --gpus 8 --workload cfg3 --steps 30 --warmup 5
FINDB200_COLLECTIVE=nccl run 8 sc_n8_nccl bench.py --gpus 8 --steps 100 --warmup 5 --skip-cpu-baseline
for f in sc_n8 sc_n4 sc_cfg5_n8 sc_cfg3_n8 sc_n8_nccl; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]); print('$f', round(d['value']/1e6,1), round(d['ms_per_step'],4), d['config'].get('collective'))
except Exception as e: print('$f', 'ERR', e)
PY
done
