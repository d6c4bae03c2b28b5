# weak-scaling check of the default workload at N = 2, 4, 8 (one rank per GPU over NCCL), as the driver launches it
for n in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 100 --warmup 5 > gpurun_out/sc_n$n.json 2> gpurun_out/sc_n$n.err; echo n$n=$?; tail -1 gpurun_out/sc_n$n.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 8 --workload cfg3 --steps 50 --warmup 5 > gpurun_out/sc_cfg3_n8.json 2> gpurun_out/sc_cfg3_n8.err; echo cfg3=$?
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29601 bench.py --gpus 8 --workload cfg4 --steps 30 --warmup 5 > gpurun_out/sc_cfg4_n8.json 2> gpurun_out/sc_cfg4_n8.err; echo cfg4=$?
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29602 bench.py --gpus 8 --workload cfg5 --steps 30 --warmup 5 > gpurun_out/sc_cfg5_n8.json 2> gpurun_out/sc_cfg5_n8.err; echo cfg5=$?
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29603 tools/dp_check.py > gpurun_out/dp_check_n8.json 2> gpurun_out/dp_check_n8.err; echo dp8=$?; cat gpurun_out/dp_check_n8.json
timeout 120 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/sc_ref.json 2>/dev/null; echo ref=$?
