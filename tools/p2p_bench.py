"""tools/p2p_bench.py — findb200_allreduce_p2p vs NCCL all_reduce on the 868 358-float bucket (run under torchrun)."""
import json, os, sys
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200.parallel import P2PAllReduce

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", lr)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
n = 868358
res = {}
for ctas in (32, 64, 128, 148):
    ar = P2PAllReduce(n, dev, n_ctas=ctas)
    ar.bucket.fill_(rank + 1.0)
    ar.enqueue(1.0)
    torch.cuda.synchronize()
    ok = bool(torch.all(ar.bucket == world * (world + 1) / 2).item())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            ar.enqueue(1.0 / world)
    g.replay(); torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    res[f"p2p_ctas{ctas}_us"] = e0.elapsed_time(e1) / 200 * 1e3
    res[f"p2p_ctas{ctas}_ok"] = ok
# the NVLS variant (sum inside the switch), when the ranks share a multicast object
for ctas in (32, 64, 128, 148):
    ar = P2PAllReduce(n, dev, n_ctas=ctas)
    mc = torch.tensor([1 if ar.mc_ptr else 0], device=dev)
    dist.all_reduce(mc, op=dist.ReduceOp.MIN)
    if int(mc.item()) == 0:
        res["nvls"] = "no multicast mapping"
        break
    ar.use_mc = True
    ar.bucket.fill_(rank + 1.0)
    ar.enqueue(1.0)
    torch.cuda.synchronize()
    ok = bool(torch.all(ar.bucket == world * (world + 1) / 2).item())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            ar.enqueue(1.0 / world)
    g.replay(); torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    res[f"nvls_ctas{ctas}_us"] = e0.elapsed_time(e1) / 200 * 1e3
    res[f"nvls_ctas{ctas}_ok"] = ok
x = torch.ones(n, device=dev)
dist.all_reduce(x); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(20):
        dist.all_reduce(x, op=dist.ReduceOp.AVG)
g.replay(); torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    g.replay()
e1.record(); torch.cuda.synchronize()
res["nccl_us"] = e0.elapsed_time(e1) / 200 * 1e3
res["world"] = world
if rank == 0:
    print(json.dumps(res), flush=True)
dist.barrier(); torch.cuda.synchronize()
os._exit(0)
