"""tools/chain_stress.py — repeat the cfg2 step (no graph) to flush out intermittent protocol failures of the chain kernels."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from find_b200.step import TrainStep
from test_step_gpu import _setup

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
B, V, P2 = 8, 10242, 10000
dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, B, seed=52)
step = TrainStep(model, B, P2, with_colour=True, use_graph=False)
step.load_inputs(torch.arange(B), target, colt)
ref = None
try:
    for i in range(n):
        step.run()
        if i % 20 == 19:
            torch.cuda.synchronize()
            g = step.flat_grad.clone()
            if ref is None:
                ref = g
            elif not torch.equal(ref, g):
                print("iteration", i, "flat_grad differs: rel", float((g - ref).norm() / ref.norm()))
    torch.cuda.synchronize()
    print("ok", n, "steps; loss", step.loss3.tolist())
except Exception as e:
    print("FAILED at iteration", i, str(e)[:200])
