"""tools/dw_trace.py — per-CTA lifetimes of the dW launch (findb200_chain_trace buffer, csrc/field_dw.cu) by GEMM."""
import ctypes, os, re, sys, subprocess
os.environ["FINDB200_CHAIN_DBG"] = "512"
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from find_b200 import _lib
from find_b200.step import TrainStep
from test_step_gpu import _setup

lib = _lib.load()
B, V, P2 = 8, 10242, 10000
dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, B, seed=52)
step = TrainStep(model, B, P2, with_colour=True, use_graph=False)
step.load_inputs(torch.arange(B), target, colt)
for _ in range(3):
    step.run()
torch.cuda.synchronize()
buf = torch.zeros(512, dtype=torch.int64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
flush.zero_()
lib.findb200_chain_trace(ctypes.c_void_p(buf.data_ptr() | 1))
step.run()
torch.cuda.synchronize()
lib.findb200_chain_trace(None)
t = buf.cpu().tolist()
n = 148
st = [t[2 * i] for i in range(n)]
en = [t[2 * i + 1] for i in range(n)]
t0 = min(x for x in st if x > 0)
print("CTA start spread %.1f us, kernel %.1f us" % ((max(st) - t0) / 1e3, (max(en) - t0) / 1e3))
print("per CTA (start, end) us:")
for i in range(0, n, 8):
    print("  " + "  ".join(f"{i+k:3d}:{(st[i+k]-t0)/1e3:5.1f}-{(en[i+k]-t0)/1e3:5.1f}" for k in range(8) if i + k < n))

print("cloud_sort phases (us from kernel start):", [round((t[400 + k] - t[400]) / 1e3, 2) for k in range(8)])
