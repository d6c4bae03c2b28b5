"""tools/umma_rate.py — cycles per K = 16 tcgen05.mma of the chain kernels' shapes (findb200_umma_rate)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib
lib = _lib.load()
dev = torch.device("cuda:0")
names = {0: "cg1 M128 N256 SS", 1: "cg1 M128 N128 SS", 2: "cg2 M256 N256 SS", 3: "cg1 M128 N256 TS", 4: "cg2 M256 N256 TS"}
out = {}
src = torch.zeros(1 << 20, dtype=torch.uint8, device=dev)
buf = torch.zeros(2, dtype=torch.int64, device=dev)


def run(mode, iters, flags):
    _lib.check(lib.findb200_umma_rate(mode, iters, flags, _lib.ptr(src), _lib.ptr(buf), _lib.stream_ptr(dev)), "umma_rate")
    torch.cuda.synchronize()
    cyc, n = buf.tolist()
    return round(cyc / n, 1)


for mode in range(5):
    out[names[mode]] = run(mode, 64, 0)
fl = {1: "commit/4", 2: "fence/4", 4: "wait/4", 8: "bg bulk copies", 16: "bg st.shared"}
for flags in (1, 3, 7, 8, 16, 24, 31):
    out["cg2 SS + " + " + ".join(v for k, v in fl.items() if flags & k)] = run(2, 64, flags)
for flags in (7, 31):
    out["cg2 TS + " + " + ".join(v for k, v in fl.items() if flags & k)] = run(4, 64, flags)
print(json.dumps(out, indent=1))
