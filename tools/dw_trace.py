"""tools/dw_trace.py — per-CTA lifetimes of the dW launch (findb200_chain_trace buffer, csrc/field_dw.cu) by GEMM."""
import ctypes, os, re, sys, subprocess, tempfile
os.environ["FINDB200_CHAIN_DBG"] = "512"
# the library prints its static allocation on stderr ("dw gemm ..." / "dw cta ..."): keep it in a file to join with the lifetimes
_errf = tempfile.TemporaryFile(mode="w+")
_saved = os.dup(2)
os.dup2(_errf.fileno(), 2)
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

os.dup2(_saved, 2)
_errf.seek(0)
alloc = {}
for line in _errf.read().splitlines():
    m = re.match(r"dw cta (\d+): gemm (\d+) units (\d+) (\d+) cost (\d+) early (\d+)", line)
    if m:
        b, g, u0, u1, c, e = map(int, m.groups())
        alloc.setdefault(b, [])
        if (g, u0, u1, c, e) not in alloc[b]:
            alloc[b].append((g, u0, u1, c, e))
per = {}
for b, segs in alloc.items():
    if len(segs) != 1 or b >= n:
        continue
    g, u0, u1, c, e = segs[0]
    per.setdefault((g, e), []).append(((en[b] - st[b]) / 1e3 / (u1 - u0), c / 1e3, (en[b] - t0) / 1e3))
print("single-GEMM CTAs: (gemm, early) -> n, measured us/unit, model us/unit, mean end us")
for k in sorted(per):
    v = per[k]
    print("  gemm %2d early %d: n %3d  measured %.3f  model %.3f  end %.1f" % (k[0], k[1], len(v), sum(x[0] for x in v) / len(v), v[0][1], sum(x[2] for x in v) / len(v)))
print("latest CTAs: cta end_us start_us [(gemm, units, early)]")
for b in sorted(range(n), key=lambda i: -en[i])[:24]:
    print("  %3d  %.1f  %.1f  %s" % (b, (en[b] - t0) / 1e3, (st[b] - t0) / 1e3, [(g, u1 - u0, e) for g, u0, u1, c, e in alloc.get(b, [])]))
print("earliest CTAs:")
for b in sorted(range(n), key=lambda i: en[i])[:10]:
    print("  %3d  %.1f  %.1f  %s" % (b, (en[b] - t0) / 1e3, (st[b] - t0) / 1e3, [(g, u1 - u0, e) for g, u0, u1, c, e in alloc.get(b, [])]))
print("cloud_sort phases (us from kernel start):", [round((t[400 + k] - t[400]) / 1e3, 2) for k in range(8)])
