"""tools/chain2_trace.py — timeline of cluster 0's first head_fwd item on the CTA-pair chain kernels (findb200_chain_trace,
stamp slots in csrc/field_chain2.cu).  Times in us from the first stamp."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib, synthetic as syn
from find_b200.field import NeuralDisplacementField

dev = torch.device("cuda:0")
lib = _lib.load()
B, V = int(os.environ.get('TRACE_B', '8')), int(os.environ.get('TRACE_V', '10242'))
model = NeuralDisplacementField(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=B, val_size=1,
                                shapevec_size=100, texvec_size=100, posevec_size=100, device=dev)
tv, tf = syn.make_template(V)
model.set_template(tv, tf)
lat = syn.make_latents(B, 0)
a = {k: lat[k].to(dev).requires_grad_(True) for k in ("shapevec", "posevec", "texvec", "reg")}
call = lambda: model.get_meshes(shapevec=a["shapevec"], reg=a["reg"], texvec=a["texvec"], posevec=a["posevec"])
for _ in range(3):
    call()
import ctypes
buf = torch.zeros(512, dtype=torch.int64, device=dev)
# pass 1: cycle accounting of the issuing thread only (no per-chunk stamps)
lib.findb200_chain_trace(ctypes.c_void_p(buf.data_ptr() | 1))
call()
(a["shapevec"].sum() * 0).backward() if False else None
torch.cuda.synchronize()
acc = buf.cpu().tolist()
for mode, name in ((0, "trunk_fwd"), (1, "head_fwd")):
    o = acc[448 + mode * 8: 448 + mode * 8 + 7]
    if o[5]:
        print(f"{name}: issuer thread of cluster 0: total {o[0]} cyc, wait ready_cnt {o[1]}, issue+commit {o[3]}, "
              f"wait act_ready {o[4]}, chunks {o[5]}, items {o[6]}  -> per chunk: ready {o[1]/o[5]:.0f} issue {o[3]/o[5]:.0f}; "
              f"act wait per item {o[4]/max(o[6],1):.0f}; total per item {o[0]/max(o[6],1):.0f}")
if os.environ.get("TRACE_ACCOUNT_ONLY"):
    lib.findb200_chain_trace(None)
    sys.exit(0)
buf.zero_()
lib.findb200_chain_trace(_lib.ptr(buf))
call()
torch.cuda.synchronize()
lib.findb200_chain_trace(None)
t = buf.cpu().tolist()
t0 = t[502] if t[502] else min(x for x in t if x > 0)          # head_fwd kernel start
r = lambda i: (t[i] - t0) / 1e3 if t[i] else float("nan")
print("head_fwd kernel start/end (us):", r(502), r(503), "  trunk_fwd:", r(500), r(501))
print("leader MMA thread, per (layer, tile): operand ready, all chunks issued + committed")
for l in range(4):
    print("  " + "   ".join(f"L{l} T{j}: {r((l*2+j)*2):6.2f} {r((l*2+j)*2+1):6.2f}" for j in range(2)))
for who, base in (("leader", 16), ("peer", 272)):
    print(f"{who} epilogue (elected thread): acc_ready, overwrite_ok, epi_done, published")
    for l in range(3):
        print("  " + "   ".join(f"T{j} L{l}: " + " ".join(f"{r(base+(j*3+l)*4+k):6.2f}" for k in range(4)) for j in range(2)))
n = 32
print("leader producer, half-chunk issued :", " ".join(f"{r(64+i):.2f}" for i in range(n)))
print("peer   producer, half-chunk issued :", " ".join(f"{r(320+i):.2f}" for i in range(n)))
print("leader MMA, own half landed        :", " ".join(f"{r(112+i):.2f}" for i in range(n)))
print("peer relay, its half landed        :", " ".join(f"{r(368+i):.2f}" for i in range(n)))
print("leader scout, peer half known landed:", " ".join(f"{r(160+i):.2f}" for i in range(n)))
print("leader issuer, chunk event seen     :", " ".join(f"{r(208+i):.2f}" for i in range(n)))
