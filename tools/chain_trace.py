"""tools/chain_trace.py — print the timeline of CTA 0's first head_fwd item (findb200_chain_trace)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib, synthetic as syn
from find_b200.field import NeuralDisplacementField

dev = torch.device("cuda:0")
lib = _lib.load()
B, V = 8, 10242
model = NeuralDisplacementField(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=B, val_size=1,
                                shapevec_size=100, texvec_size=100, posevec_size=100, device=dev)
tv, tf = syn.make_template(V)
model.set_template(tv, tf)
lat = syn.make_latents(B, 0)
a = {k: lat[k].to(dev).requires_grad_(True) for k in ("shapevec", "posevec", "texvec", "reg")}
call = lambda: model.get_meshes(shapevec=a["shapevec"], reg=a["reg"], texvec=a["texvec"], posevec=a["posevec"])
for _ in range(3):
    call()
buf = torch.zeros(256, dtype=torch.int64, device=dev)
lib.findb200_chain_trace(_lib.ptr(buf))
call()
torch.cuda.synchronize()
lib.findb200_chain_trace(None)
t = buf.cpu().tolist()
t0 = min(x for x in t if x > 0)
r = lambda i: (t[i] - t0) / 1e3 if t[i] else float("nan")
print("MMA warp (us): per (layer, tile): ready-to-issue, issued")
for l in range(4):
    for tt in range(2):
        print(f"  L{l} T{tt}: {r((l*2+tt)*2):7.2f} {r((l*2+tt)*2+1):7.2f}")
print("epilogue leader (us): acc_ready, overwrite_ok, epi_done, published")
for tl in range(2):
    for l in range(3):
        b = 32 + (tl * 3 + l) * 4
        print(f"  T{tl} L{l}: {r(b):7.2f} {r(b+1):7.2f} {r(b+2):7.2f} {r(b+3):7.2f}")
print("producer chunk issue (us):", " ".join(f"{r(64+i):.2f}" for i in range(32)))
print("mma chunk full   (us):", " ".join(f"{r(96+i):.2f}" for i in range(32)))
print("kernel start/end (us): trunk_fwd", r(200), r(201), " head_fwd", r(202), r(203))
print("trunk_fwd MMA warp (us): [enc pass 0..2, layers 1..4] ready, issued")
for i in range(7):
    print(f"  stage {i}: {r(128+i*2):7.2f} {r(128+i*2+1):7.2f}")
print("trunk_fwd epilogue leader: encode pass start / signalled")
for p_ in range(3):
    print(f"  enc {p_}: {r(160+p_*2):7.2f} {r(160+p_*2+1):7.2f}")
print("trunk_fwd epilogue leader per layer: loop top, acc_ready, published")
for l in range(5):
    print(f"  L{l}: {r(170+l*3):7.2f} {r(170+l*3+1):7.2f} {r(170+l*3+2):7.2f}")
