"""tools/chain_probe.py — head_fwd / trunk_fwd time with and without saving activations (is the chain or the store the limiter?)"""
import ctypes, os, sys
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
args = {k: lat[k].to(dev) for k in ("shapevec", "posevec", "texvec", "reg")}
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def run(grad):
    for k in args:
        args[k].requires_grad_(grad)
    ctx = torch.enable_grad() if grad else torch.no_grad()
    with ctx:
        for _ in range(3):
            model.get_meshes(shapevec=args["shapevec"], reg=args["reg"], texvec=args["texvec"], posevec=args["posevec"])
        torch.cuda.synchronize()
        lib.findb200_profile_enable(1)
        n = 10
        for _ in range(n):
            flush.zero_()
            model.get_meshes(shapevec=args["shapevec"], reg=args["reg"], texvec=args["texvec"], posevec=args["posevec"])
        torch.cuda.synchronize()
        buf = ctypes.create_string_buffer(1 << 14)
        lib.findb200_profile_read(buf, len(buf))
        lib.findb200_profile_enable(0)
    out = {}
    for l in buf.value.decode().strip().splitlines():
        name, cnt, tot = l.split(",")
        out[name] = round(float(tot) / n * 1e3, 1)
    return out

print("save=0", run(False))
print("save=1", run(True))
