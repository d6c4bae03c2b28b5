"""oracle/make_golden.py — generates tests/golden/field_golden.npz by running the reference's OWN
NeuralDisplacementField (model.py:206-453, imported by oracle/ref_field.py) on CPU with torch autograd.

Run in the build container only (needs /root/reference):   python oracle/make_golden.py
The produced fixture travels to the GPU box; this script and /root/reference do not need to.
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_field  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                   "field_golden.npz")


def state_hash(sd) -> str:
    h = hashlib.sha256()
    for k in sorted(sd):
        h.update(k.encode())
        h.update(sd[k].detach().cpu().contiguous().numpy().tobytes())
    return h.hexdigest()


def grad_summary(g: torch.Tensor):
    g64 = g.double()
    return np.array([g64.sum().item(), g64.norm().item()], np.float64), g[:4].detach().numpy().copy()


def main():
    torch.set_num_threads(1)          # deterministic reduction order for the fixture
    model = ref_field.make_reference_field(train_size=8, val_size=2)
    out = {}
    B = model.encoder[0]._B
    out["B"] = B.numpy().copy()
    out["B_sha256"] = np.array(hashlib.sha256(B.numpy().tobytes()).hexdigest())
    out["ctor_state_sha256"] = np.array(state_hash(model.state_dict()))

    # SURVEY.md §8(d) parameter recipe: non-degenerate last disp layer, all under manual_seed(0)
    torch.manual_seed(0)
    with torch.no_grad():
        model.mlp_disp[-1].weight.normal_(0, 0.05)
        model.mlp_disp[-1].bias.normal_(0, 0.05)
    for k, v in model.state_dict().items():
        if k.startswith(("base.", "mlp_disp.", "mlp_col.")):
            out["w/" + k] = v.detach().numpy().copy()

    def run_case(tag, pos, zs, zt, zp, gd, gc):
        zs, zt, zp = [z.clone().requires_grad_(True) for z in (zs, zt, zp)]
        model.zero_grad()
        res = model(pos, shapevec=zs, texvec=zt, posevec=zp)
        loss = (res["disp"] * gd).sum() + (res["col"] * gc).sum()
        loss.backward()
        out[f"{tag}/pos"] = pos.numpy().copy()
        for n, z in (("shapevec", zs), ("texvec", zt), ("posevec", zp)):
            out[f"{tag}/{n}"] = z.detach().numpy().copy()
            out[f"{tag}/grad_{n}"] = z.grad.numpy().copy()
        out[f"{tag}/gd"], out[f"{tag}/gc"] = gd.numpy().copy(), gc.numpy().copy()
        out[f"{tag}/disp"] = res["disp"].detach().numpy().copy()
        out[f"{tag}/col"] = res["col"].detach().numpy().copy()
        for k, prm in model.named_parameters():
            if not k.startswith(("base.", "mlp_disp.", "mlp_col.")):
                continue
            if prm.grad.numel() <= 1024:
                out[f"{tag}/grad/{k}"] = prm.grad.numpy().copy()
            else:
                s, head = grad_summary(prm.grad)
                out[f"{tag}/gradsum/{k}"] = s
                out[f"{tag}/gradhead/{k}"] = head

    g = torch.Generator().manual_seed(1234)
    rn = lambda *s, sc=1.0: torch.randn(*s, generator=g) * sc
    # case A: template shared by all feet (get_meshes call site, model.py:473-477)
    run_case("A", rn(1, 200, 3, sc=0.05), rn(3, 100, sc=0.1), rn(3, 100, sc=0.1), rn(3, 100, sc=0.1),
             rn(3, 200, 3), rn(3, 200, 3))
    # case B: per-foot distinct points (TextureLossGTSpace call site, losses.py:50)
    run_case("B", rn(2, 130, 3, sc=0.05), rn(2, 100, sc=0.1), rn(2, 100, sc=0.1), rn(2, 100, sc=0.1),
             rn(2, 130, 3), rn(2, 130, 3))

    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes;", len(out), "arrays; torch", torch.__version__)


if __name__ == "__main__":
    main()
