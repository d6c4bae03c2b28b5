"""oracle/step_oracle.py — TEST INFRASTRUCTURE ONLY (never imported by find_b200/).

CPU restatement of FIND's 3-D training step on the hot path (chamf + colour configuration):
    sample_latent_vectors            /root/reference/src/train/trainer.py:29-46, 97
    ModelWithLoss.forward            /root/reference/src/model/model.py:1009, 1026-1030, 1157-1158
      get_meshes                     /root/reference/src/model/model.py:455-504   (oracle/field_oracle.py)
      DisplacementLoss, plain branch /root/reference/src/model/losses.py:87-88    (oracle/chamfer_oracle.c)
    loss.backward()                  /root/reference/src/train/trainer.py:121     (torch CPU autograd)
Chamfer is taken between the deformed template vertices and the target cloud (BASELINE.json configs), the colour
term is a per-vertex MSE (stand-in for TextureLossGTSpace, losses.py:22-57).  Field part pinned by
tests/golden/field_golden.npz (reference's own class); Chamfer / reg-transform parts: PARITY UNPINNED (PyTorch3D
absent) — see the headers of field_oracle.py / chamfer_oracle.c.

Also the timed CPU baseline of bench.py (kind "port"): torch CPU ops use all host threads, the C Chamfer uses OpenMP.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np
import torch

from . import chamfer_oracle as co
from . import field_oracle as fo
from . import mesh_oracle as mo


def train_step(p: Dict[str, torch.Tensor], Bmat: torch.Tensor, template: torch.Tensor, tables: Dict[str, torch.Tensor],
               idx: torch.Tensor, target: torch.Tensor, col_target: Optional[torch.Tensor] = None,
               w_chamf: float = 1e4, w_tex: float = 1.0, avg_col: Optional[torch.Tensor] = None,
               threads: int = 1, need_param_grads: bool = True, forward_only: bool = False,
               faces: Optional[np.ndarray] = None, num_samples: int = 0, seed: int = 0, counter: int = 0,
               w_smooth: float = 0.0, draw_from_verts: Optional[np.ndarray] = None) -> Dict[str, object]:
    """p: the 26 MLP tensors (field_oracle.WEIGHT_KEYS); tables: 'shapevec','posevec','texvec','reg' [N,·];
    idx: int64 [B] rows of this batch; template [V,3]; target [B,P2,3]; col_target [B,V,3] or None.

    Mesh terms (cfgs/train_3d.yaml: chamf + smooth + tex): ``faces`` [F,3] template topology; ``num_samples`` > 0 takes
    Chamfer on surface samples of the deformed template (losses.py:67; draw = mesh_oracle with ``seed`` / ``counter``,
    made from ``draw_from_verts`` when given so a test can reuse the GPU's own draw); ``w_smooth`` adds
    w_smooth * MeshSmoothnessLoss (model.py:1032-1033, losses.py:93-99).

    -> dict(loss, chamf, tex, smooth, verts, disp, col, idx_x, idx_y, grads={param: grad}, table_grads={name: dense grad})"""
    pp = {k: v.detach().clone().requires_grad_(need_param_grads and not forward_only) for k, v in p.items()}
    tt = {k: v.detach().clone().requires_grad_(not forward_only) for k, v in tables.items()}
    rows = {k: v[idx] for k, v in tt.items()}                       # LatentVector.__getitem__ (model.py:130-152)
    res = fo.get_verts(pp, Bmat, template[None], shapevec=rows.get("shapevec"), reg=rows.get("reg"),
                       texvec=rows.get("texvec"), posevec=rows.get("posevec"), avg_col=avg_col)
    verts = res["verts"]
    pred = verts
    if num_samples > 0:
        dv = verts.detach().numpy() if draw_from_verts is None else np.asarray(draw_from_verts, dtype=np.float32)
        d = mo.sample_points(dv, faces, num_samples, seed=seed, counter=counter)
        f = torch.as_tensor(np.asarray(faces).astype(np.int64))
        fi, w = torch.as_tensor(d["face_idx"]), torch.as_tensor(d["bary"])
        corner = [torch.stack([verts[n][f[fi[n], k]] for n in range(verts.shape[0])]) for k in range(3)]
        pred = (w[..., 0:1] * corner[0] + w[..., 1:2] * corner[1]) + w[..., 2:3] * corner[2]
    ch = co.chamfer(pred.detach().numpy(), target.numpy(), threads=threads, with_grad=not forward_only)
    chamf = float(ch["loss"])
    tex = 0.0
    out = dict(chamf=chamf, verts=verts.detach(), disp=res["disp"].detach(), col=res["col"].detach(),
               idx_x=ch["idx_x"], idx_y=ch["idx_y"])
    mse = None
    if col_target is not None and w_tex:
        mse = torch.nn.functional.mse_loss(res["col"], col_target)
        tex = float(mse.detach())
    out["tex"] = tex
    smooth_t = None
    if w_smooth:
        smooth_t, e, l = mo.mesh_smoothness_loss(verts, faces)
        out["smooth"], out["smooth_edge"], out["smooth_lap"] = float(smooth_t.detach()), float(e.detach()), float(l.detach())
    out["loss"] = w_chamf * chamf + w_tex * tex + (w_smooth * out["smooth"] if w_smooth else 0.0)
    if forward_only:
        return out
    # d(w_chamf * chamfer)/d(verts) comes from the C oracle; the rest is torch autograd
    gx = torch.from_numpy(np.ascontiguousarray(ch["grad_x"])) * w_chamf
    surrogate = (pred * gx).sum()
    if smooth_t is not None:
        surrogate = surrogate + w_smooth * smooth_t
    if mse is not None:
        surrogate = surrogate + w_tex * mse
    surrogate.backward()
    out["grads"] = {k: v.grad for k, v in pp.items()} if need_param_grads else {}
    out["table_grads"] = {k: v.grad for k, v in tt.items()}
    return out
