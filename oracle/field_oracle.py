"""oracle/field_oracle.py — TEST INFRASTRUCTURE ONLY (never imported by find_b200/).

Plain-torch CPU restatement (any dtype: fp32 as the reference computes, fp64 for gradient checks) of
FIND's neural displacement field and the registration transform that follows it:

  * Fourier features        /root/reference/src/utils/fourier_feature_transform.py:17-55
  * field forward           /root/reference/src/model/model.py:393-453
  * layer sizes             /root/reference/src/model/model.py:224-259,350-371
  * registration transform  /root/reference/src/model/model.py:481-491  (+ PyTorch3D
                            ``euler_angles_to_matrix(.,'XYZ')`` = Rx·Ry·Rz and the row-vector
                            ``Transform3d().scale(S).rotate(R).translate(t)``:  X = ((v+off)*S) @ R + t)
  * 3-D loss wiring         /root/reference/src/model/losses.py:59-90 (plain branch, chamfer on points)

Pinned: ``tests/test_oracle_field.py`` checks this file against ``tests/golden/field_golden.npz``,
which ``oracle/make_golden.py`` produced by running the reference's OWN ``NeuralDisplacementField``
(imported under stubs by ``oracle/ref_field.py``) in the build container.
The registration transform and Chamfer parts live in PyTorch3D (absent) => PARITY UNPINNED for those.
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import numpy as np
import torch

TRUNK = ["base.0", "base.2", "base.4", "base.6", "base.8"]
DISP = ["mlp_disp.0", "mlp_disp.2", "mlp_disp.4", "mlp_disp.6"]
COL = ["mlp_col.0", "mlp_col.2", "mlp_col.4", "mlp_col.6"]
WEIGHT_KEYS = [f"{l}.{s}" for l in TRUNK + DISP + COL for s in ("weight", "bias")]


def fourier_B(num_input_channels: int = 3, mapping_size: int = 256, scale: float = 10.0) -> torch.Tensor:
    """fourier_feature_transform.py:23-26 — NOTE: reseeds the global torch RNG exactly as the reference."""
    torch.manual_seed(1)
    B = torch.randn((num_input_channels, mapping_size)) * scale
    rows = sorted(B, key=lambda r: torch.norm(r, p=2))      # iterating [3,256] yields the 3 rows
    return torch.stack(rows)


def fourier_encode(x: torch.Tensor, B: torch.Tensor) -> torch.Tensor:
    """fourier_feature_transform.py:28-55: [x, sin(2*pi*xB), cos(2*pi*xB)] on the last dim."""
    shp = x.shape
    x2 = x.reshape(-1, shp[-1])
    res = x2 @ B.to(x2.dtype)
    res = 2 * np.pi * res
    out = torch.cat([x2, torch.sin(res), torch.cos(res)], dim=1)
    return out.reshape(*shp[:-1], out.shape[-1])


def _q(x: torch.Tensor, operand_dtype) -> torch.Tensor:
    """Round a GEMM operand to `operand_dtype` with a straight-through gradient (identity when None).

    operand_dtype=torch.float16 restates what the sm_100a kernels feed the tensor cores (fp16 operands, fp32
    accumulate; csrc/field_chain.cu): the "rounding-matched" oracle used to check the backward kernels exactly,
    because ReLU masks of a reduced-precision forward legitimately differ from the fp32 forward's near z = 0."""
    if operand_dtype is None:
        return x
    return x + (x.to(operand_dtype).to(x.dtype) - x).detach()


def _lin(x, p, n, od):
    return torch.nn.functional.linear(_q(x, od), _q(p[n + ".weight"], od), p[n + ".bias"])


def _head(x, lat, p, names, od):
    """mlp_disp / mlp_col on cat([x, lat]) (model.py:428-440); the latent columns stay fp32 in the kernels."""
    W = p[names[0] + ".weight"]
    h = torch.nn.functional.linear(_q(x, od), _q(W[:, :x.shape[-1]], od), p[names[0] + ".bias"])
    if lat is not None:
        h = h + torch.nn.functional.linear(lat, W[:, x.shape[-1]:])
    h = torch.relu(h)
    for n in names[1:-1]:
        h = torch.relu(_lin(h, p, n, od))
    return _lin(h, p, names[-1], od)


def field_forward(p: Dict[str, torch.Tensor], B: torch.Tensor, pos: torch.Tensor,
                  shapevec: Optional[torch.Tensor] = None, texvec: Optional[torch.Tensor] = None,
                  posevec: Optional[torch.Tensor] = None, avg_col: Optional[torch.Tensor] = None,
                  operand_dtype=None):
    """model.py:393-453.  pos [B or 1, V, 3]; latents [B, 100] or None -> {'disp','col'} [B,V,3].
    operand_dtype=None is the reference arithmetic (fp32 / whatever dtype the inputs have); see _q."""
    batch, npts, _ = pos.shape
    if batch == 1 and shapevec is not None:
        batch = shapevec.shape[0]
        pos = pos.expand(batch, -1, -1)
    ex = lambda z: None if z is None else z.unsqueeze(1).expand(-1, npts, -1)
    shapevec, posevec, texvec = ex(shapevec), ex(posevec), ex(texvec)
    x = fourier_encode(pos, B)
    for n in TRUNK:                                         # every trunk Linear is followed by ReLU
        x = torch.relu(_lin(x, p, n, operand_dtype))
    lat_d = [z for z in (shapevec, posevec) if z is not None]    # disp_in = [x, shape, pose] (model.py:428-433)
    lat_d = torch.cat(lat_d, dim=-1) if lat_d else None
    disp = 0.1 * torch.tanh(_head(x, lat_d, p, DISP, operand_dtype))
    col = 0.5 * (1 + torch.tanh(_head(x, texvec, p, COL, operand_dtype)))   # col_in = [x, tex] (model.py:435-437)
    if avg_col is not None:
        col = avg_col[None, None, :] + col
    return {"disp": disp, "col": col}


def euler_xyz_to_matrix(e: torch.Tensor) -> torch.Tensor:
    """PyTorch3D euler_angles_to_matrix(e, 'XYZ') = Rx(e0) @ Ry(e1) @ Rz(e2), e [...,3] -> [...,3,3]."""
    a, b, c = e.unbind(-1)
    one, zero = torch.ones_like(a), torch.zeros_like(a)
    ca, sa, cb, sb, cc, sc = torch.cos(a), torch.sin(a), torch.cos(b), torch.sin(b), torch.cos(c), torch.sin(c)
    Rx = torch.stack([one, zero, zero, zero, ca, -sa, zero, sa, ca], -1).reshape(e.shape[:-1] + (3, 3))
    Ry = torch.stack([cb, zero, sb, zero, one, zero, -sb, zero, cb], -1).reshape(e.shape[:-1] + (3, 3))
    Rz = torch.stack([cc, -sc, zero, sc, cc, zero, zero, zero, one], -1).reshape(e.shape[:-1] + (3, 3))
    return Rx @ Ry @ Rz


def reg_transform(points: torch.Tensor, reg: Optional[torch.Tensor]) -> torch.Tensor:
    """model.py:481-491: reg = [t(3), euler(3), S(3)];  X = (points * S) @ R + t  (row vectors)."""
    if reg is None:
        return points
    t, e, S = reg[..., :3], reg[..., 3:6], reg[..., 6:9]
    R = euler_xyz_to_matrix(e)
    return torch.matmul(points * S[:, None, :], R) + t[:, None, :]


def get_verts(p, B, template_verts, shapevec=None, reg=None, texvec=None, posevec=None, avg_col=None):
    """model.py:455-504 minus the Meshes packaging -> dict(offsets, verts, disp, col)."""
    n = 0 if shapevec is None else shapevec.shape[0]
    verts = template_verts.expand(n, -1, -1)
    res = field_forward(p, B, verts, shapevec=shapevec, texvec=texvec, posevec=posevec, avg_col=avg_col)
    X = reg_transform(verts + res["disp"], reg)
    return dict(offsets=res["disp"], verts=X, **res)


def chamfer_torch(x: torch.Tensor, y: torch.Tensor, x_lengths=None, y_lengths=None, chunk: int = 2048):
    """Differentiable brute-force chamfer_distance(x, y) (mean/mean) in plain torch, for gradient checks.
    Distances are the direct difference form; ties resolve to the lowest index (torch.min on CPU)."""
    N, P1, _ = x.shape
    P2 = y.shape[1]
    xl = torch.full((N,), P1, dtype=torch.long) if x_lengths is None else x_lengths
    yl = torch.full((N,), P2, dtype=torch.long) if y_lengths is None else y_lengths

    def one_dir(a, b, la, lb):
        tot = a.new_zeros(())
        for n in range(a.shape[0]):
            an, bn = a[n, : int(la[n])], b[n, : int(lb[n])]
            if an.shape[0] == 0 or bn.shape[0] == 0:
                continue
            s = a.new_zeros(())
            for s0 in range(0, an.shape[0], chunk):
                d = (an[s0:s0 + chunk, None, :] - bn[None, :, :]).square().sum(-1)
                s = s + d.min(dim=1).values.sum()
            tot = tot + s / max(int(la[n]), 1)
        return tot / max(N, 1)

    return one_dir(x, y, xl, yl) + one_dir(y, x, yl, xl)


# ---------------------------------------------------------------------------------------------
# Synthetic workload — shared by tests, smoke and the bench's CPU leg.
def random_params(seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """nn.Linear-style init for all 13 layers + a NON-zero last disp layer (the reference zero-inits it,
    model.py:516-518, which makes every upstream gradient vanish)."""
    g = torch.Generator().manual_seed(seed)
    shapes = {"base.0": (256, 515), "base.2": (256, 256), "base.4": (256, 256), "base.6": (256, 256),
              "base.8": (256, 256), "mlp_disp.0": (256, 456), "mlp_disp.2": (256, 256),
              "mlp_disp.4": (256, 256), "mlp_disp.6": (3, 256), "mlp_col.0": (256, 356),
              "mlp_col.2": (256, 256), "mlp_col.4": (256, 256), "mlp_col.6": (3, 256)}
    p = {}
    for n, (o, i) in shapes.items():
        k = 1.0 / math.sqrt(i)
        p[n + ".weight"] = ((torch.rand(o, i, generator=g) * 2 - 1) * k).to(dtype)
        p[n + ".bias"] = ((torch.rand(o, generator=g) * 2 - 1) * k).to(dtype)
    p["mlp_disp.6.weight"] = (torch.randn(3, 256, generator=g) * 0.05).to(dtype)
    return p


def kinkfree_params(seed: int, Bmat: torch.Tensor, pos: torch.Tensor, shapevec, posevec, texvec,
                    sigma: float = 0.2, margin: float = 6.0) -> Dict[str, torch.Tensor]:
    """Parameters for which NO ReLU pre-activation of the given inputs is near zero (every unit sits `margin`
    standard deviations on one side), so that a reduced-precision forward has exactly the masks of the fp32 forward
    and gradients can be compared at rounding-error level.  (With ordinary weights ~1e-4 of the pre-activations
    change sign under fp16/tf32/bf16 operand rounding and each flip changes a gradient term by 100 %.)"""
    p = random_params(seed)
    g = torch.Generator().manual_seed(seed + 1)
    npts = pos.shape[1]
    nb = shapevec.shape[0]
    ex = lambda z: z.unsqueeze(1).expand(-1, npts, -1)
    x = fourier_encode(pos.expand(nb, -1, -1), Bmat)

    def fix(name, inp):
        W = p[name + ".weight"]
        z = torch.nn.functional.linear(inp, W)
        s = z.reshape(-1, z.shape[-1]).std(0).clamp_min(1e-6)
        W = W * (sigma / s)[:, None]
        z = torch.nn.functional.linear(inp, W)
        mu = z.reshape(-1, z.shape[-1]).mean(0)
        sign = (torch.rand(W.shape[0], generator=g) < 0.5).float() * 2 - 1
        p[name + ".weight"] = W.contiguous()
        p[name + ".bias"] = (sign * margin * sigma - mu).contiguous()
        return torch.relu(z + p[name + ".bias"])

    for n in TRUNK:
        x = fix(n, x)
    for names, lat in ((DISP, torch.cat([ex(shapevec), ex(posevec)], -1)), (COL, ex(texvec))):
        h = fix(names[0], torch.cat([x, lat], -1))
        for n in names[1:-1]:
            h = fix(n, h)
        W = p[names[-1] + ".weight"]
        z = torch.nn.functional.linear(h, W)
        p[names[-1] + ".weight"] = (W * (0.5 / z.std().clamp_min(1e-6))).contiguous()
        p[names[-1] + ".bias"] = (-torch.nn.functional.linear(h, p[names[-1] + ".weight"]).reshape(-1, 3).mean(0)).contiguous()
    return p
