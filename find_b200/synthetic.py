"""find_b200/synthetic.py — synthetic workload of SURVEY.md §8(d) (the Foot3D dataset is not shipped).

Template: icosphere (subdivision 5 -> 10 242 verts / 20 480 faces) or a (u,v) grid sphere for the 50 k-vertex
configuration, scaled to foot-like half-extents (0.125, 0.05, 0.05) m and centred (the reference centres its
template, /root/reference/src/model/model.py:274-275).  Targets: per-foot ellipsoid surface points with a radius
scale, a low-frequency bump and jitter.  Everything is seeded and generated on the CPU with numpy/torch.
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np
import torch

HALF_EXTENTS = (0.125, 0.05, 0.05)


def icosphere(subdiv: int) -> Tuple[np.ndarray, np.ndarray]:
    """Unit icosphere: verts [10*4^s+2, 3] float64, faces [20*4^s, 3] int64."""
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    verts = [np.array(p, np.float64) / np.linalg.norm(p) for p in v]
    faces = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
             (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
             (8, 6, 7), (9, 8, 1)]
    for _ in range(subdiv):
        cache: Dict[Tuple[int, int], int] = {}

        def mid(a, b):
            key = (a, b) if a < b else (b, a)
            if key not in cache:
                m = verts[a] + verts[b]
                verts.append(m / np.linalg.norm(m))
                cache[key] = len(verts) - 1
            return cache[key]

        nf = []
        for a, b, c in faces:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        faces = nf
    return np.stack(verts), np.array(faces, np.int64)


def grid_sphere(nu: int, nv: int) -> Tuple[np.ndarray, np.ndarray]:
    """(u,v) grid sphere with nu*nv vertices (poles excluded), quads split into triangles."""
    u = (np.arange(nu) + 0.5) / nu * math.pi
    v = np.arange(nv) / nv * 2 * math.pi
    uu, vv = np.meshgrid(u, v, indexing="ij")
    verts = np.stack([np.sin(uu) * np.cos(vv), np.sin(uu) * np.sin(vv), np.cos(uu)], -1).reshape(-1, 3)
    idx = np.arange(nu * nv).reshape(nu, nv)
    a, b = idx[:-1, :], idx[1:, :]
    a2, b2 = np.roll(a, -1, 1), np.roll(b, -1, 1)
    faces = np.concatenate([np.stack([a, b, a2], -1).reshape(-1, 3), np.stack([a2, b, b2], -1).reshape(-1, 3)])
    return verts, faces.astype(np.int64)


def make_template(V: int = 10242) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> verts [V,3] fp32 (foot-like ellipsoid, centroid at the origin), faces [F,3] int64."""
    if V == 10242:
        verts, faces = icosphere(5)
    elif V == 50000:
        verts, faces = grid_sphere(250, 200)
    else:
        s = 0
        while 10 * 4 ** s + 2 < V:
            s += 1
        verts, faces = icosphere(s)
        verts = verts[:V]
        faces = faces[(faces < V).all(1)]
    verts = verts * np.array(HALF_EXTENTS)[None]
    verts = verts - verts.mean(0, keepdims=True)
    return torch.from_numpy(verts.astype(np.float32)), torch.from_numpy(faces)


def make_targets(B: int, P: int, foot0: int = 0) -> torch.Tensor:
    """[B,P,3] fp32 target clouds; foot b uses torch.Generator(seed=1000 + foot0 + b), so a rank that owns feet
    [foot0, foot0+B) generates exactly the rows the single-rank job would."""
    out = torch.empty(B, P, 3, dtype=torch.float32)
    he = torch.tensor(HALF_EXTENTS, dtype=torch.float32)
    for b in range(B):
        g = torch.Generator().manual_seed(1000 + foot0 + b)
        d = torch.randn(P, 3, generator=g)
        d = d / d.norm(dim=1, keepdim=True).clamp_min(1e-12)
        scale = 1.0 + 0.1 * torch.randn(1, generator=g)
        phi = torch.rand(1, generator=g) * 2 * math.pi
        pts = d * he[None] * scale
        nrm = d / he[None]
        nrm = nrm / nrm.norm(dim=1, keepdim=True)
        pts = pts + nrm * (0.005 * torch.sin(20 * pts[:, :1] + phi))
        pts = pts + torch.randn(P, 3, generator=g) * 2e-4
        out[b] = pts
    return out


def make_latents(B: int, foot0: int = 0, size: int = 100) -> Dict[str, torch.Tensor]:
    """Per-foot latent rows / registration rows (§8(d): latents ~ N(0, 0.1^2); reg = [t ~ N(0,0.005^2),
    euler ~ N(0,0.05^2), S = 1 + N(0,0.02^2)]), seeded per foot."""
    zs, zp, zt, reg, colt = [], [], [], [], []
    for b in range(B):
        g = torch.Generator().manual_seed(5000 + foot0 + b)
        zs.append(torch.randn(size, generator=g) * 0.1)
        zp.append(torch.randn(size, generator=g) * 0.1)
        zt.append(torch.randn(size, generator=g) * 0.1)
        reg.append(torch.cat([torch.randn(3, generator=g) * 0.005, torch.randn(3, generator=g) * 0.05,
                              1 + torch.randn(3, generator=g) * 0.02]))
    return dict(shapevec=torch.stack(zs), posevec=torch.stack(zp), texvec=torch.stack(zt), reg=torch.stack(reg))


def make_col_target(B: int, V: int, foot0: int = 0) -> torch.Tensor:
    out = torch.empty(B, V, 3, dtype=torch.float32)
    for b in range(B):
        g = torch.Generator().manual_seed(9000 + foot0 + b)
        out[b] = torch.rand(V, 3, generator=g)
    return out


def randomise_last_disp_layer(mlp_disp_last: torch.nn.Linear, seed: int = 0) -> None:
    """The reference zero-initialises mlp_disp[-1] (model.py:516-518), which makes every upstream gradient vanish;
    benchmarks / tests use N(0, 0.05^2) instead."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        mlp_disp_last.weight.copy_((torch.randn(3, 256, generator=g) * 0.05).to(mlp_disp_last.weight.device))
        mlp_disp_last.bias.copy_((torch.randn(3, generator=g) * 0.05).to(mlp_disp_last.bias.device))
