"""oracle/mesh_oracle.py — CPU restatement of the mesh side of FIND's 3-D loss.  TEST INFRASTRUCTURE ONLY: imported
by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg; never by the product path.

Follows
  * /root/reference/src/model/losses.py:39,63,67 (``sample_points_from_meshes`` call sites), :93-99
    (``MeshSmoothnessLoss`` = 0.1 * mesh_laplacian_smoothing(method="cot") + 10 * mesh_edge_loss);
  * PyTorch3D @ 1706eb82 (requirements_mac_linux.txt:31, NOT on this box) for the arithmetic, restated from its published
    algorithm: ops/sample_points_from_meshes.py (area-weighted multinomial over faces, `_rand_barycentric_coords`:
    w0 = 1 - sqrt(u), w1 = sqrt(u)(1 - v), w2 = sqrt(u) v, samples = w0*a + w1*b + w2*c), loss/mesh_edge_loss.py,
    loss/mesh_laplacian_smoothing.py and ops/laplacian_matrices.py::cot_laplacian (Heron area clamped at 1e-12,
    cot / 4, symmetric, built under no_grad).

PARITY UNPINNED against PyTorch3D itself (absent, the reference ships no vectors); pinned by known-answer tests in
tests/test_oracle_mesh.py (equilateral / right triangles, regular grids, fp64 dense-matrix cross-check, chi-square of the
face draw against the areas).  The random draw is this repo's own definition (csrc/mesh.cu header): the oracle
restates it bit for bit (Philox4x32-10, integer CDF), so CUDA-vs-oracle parity on face ids and samples is exact.
"""
from __future__ import annotations

import numpy as np
import torch

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
U32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32-10 on uint64-held 32-bit lanes (vectorised).  Known-answer vectors: tests/test_oracle_mesh.py."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & U32 for c in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0 & 0xFFFFFFFF), np.uint64(k1 & 0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = np.uint64(M0) * c0, np.uint64(M1) * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & U32, p1 >> np.uint64(32), p1 & U32
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + np.uint64(W0)) & U32, (k1 + np.uint64(W1)) & U32
    return c0, c1, c2, c3


def face_areas(verts: np.ndarray, faces: np.ndarray) -> np.ndarray:
    """0.5 |e1 x e2| in fp32 with every operation rounded once (numpy float32 arithmetic does exactly that)."""
    v = verts.astype(np.float32)
    a, b, c = v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]]
    e1, e2 = b - a, c - a
    cx = e1[:, 1] * e2[:, 2] - e1[:, 2] * e2[:, 1]
    cy = e1[:, 2] * e2[:, 0] - e1[:, 0] * e2[:, 2]
    cz = e1[:, 0] * e2[:, 1] - e1[:, 1] * e2[:, 0]
    n2 = (cx * cx + cy * cy) + cz * cz
    area = np.float32(0.5) * np.sqrt(n2, dtype=np.float32)
    area[~(area >= 0) | np.isinf(area)] = 0
    return area.astype(np.float32)


def area_cdf(area: np.ndarray) -> np.ndarray:
    amax = float(area.max()) if area.size else 0.0
    if amax <= 0.0:
        return np.zeros(area.shape, dtype=np.uint64)
    q = np.floor(area.astype(np.float64) / np.float64(amax) * 4294967296.0).astype(np.uint64)
    return np.cumsum(q, dtype=np.uint64)


def draw(cdf: np.ndarray, mesh: int, S: int, seed: int, counter: int = 0):
    """-> face [S] int64, u [S] f32, v [S] f32 for mesh index `mesh`."""
    s = np.arange(S, dtype=np.uint64)
    r0, r1, r2, r3 = philox4x32_10(s, np.full(S, mesh, np.uint64), np.full(S, counter & 0xFFFFFFFF, np.uint64),
                                   np.full(S, (counter >> 32) & 0xFFFFFFFF, np.uint64),
                                   seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    total = int(cdf[-1])
    r64 = [(int(a) << 32) | int(b) for a, b in zip(r0, r1)]
    target = np.array([(r * total) >> 64 for r in r64], dtype=np.uint64)
    face = np.searchsorted(cdf, target, side="right").astype(np.int64) if total else np.zeros(S, np.int64)
    face = np.minimum(face, len(cdf) - 1)
    u = (r2 >> np.uint64(8)).astype(np.float32) * np.float32(2.0 ** -24)
    v = (r3 >> np.uint64(8)).astype(np.float32) * np.float32(2.0 ** -24)
    return face, u, v


def barycentric(u: np.ndarray, v: np.ndarray):
    u, v = u.astype(np.float32), v.astype(np.float32)
    us = np.sqrt(u, dtype=np.float32)
    one = np.float32(1.0)
    return one - us, us * (one - v), us * v


def interpolate(values: np.ndarray, faces: np.ndarray, face: np.ndarray, w0, w1, w2) -> np.ndarray:
    """w0*a + w1*b + w2*c with PyTorch3D's operation order; values [V,C], -> [S,C] fp32."""
    x = values.astype(np.float32)
    a, b, c = x[faces[face, 0]], x[faces[face, 1]], x[faces[face, 2]]
    return (w0[:, None] * a + w1[:, None] * b) + w2[:, None] * c


def sample_points(verts: np.ndarray, faces, num_samples: int, seed: int = 0, face_idx=None, uv=None, num_faces=None,
                  attr=None, counter: int = 0):
    """verts [N,V,3]; faces [F,3] shared or [N,F,3]; -> dict(samples [N,S,3], face_idx [N,S], bary [N,S,3], attr [N,S,C])."""
    N = verts.shape[0]
    out = dict(samples=[], face_idx=[], bary=[], attr=[])
    for n in range(N):
        fc = np.asarray(faces if np.ndim(faces) == 2 else faces[n]).astype(np.int64)
        F = fc.shape[0]
        if face_idx is None or uv is None:
            area = face_areas(verts[n], fc)
            if num_faces is not None:
                area[int(num_faces[n]):] = 0
            f_d, u_d, v_d = draw(area_cdf(area), n, num_samples, seed, counter)
        f = np.clip(np.asarray(face_idx[n]).astype(np.int64), 0, F - 1) if face_idx is not None else f_d
        u, v = (np.asarray(uv[0][n]), np.asarray(uv[1][n])) if uv is not None else (u_d, v_d)
        w0, w1, w2 = barycentric(u, v)
        out["samples"].append(interpolate(verts[n], fc, f, w0, w1, w2))
        out["face_idx"].append(f)
        out["bary"].append(np.stack([w0, w1, w2], -1))
        if attr is not None:
            out["attr"].append(interpolate(attr[n], fc, f, w0, w1, w2))
    return {k: np.stack(v) for k, v in out.items() if v}


def sample_points_backward(grad_samples: np.ndarray, faces, face_idx: np.ndarray, bary: np.ndarray, V: int) -> np.ndarray:
    N, S, _ = grad_samples.shape
    g = np.zeros((N, V, 3), dtype=np.float64)
    for n in range(N):
        fc = np.asarray(faces if np.ndim(faces) == 2 else faces[n]).astype(np.int64)
        for k in range(3):
            np.add.at(g[n], fc[face_idx[n], k], bary[n][:, k:k + 1].astype(np.float64) * grad_samples[n].astype(np.float64))
    return g


# ---- topology helpers (what Meshes.edges_packed() gives: unique undirected edges) -------------------------------------
def unique_edges(faces: np.ndarray) -> np.ndarray:
    f = np.asarray(faces).astype(np.int64)
    e = np.concatenate([f[:, [1, 2]], f[:, [2, 0]], f[:, [0, 1]]], 0)
    e = np.sort(e, axis=1)
    return np.unique(e, axis=0)


def mesh_edge_loss(verts: torch.Tensor, faces: np.ndarray, target_length: float = 0.0) -> torch.Tensor:
    """verts [N,V,3] torch fp32 (autograd), one shared topology."""
    e = torch.as_tensor(unique_edges(faces))
    N = verts.shape[0]
    v0, v1 = verts[:, e[:, 0]], verts[:, e[:, 1]]
    loss = ((v0 - v1).norm(dim=2, p=2) - target_length) ** 2.0
    return (loss / e.shape[0]).sum() / N


def cot_laplacian_dense(v: torch.Tensor, faces: torch.Tensor, eps: float = 1e-12) -> torch.Tensor:
    """cot_laplacian of ONE mesh as a dense [V,V] matrix (same dtype as v)."""
    V = v.shape[0]
    fv = v[faces]
    v0, v1, v2 = fv[:, 0], fv[:, 1], fv[:, 2]
    A, B, C = (v1 - v2).norm(dim=1), (v0 - v2).norm(dim=1), (v0 - v1).norm(dim=1)
    s = 0.5 * (A + B + C)
    area = (s * (s - A) * (s - B) * (s - C)).clamp(min=eps).sqrt()
    A2, B2, C2 = A * A, B * B, C * C
    cot = torch.stack([(B2 + C2 - A2) / area, (A2 + C2 - B2) / area, (A2 + B2 - C2) / area], dim=1) / 4.0
    ii, jj = faces[:, [1, 2, 0]].reshape(-1), faces[:, [2, 0, 1]].reshape(-1)
    L = torch.zeros(V, V, dtype=v.dtype)
    L.index_put_((ii, jj), cot.reshape(-1), accumulate=True)
    return L + L.t()


def mesh_laplacian_smoothing_cot(verts: torch.Tensor, faces: np.ndarray) -> torch.Tensor:
    fc = torch.as_tensor(np.asarray(faces).astype(np.int64))
    N, V = verts.shape[0], verts.shape[1]
    total = verts.new_zeros(())
    for n in range(N):
        with torch.no_grad():
            L = cot_laplacian_dense(verts[n].detach(), fc)
            norm_w = L.sum(dim=1, keepdim=True)
            idx = norm_w > 0
            norm_w[idx] = 1.0 / norm_w[idx]
        loss = (L.mm(verts[n]) * norm_w - verts[n]).norm(dim=1)
        total = total + (loss / V).sum()
    return total / N


def mesh_smoothness_loss(verts: torch.Tensor, faces: np.ndarray, w_lap: float = 0.1, w_edge: float = 10.0):
    """losses.py:93-99 -> (loss, edge, laplacian)."""
    e, l = mesh_edge_loss(verts, faces), mesh_laplacian_smoothing_cot(verts, faces)
    return w_lap * l + w_edge * e, e, l
