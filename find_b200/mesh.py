"""find_b200/mesh.py — drop-ins for the PyTorch3D mesh calls on FIND's 3-D loss path, on the sm_100a kernels of csrc/mesh.cu.

Mirrors (names, argument meaning, return structure):
  * ``pytorch3d.ops.sample_points_from_meshes``  — /root/reference/src/model/losses.py:39,63,67, src/eval/eval_3d.py:149-150
  * ``pytorch3d.loss.mesh_edge_loss`` / ``mesh_laplacian_smoothing(method="cot")`` and the reference's
    ``MeshSmoothnessLoss`` — /root/reference/src/model/losses.py:93-99 (enabled by cfgs/train_3d.yaml:21)
  * ``pytorch3d.structures.Meshes`` — only what those calls read (padded verts / faces, per-mesh face counts, textures).

All arithmetic happens in libfindb200.so through the C ABI; there is no torch/CPU fallback (CPU tensors raise).
The random draw is this library's own (Philox4x32-10 over an integer area CDF, see csrc/mesh.cu): same distribution as
PyTorch3D's multinomial + rand, different stream.  ``seed=None`` takes a seed from torch's global generator, so
``torch.manual_seed`` makes runs repeatable; ``face_idx=`` / ``uv=`` inject an external draw.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Union

import torch
from torch import nn

from . import _lib


class MeshTopology:
    """Static connectivity of one triangle mesh, prepared once (what ``Meshes.edges_packed()`` and the sparse pattern of
    ``cot_laplacian`` are in PyTorch3D): int32 faces [F,3], unique undirected edges [E,2], and the vertex -> (face, corner)
    incidence list in CSR form that lets one thread own one vertex (no atomics in the Laplacian kernels)."""

    def __init__(self, faces: torch.Tensor, num_verts: int):
        if faces.dim() != 2 or faces.shape[1] != 3:
            raise ValueError("faces must be [F,3]")
        dev = faces.device
        f = faces.to(torch.int64)
        self.V, self.F = int(num_verts), int(f.shape[0])
        if self.F and (int(f.min()) < 0 or int(f.max()) >= self.V):
            raise ValueError("faces index outside [0, V)")
        self.faces = f.to(torch.int32).contiguous()
        e = torch.cat([f[:, [1, 2]], f[:, [2, 0]], f[:, [0, 1]]], 0)
        e, _ = e.sort(dim=1)
        key = torch.unique(e[:, 0] * self.V + e[:, 1])
        self.edges = torch.stack([key // self.V, key % self.V], 1).to(torch.int32).contiguous()
        self.E = int(self.edges.shape[0])
        flat = f.reshape(-1)                                   # entry 3*face + corner
        order = torch.argsort(flat, stable=True)
        counts = torch.bincount(flat, minlength=self.V)
        self.vf_ptr = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), counts.cumsum(0)]).to(torch.int32).contiguous()
        self.vf_face = (order // 3).to(torch.int32).contiguous()
        self.vf_corner = (order % 3).to(torch.int32).contiguous()
        self.nnz = int(order.numel())

    def to(self, device):
        if self.faces.device == torch.device(device):
            return self
        t = object.__new__(MeshTopology)
        t.__dict__.update({k: (v.to(device) if torch.is_tensor(v) else v) for k, v in self.__dict__.items()})
        return t


class Meshes:
    """Minimal ``pytorch3d.structures.Meshes``: a batch of triangle meshes, ragged (lists) or padded tensors."""

    def __init__(self, verts: Union[torch.Tensor, Sequence[torch.Tensor]], faces: Union[torch.Tensor, Sequence[torch.Tensor]],
                 textures=None):
        if torch.is_tensor(verts):
            if verts.dim() != 3 or verts.shape[-1] != 3:
                raise ValueError("Verts tensor has incorrect dimensions.")
            self._verts = verts
            self._nv = None
        else:
            vl = list(verts)
            vmax = max(v.shape[0] for v in vl)
            self._verts = vl[0].new_zeros((len(vl), vmax, 3))
            for i, v in enumerate(vl):
                self._verts[i, : v.shape[0]] = v
            self._nv = torch.tensor([v.shape[0] for v in vl], dtype=torch.int64, device=vl[0].device)
        N = self._verts.shape[0]
        if torch.is_tensor(faces):
            if faces.dim() == 2:
                faces = faces.unsqueeze(0).expand(N, -1, -1)
            if faces.dim() != 3 or faces.shape[-1] != 3:
                raise ValueError("Faces tensor has incorrect dimensions.")
            self._faces = faces
            self._nf = None
        else:
            fl = list(faces)
            fmax = max(f.shape[0] for f in fl)
            self._faces = fl[0].new_full((len(fl), fmax, 3), 0)
            for i, f in enumerate(fl):
                self._faces[i, : f.shape[0]] = f
            self._nf = torch.tensor([f.shape[0] for f in fl], dtype=torch.int64, device=fl[0].device)
        if self._faces.shape[0] != N:
            raise ValueError("verts and faces must hold the same number of meshes")
        self.textures = textures
        self._topo: Optional[MeshTopology] = None

    def __len__(self):
        return self._verts.shape[0]

    @property
    def device(self):
        return self._verts.device

    def verts_padded(self):
        return self._verts

    def faces_padded(self):
        return self._faces

    def num_faces_per_mesh(self):
        if self._nf is None:
            return torch.full((len(self),), self._faces.shape[1], dtype=torch.int64, device=self._verts.device)
        return self._nf

    def num_verts_per_mesh(self):
        if self._nv is None:
            return torch.full((len(self),), self._verts.shape[1], dtype=torch.int64, device=self._verts.device)
        return self._nv

    def update_padded(self, new_verts_padded):
        m = Meshes(new_verts_padded, self._faces, self.textures)
        m._nf, m._nv, m._topo = self._nf, self._nv, self._topo
        return m


def _faces_shared(faces: torch.Tensor) -> bool:
    return faces.dim() == 2 or faces.shape[0] == 1 or faces.stride(0) == 0


def _as_i32_faces(meshes):
    """-> (int32 faces tensor [F,3] or [N,F,3] contiguous, shared?)  cached on the mesh object when the topology is static."""
    cached = getattr(meshes, "_faces_i32", None)
    if cached is not None:
        return cached
    faces = meshes.faces_padded()
    shared = _faces_shared(faces)
    f = (faces[0] if (shared and faces.dim() == 3) else faces).to(torch.int32).contiguous()
    try:
        meshes._faces_i32 = (f, shared)
    except AttributeError:
        pass
    return f, shared


def _draw_seed() -> int:
    return int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item())


class _SamplePoints(torch.autograd.Function):
    @staticmethod
    def forward(ctx, verts, faces_i32, faces_shared, num_faces, S, seed, face_idx_in, uv_in, attr):
        lib = _lib.load()
        dev = _lib.require_cuda(verts, faces_i32, num_faces, face_idx_in, uv_in, attr)
        N, V = verts.shape[0], verts.shape[1]
        F = faces_i32.shape[-2]
        verts_shared = N > 1 and verts.stride(0) == 0
        vc = verts[0].contiguous() if verts_shared else verts.contiguous()
        samples = torch.empty(N, S, 3, dtype=torch.float32, device=dev)
        fidx = torch.empty(N, S, dtype=torch.int64, device=dev)
        bary = torch.empty(N, S, 3, dtype=torch.float32, device=dev)
        C = int(attr.shape[-1]) if attr is not None else 0
        attr_shared = attr is not None and N > 1 and attr.stride(0) == 0
        ac = None if attr is None else (attr[0].contiguous() if attr_shared else attr.contiguous())
        attr_out = torch.empty(N, S, C, dtype=torch.float32, device=dev) if attr is not None else None
        ws = None
        if face_idx_in is None:
            ws = torch.empty(int(lib.findb200_sample_workspace_bytes(N, F)), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_sample_points_forward(
                _lib.ptr(vc), int(verts_shared), _lib.ptr(faces_i32), int(faces_shared), _lib.ptr(num_faces), N, V, F, S,
                seed, None, _lib.ptr(face_idx_in), _lib.ptr(uv_in), _lib.ptr(samples), _lib.ptr(fidx), _lib.ptr(bary), _lib.ptr(ac),
                int(attr_shared), C, _lib.ptr(attr_out), _lib.ptr(ws), ws.numel() if ws is not None else 0, _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_sample_points_forward")
        ctx.save_for_backward(faces_i32, fidx, bary)
        ctx.shape = (N, V, F, S, faces_shared, verts_shared)
        ctx.mark_non_differentiable(fidx, bary)
        if attr_out is not None:
            ctx.mark_non_differentiable(attr_out)
        return samples, fidx, bary, attr_out

    @staticmethod
    def backward(ctx, g_samples, _gf, _gb, _ga):
        faces_i32, fidx, bary = ctx.saved_tensors
        N, V, F, S, faces_shared, verts_shared = ctx.shape
        lib = _lib.load()
        dev = g_samples.device
        gv = torch.empty(N, V, 3, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_sample_points_backward(_lib.ptr(g_samples.contiguous()), _lib.ptr(faces_i32), int(faces_shared),
                                                     _lib.ptr(fidx), _lib.ptr(bary), N, V, F, S, 1.0, 0, _lib.ptr(gv),
                                                     _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_sample_points_backward")
        if verts_shared:
            gv = gv.sum(0, keepdim=True).expand(N, -1, -1)
        return gv, None, None, None, None, None, None, None, None


def sample_points_from_meshes(meshes, num_samples: int = 10000, return_normals: bool = False, return_textures: bool = False,
                              *, seed: Optional[int] = None, face_idx: Optional[torch.Tensor] = None,
                              uv: Optional[torch.Tensor] = None, return_draw: bool = False):
    """pytorch3d.ops.sample_points_from_meshes: -> samples [N,S,3] (+ textures [N,S,C] if return_textures).
    Keyword extras: ``seed`` (int), ``face_idx`` [N,S] int64 mesh-local + ``uv`` [2,N,S] to inject a draw,
    ``return_draw`` to also get (face_idx, bary)."""
    if return_normals:
        raise NotImplementedError("return_normals is not used by FIND (losses.py:39,63,67) and is not built")
    verts = meshes.verts_padded()
    if verts.dtype != torch.float32:
        raise RuntimeError("sample_points_from_meshes: verts must be float32")
    faces_i32, shared = _as_i32_faces(meshes)
    nf = getattr(meshes, "_nf", None)
    attr = None
    if return_textures:
        tex = meshes.textures
        if tex is None:
            raise ValueError("Meshes do not contain textures.")
        attr = tex.verts_features_padded() if hasattr(tex, "verts_features_padded") else tex
        if attr.dim() == 2:
            attr = attr.unsqueeze(0).expand(verts.shape[0], -1, -1)
        attr = attr.to(torch.float32)
    if (face_idx is None) != (uv is None):
        raise ValueError("face_idx and uv must be given together")
    if face_idx is not None:
        face_idx = face_idx.to(torch.int64).contiguous()
        uv = uv.to(torch.float32).contiguous()
    if seed is None:
        seed = _draw_seed() if face_idx is None else 0
    samples, fidx, bary, attr_out = _SamplePoints.apply(verts, faces_i32, shared, nf, int(num_samples), int(seed), face_idx,
                                                        uv, attr)
    out = (samples,)
    if return_textures:
        out = out + (attr_out,)
    if return_draw:
        out = out + (fidx, bary)
    return out[0] if len(out) == 1 else out


# ---- MeshSmoothnessLoss ------------------------------------------------------------------------------------------
def _topology_of(meshes) -> MeshTopology:
    topo = getattr(meshes, "_topo", None)
    if topo is None:
        faces = meshes.faces_padded()
        if not _faces_shared(faces):
            # per-mesh topologies: FIND only smooths the deformed template (model.py:1032-1033), one topology per batch
            raise NotImplementedError("mesh smoothness needs one topology shared by the batch (the deformed template)")
        topo = MeshTopology(faces[0] if faces.dim() == 3 else faces, meshes.verts_padded().shape[1])
        try:
            meshes._topo = topo
        except AttributeError:
            pass
    return topo


class _MeshSmooth(torch.autograd.Function):
    """forward computes both loss terms AND d(w_edge*edge + w_lap*lap)/d verts in one enqueue (the weights are known
    up front); backward scales the stored gradient by the upstream scalar."""

    @staticmethod
    def forward(ctx, verts, topo: MeshTopology, target_length, w_edge, w_lap):
        lib = _lib.load()
        dev = _lib.require_cuda(verts)
        vc = verts.contiguous()
        N, V = vc.shape[0], vc.shape[1]
        topo = topo.to(dev)
        need_grad = ctx.needs_input_grad[0]
        loss2 = torch.empty(2, dtype=torch.float32, device=dev)
        gv = torch.empty_like(vc) if need_grad else None
        ws = torch.empty(int(lib.findb200_mesh_smooth_workspace_bytes(N, V, topo.nnz)), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_mesh_smooth(_lib.ptr(vc), N, V, _lib.ptr(topo.faces), topo.F, _lib.ptr(topo.edges), topo.E,
                                          _lib.ptr(topo.vf_ptr), _lib.ptr(topo.vf_face), _lib.ptr(topo.vf_corner), topo.nnz,
                                          float(target_length), float(w_edge), float(w_lap), None, _lib.ptr(loss2),
                                          _lib.ptr(gv), 0, _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_mesh_smooth")
        ctx.save_for_backward(gv)
        total = float(w_edge) * loss2[0] + float(w_lap) * loss2[1]
        ctx.mark_non_differentiable(loss2)
        return total, loss2

    @staticmethod
    def backward(ctx, g_total, _g2):
        (gv,) = ctx.saved_tensors
        return gv * g_total, None, None, None, None


def mesh_edge_loss(meshes, target_length: float = 0.0) -> torch.Tensor:
    """pytorch3d.loss.mesh_edge_loss (losses.py:95)."""
    total, _ = _MeshSmooth.apply(meshes.verts_padded(), _topology_of(meshes), target_length, 1.0, 0.0)
    return total


def mesh_laplacian_smoothing(meshes, method: str = "cot") -> torch.Tensor:
    """pytorch3d.loss.mesh_laplacian_smoothing — only method="cot" is on FIND's path (losses.py:96)."""
    if method != "cot":
        raise NotImplementedError('only method="cot" is built (what MeshSmoothnessLoss uses, losses.py:96)')
    total, _ = _MeshSmooth.apply(meshes.verts_padded(), _topology_of(meshes), 0.0, 0.0, 1.0)
    return total


class MeshSmoothnessLoss(nn.Module):
    """/root/reference/src/model/losses.py:93-99: 0.1 * laplacian(cot) + 10 * edge, one fused enqueue."""

    def forward(self, meshes) -> torch.Tensor:
        total, _ = _MeshSmooth.apply(meshes.verts_padded(), _topology_of(meshes), 0.0, 10.0, 0.1)
        return total


__all__ = ["Meshes", "MeshTopology", "sample_points_from_meshes", "mesh_edge_loss", "mesh_laplacian_smoothing",
           "MeshSmoothnessLoss"]
