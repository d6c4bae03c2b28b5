"""find_b200/chamfer.py — drop-in for the PyTorch3D calls on FIND's 3-D loss path.

Mirrors (names, argument meaning, return structure, error behaviour):
  * ``pytorch3d.ops.knn_points``          as used at /root/reference/src/eval/eval_3d.py:176-177
  * ``pytorch3d.loss.chamfer_distance``   as used at /root/reference/src/model/losses.py:77,85,88 and
                                          /root/reference/src/eval/eval_3d.py:151,159
  * ``pytorch3d.structures.Pointclouds``  (only what losses.py:69-85 needs: list of ragged clouds)

All arithmetic happens in libfindb200.so (csrc/chamfer.cu) through the C ABI; there is no torch/CPU fallback.
"""
from __future__ import annotations

from collections import namedtuple
from typing import List, Optional, Sequence, Union

import torch

from . import _lib

_KNN = namedtuple("KNN", "dists idx knn")


class Pointclouds:
    """Minimal stand-in for pytorch3d.structures.Pointclouds (ragged list of [P_i, 3] clouds)."""

    def __init__(self, points: Union[Sequence[torch.Tensor], torch.Tensor]):
        if torch.is_tensor(points):
            if points.dim() != 3 or points.shape[-1] != 3:
                raise ValueError("Points tensor has incorrect dimensions.")
            points = list(points.unbind(0))
        self._list: List[torch.Tensor] = list(points)
        for p in self._list:
            if p.dim() != 2 or p.shape[1] != 3:
                raise ValueError("Clouds in list must be of shape Px3 or empty")

    def __len__(self):
        return len(self._list)

    def points_list(self):
        return self._list

    def num_points_per_cloud(self) -> torch.Tensor:
        dev = self._list[0].device if self._list else torch.device("cpu")
        return torch.tensor([p.shape[0] for p in self._list], dtype=torch.int64, device=dev)

    def points_padded(self) -> torch.Tensor:
        n = len(self._list)
        pmax = max([p.shape[0] for p in self._list], default=0)
        ref = self._list[0] if self._list else torch.zeros(0, 3)
        out = ref.new_zeros((n, pmax, 3))
        for i, p in enumerate(self._list):
            out[i, : p.shape[0]] = p
        return out


def _prep_lengths(lengths: Optional[torch.Tensor], n: int, p: int, device) -> Optional[torch.Tensor]:
    if lengths is None:
        return None
    if lengths.shape != (n,):
        raise ValueError("Expected lengths to be of shape (N,)")
    lengths = lengths.to(device=device, dtype=torch.int64).contiguous()
    return lengths


def _workspace(n: int, p1: int, p2: int, device) -> torch.Tensor:
    nbytes = _lib.load().findb200_knn_workspace_bytes(n, p1, p2)
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


class _KnnFunction(torch.autograd.Function):
    """knn_points(K=1) forward/backward through findb200_knn1 / findb200_knn1_backward."""

    @staticmethod
    def forward(ctx, p1, p2, lengths1, lengths2):
        dev = _lib.require_cuda(p1, p2)
        lib = _lib.load()
        p1c, p2c = p1.detach().contiguous().float(), p2.detach().contiguous().float()
        n, P1, _ = p1c.shape
        P2 = p2c.shape[1]
        dists = torch.empty((n, P1), dtype=torch.float32, device=dev)
        idx = torch.empty((n, P1), dtype=torch.int64, device=dev)
        ws = _workspace(n, P1, P2, dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_knn1(_lib.ptr(p1c), _lib.ptr(p2c), _lib.ptr(lengths1), _lib.ptr(lengths2), n, P1, P2,
                                   _lib.ptr(dists), _lib.ptr(idx), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_knn1")
        ctx.save_for_backward(p1c, p2c, idx)
        ctx.lengths = (lengths1, lengths2)
        ctx.mark_non_differentiable(idx)
        return dists, idx

    @staticmethod
    def backward(ctx, grad_dists, _grad_idx):
        p1c, p2c, idx = ctx.saved_tensors
        lengths1, lengths2 = ctx.lengths
        dev = p1c.device
        lib = _lib.load()
        n, P1, _ = p1c.shape
        P2 = p2c.shape[1]
        g = grad_dists.contiguous().float()
        g1 = torch.empty_like(p1c)
        g2 = torch.empty_like(p2c)
        with torch.cuda.device(dev):
            rc = lib.findb200_knn1_backward(_lib.ptr(p1c), _lib.ptr(p2c), _lib.ptr(lengths1), _lib.ptr(lengths2),
                                            _lib.ptr(idx), _lib.ptr(g), n, P1, P2, _lib.ptr(g1), _lib.ptr(g2),
                                            _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_knn1_backward")
        return g1, g2, None, None


def knn_points(p1: torch.Tensor, p2: torch.Tensor, lengths1: Optional[torch.Tensor] = None,
               lengths2: Optional[torch.Tensor] = None, norm: int = 2, K: int = 1, version: int = -1,
               return_nn: bool = False, return_sorted: bool = True):
    """pytorch3d.ops.knn_points for K=1, norm=2 (what FIND uses).  -> KNN(dists [N,P1,1], idx [N,P1,1], knn)."""
    if p1.shape[0] != p2.shape[0]:
        raise ValueError("pts1 and pts2 must have the same batch dimension.")
    if p1.shape[2] != p2.shape[2]:
        raise ValueError("pts1 and pts2 must have the same point dimension.")
    if p1.shape[2] != 3:
        raise ValueError("find_b200.knn_points supports 3-D points only")
    if K != 1 or norm != 2:
        raise ValueError("find_b200.knn_points implements K=1, norm=2 only (the FIND hot path)")
    n, P1 = p1.shape[0], p1.shape[1]
    P2 = p2.shape[1]
    dev = _lib.require_cuda(p1, p2)
    lengths1 = _prep_lengths(lengths1, n, P1, dev)
    lengths2 = _prep_lengths(lengths2, n, P2, dev)
    dists, idx = _KnnFunction.apply(p1, p2, lengths1, lengths2)
    dists, idx = dists.unsqueeze(-1), idx.unsqueeze(-1)
    knn = None
    if return_nn:
        knn = torch.gather(p2, 1, idx.expand(-1, -1, 3)).unsqueeze(2)
    return _KNN(dists=dists, idx=idx, knn=knn)


class _ChamferFunction(torch.autograd.Function):
    """Fused bidirectional Chamfer (mean/mean) through findb200_chamfer_forward / _backward."""

    @staticmethod
    def forward(ctx, x, y, x_lengths, y_lengths):
        dev = _lib.require_cuda(x, y)
        lib = _lib.load()
        xc, yc = x.detach().contiguous().float(), y.detach().contiguous().float()
        n, P1, _ = xc.shape
        P2 = yc.shape[1]
        loss = torch.empty((), dtype=torch.float32, device=dev)
        idx_x = torch.empty((n, P1), dtype=torch.int64, device=dev)
        idx_y = torch.empty((n, P2), dtype=torch.int64, device=dev)
        ws = _workspace(n, P1, P2, dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_chamfer_forward(_lib.ptr(xc), _lib.ptr(yc), _lib.ptr(x_lengths), _lib.ptr(y_lengths),
                                              n, P1, P2, _lib.ptr(loss), None, _lib.ptr(idx_x), None, _lib.ptr(idx_y),
                                              _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_chamfer_forward")
        ctx.save_for_backward(xc, yc, idx_x, idx_y)
        ctx.lengths = (x_lengths, y_lengths)
        ctx.need_y = y.requires_grad
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        xc, yc, idx_x, idx_y = ctx.saved_tensors
        x_lengths, y_lengths = ctx.lengths
        dev = xc.device
        lib = _lib.load()
        n, P1, _ = xc.shape
        P2 = yc.shape[1]
        g = grad_loss.contiguous().float()
        gx = torch.empty_like(xc)
        gy = torch.empty_like(yc) if ctx.need_y else None
        with torch.cuda.device(dev):
            rc = lib.findb200_chamfer_backward(_lib.ptr(xc), _lib.ptr(yc), _lib.ptr(x_lengths), _lib.ptr(y_lengths),
                                               n, P1, P2, _lib.ptr(idx_x), _lib.ptr(idx_y), _lib.ptr(g), 1.0,
                                               _lib.ptr(gx), _lib.ptr(gy), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_chamfer_backward")
        return gx, gy, None, None


def _handle_pointcloud_input(points, lengths):
    """pytorch3d.loss.chamfer._handle_pointcloud_input: Pointclouds or padded tensor (+ optional lengths)."""
    if isinstance(points, Pointclouds):
        X = points.points_padded()
        lengths = points.num_points_per_cloud()
        return X, lengths
    if torch.is_tensor(points):
        if points.ndim != 3:
            raise ValueError("Expected points to be of shape (N, P, D)")
        if lengths is not None and (lengths.ndim != 1 or lengths.shape[0] != points.shape[0]):
            raise ValueError("Expected lengths to be of shape (N,)")
        return points, lengths
    raise ValueError("The input pointclouds should be either Pointclouds objects or torch.Tensor of shape "
                     "(minibatch, num_points, 3).")


def chamfer_distance(x, y, x_lengths=None, y_lengths=None, x_normals=None, y_normals=None, weights=None,
                     batch_reduction: Optional[str] = "mean", point_reduction: Optional[str] = "mean", norm: int = 2,
                     single_directional: bool = False, abs_cosine: bool = True):
    """pytorch3d.loss.chamfer_distance (no normals).  -> (loss, None).

    The default reductions (the only ones FIND uses) run as ONE fused bidirectional kernel pass; any other
    reduction / weights / single_directional combination is assembled from ``knn_points`` outputs."""
    if batch_reduction is not None and batch_reduction not in ("mean", "sum"):
        raise ValueError('batch_reduction must be one of ["mean", "sum"] or None')
    if point_reduction is not None and point_reduction not in ("mean", "sum"):
        raise ValueError('point_reduction must be one of ["mean", "sum"] or None')
    if point_reduction is None and batch_reduction is not None:
        raise ValueError("Batch reduction must be None if point_reduction is None")
    if norm != 2:
        raise ValueError("find_b200.chamfer_distance implements norm=2 only (the FIND hot path)")
    if x_normals is not None or y_normals is not None:
        raise ValueError("find_b200.chamfer_distance does not implement the normals term (unused by FIND)")
    x, x_lengths = _handle_pointcloud_input(x, x_lengths)
    y, y_lengths = _handle_pointcloud_input(y, y_lengths)
    if y.shape[0] != x.shape[0] or y.shape[2] != x.shape[2]:
        raise ValueError("y does not have the correct shape.")
    n, P1, P2 = x.shape[0], x.shape[1], y.shape[1]
    dev = _lib.require_cuda(x, y)
    x_lengths = _prep_lengths(x_lengths, n, P1, dev)
    y_lengths = _prep_lengths(y_lengths, n, P2, dev)

    if weights is None and batch_reduction == "mean" and point_reduction == "mean" and not single_directional:
        return _ChamferFunction.apply(x, y, x_lengths, y_lengths), None

    if weights is not None:
        if weights.size(0) != n:
            raise ValueError("weights must be of shape (N,).")
        if not (weights >= 0).all():
            raise ValueError("weights cannot be negative.")

    def one_direction(a, b, la, lb):
        P = a.shape[1]
        d = knn_points(a, b, la, lb, K=1).dists[..., 0]
        if la is not None:
            mask = torch.arange(P, device=dev)[None] >= la[:, None]
            d = d.masked_fill(mask, 0.0)
        if weights is not None:
            d = d * weights.view(n, 1)
        if point_reduction is not None:
            d = d.sum(1)
            if point_reduction == "mean":
                lens = la if la is not None else torch.full((n,), P, dtype=torch.int64, device=dev)
                d = d / lens.clamp(min=1)
            if batch_reduction is not None:
                d = d.sum()
                if batch_reduction == "mean":
                    d = d / (weights.sum() if weights is not None else max(n, 1))
        return d

    cham_x = one_direction(x, y, x_lengths, y_lengths)
    if single_directional:
        return cham_x, None
    cham_y = one_direction(y, x, y_lengths, x_lengths)
    if point_reduction is not None:
        return cham_x + cham_y, None
    return (cham_x, cham_y), None
