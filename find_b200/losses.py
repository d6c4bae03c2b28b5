"""find_b200/losses.py — the 3-D loss of FIND's training step on the sm_100a Chamfer kernels.

Mirrors /root/reference/src/model/losses.py:59-90 (``DisplacementLoss``) and the weighting of
``ModelWithLoss.forward`` (/root/reference/src/model/model.py:1026-1030,1157-1158; weights opts.py:97-99).

The reference draws 5000 random surface samples from both meshes with PyTorch3D's ``sample_points_from_meshes``
before Chamfer (losses.py:61,63,67): ``num_samples=5000`` is the default here too, so an unmodified ``ModelWithLoss`` call
keeps its meaning — both meshes are surface-sampled on the GPU first (find_b200.mesh; same
distribution as PyTorch3D's sampler, its own Philox stream — ``seed`` makes it repeatable).
BASELINE.json's configurations take Chamfer directly between the deformed template vertices and a target point cloud:
that is the EXPLICIT opt-out ``num_samples=None`` ("use the points as given"):
  * prediction = ``res['verts']`` ([B,V,3], what ``res['meshes'].verts_padded()`` holds),
  * target     = ``batch['points']`` ([B,P,3] tensor or ``Pointclouds``), else ``batch['mesh'].verts_padded()``.
``TextureLossGTSpace`` (losses.py:22-57) is the masked colour MSE at surface samples of the ground-truth mesh.
"""
from __future__ import annotations

import torch
from torch import nn

from .chamfer import Pointclouds, chamfer_distance

CHAMF_WEIGHT = 1e4      # opts.py:97  (--weight_chamf default)
TEX_WEIGHT = 1.0        # opts.py:99


def _target_points(batch):
    if "points" in batch:
        return batch["points"]
    mesh = batch["mesh"]
    if hasattr(mesh, "verts_padded"):
        return mesh.verts_padded()
    raise RuntimeError("DisplacementLoss: batch needs 'points' ([B,P,3] or Pointclouds) or 'mesh' with verts_padded()")


class DisplacementLoss(nn.Module):
    """losses.py:59-90 -> dict(loss=chamfer).  All three z-cutoff variants are kept (ragged clouds)."""

    def forward(self, model, res, batch, epoch=0, num_samples=5000, z_cutoff=None, gt_z_cutoff=None, seed=None):
        if num_samples is not None:                                # losses.py:63,67: 5000 surface samples of both meshes
            from .mesh import sample_points_from_meshes
            s_gt = None if seed is None else 2 * int(seed)
            s_pr = None if seed is None else 2 * int(seed) + 1
            gt_samples = (sample_points_from_meshes(batch["mesh"], num_samples=num_samples, seed=s_gt)
                          if "mesh" in batch and hasattr(batch["mesh"], "faces_padded") else _target_points(batch))
            pred_samples = sample_points_from_meshes(res["meshes"], num_samples=num_samples, seed=s_pr)
        else:
            gt_samples = _target_points(batch)
            pred_samples = res["verts"] if "verts" in res else res["meshes"].verts_padded()

        if z_cutoff is not None:                                   # losses.py:69-77
            gt_pad = gt_samples.points_padded() if isinstance(gt_samples, Pointclouds) else gt_samples
            pred_verts, gt_verts = [], []
            for b in range(len(pred_samples)):
                pred_verts.append(pred_samples[b][pred_samples[b, :, 2] <= z_cutoff])
                gt_verts.append(gt_pad[b][gt_pad[b, :, 2] <= z_cutoff])
            chamf_loss, _ = chamfer_distance(Pointclouds(pred_verts), Pointclouds(gt_verts))
        elif gt_z_cutoff is not None:                              # losses.py:79-85
            gt_pad = gt_samples.points_padded() if isinstance(gt_samples, Pointclouds) else gt_samples
            gt_verts = [gt_pad[b][gt_pad[b, :, 2] <= gt_z_cutoff] for b in range(len(gt_pad))]
            chamf_loss, _ = chamfer_distance(pred_samples, Pointclouds(gt_verts))
        else:                                                      # losses.py:87-88 (what training uses)
            chamf_loss, _ = chamfer_distance(pred_samples, gt_samples)
        return dict(loss=chamf_loss)


class TextureLossGTSpace(nn.Module):
    """losses.py:22-57 — colour loss in ground-truth space: sample ``num_samples`` points AND their colours on the GT meshes
    (``sample_points_from_meshes(..., return_textures=True)``), evaluate the field's colour head AT those per-foot points
    (the second call site of the field, ``pos_shared = 0``), squared error masked to non-white GT samples
    (``mask = (gt_col < 1).any(-1)``), mean over ALL elements (masked ones count as zeros, as in the reference)."""

    def forward(self, model, batch: dict, num_samples=1000, shapevec=None, texvec=None, posevec=None, seed=None) -> torch.Tensor:
        from .mesh import sample_points_from_meshes
        mesh_gt = batch["mesh"]
        sampled_verts, sampled_gt_colours = sample_points_from_meshes(mesh_gt, num_samples=num_samples, return_textures=True,
                                                                      seed=seed)
        mask = (sampled_gt_colours < 1).any(dim=-1).unsqueeze(-1).expand(-1, -1, 3)
        texvec = texvec if texvec is not None else batch.get("texvec", None)
        shapevec = shapevec if shapevec is not None else batch.get("shapevec", None)
        posevec = posevec if posevec is not None else batch.get("posevec", None)
        res = model(sampled_verts.detach(), texvec=texvec, shapevec=shapevec, posevec=posevec)
        loss = torch.nn.functional.mse_loss(res["col"], sampled_gt_colours[..., :3].detach(), reduction="none")
        loss = loss * mask                                       # multiply out by mask to disregard white points
        return loss.mean()


class FieldChamferObjective(nn.Module):
    """The bench / training objective of SURVEY.md §8(d): ``1e4 * Chamfer(verts, target) + 1 * MSE(col, col_target)``
    (the colour term stands in for TextureLossGTSpace, losses.py:22-57, so that both heads carry gradient)."""

    def __init__(self, weight_chamf: float = CHAMF_WEIGHT, weight_tex: float = TEX_WEIGHT):
        super().__init__()
        self.disp_loss = DisplacementLoss()
        self.weight_chamf, self.weight_tex = weight_chamf, weight_tex

    def forward(self, model, batch, is_train=True):
        res = model.get_meshes_from_batch(batch, is_train=is_train)
        losses = {"chamf": self.disp_loss(model, res, batch, 0, num_samples=None)["loss"]}
        total = self.weight_chamf * losses["chamf"]
        if "col_target" in batch and self.weight_tex:
            losses["tex"] = torch.nn.functional.mse_loss(res["col"], batch["col_target"])
            total = total + self.weight_tex * losses["tex"]
        return total, losses


__all__ = ["DisplacementLoss", "TextureLossGTSpace", "FieldChamferObjective", "CHAMF_WEIGHT", "TEX_WEIGHT"]
