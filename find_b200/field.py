"""find_b200/field.py — drop-in for FIND's field module: same names, ctor kwargs, attributes, ``state_dict`` keys.

Mirrors /root/reference/src/model/model.py:
  * ``LatentVector``               model.py:92-152   (zero-init nn.Parameter table, int / tensor / label indexing)
  * ``NeuralDisplacementField``    model.py:206-534  (ctor kwargs 207-216, forward 393-453, get_meshes 455-504,
                                                      get_meshes_from_batch 506-514, reset_weights 516-518,
                                                      configure_template 520-530, save_model/load/freeze 155-196)
  * ``FourierFeatureTransform``    /root/reference/src/utils/fourier_feature_transform.py:7-55
  * ``sample_latent_vectors``      /root/reference/src/train/trainer.py:29-46

The arithmetic (Fourier encoding, 13 Linear layers, latent concat, tanh heads, registration transform and the whole
autograd) runs in libfindb200.so through ``field_ops.field_apply``; the nn.Linear modules below only hold the fp32
master parameters so optimisers and checkpoints see exactly the reference's tensors.  There is no CPU path.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import numpy as np
import torch
from torch import nn

from .field_ops import PARAM_NAMES, field_apply


def make_params_list(*params):
    """model.py:47-60"""
    out = []
    for param in params:
        if hasattr(param, "parameters"):
            out += list(param.parameters())
        elif isinstance(param, nn.Parameter):
            out += [param]
        elif param is None:
            pass
        else:
            raise ValueError(f"Param type {type(param)} not understood.")
    return out


class FourierFeatureTransform(nn.Module):
    """Holder of the Gaussian matrix ``_B`` (fourier_feature_transform.py:17-26).  Like the reference it reseeds the
    GLOBAL torch RNG with 1, sorts the 3 rows of B by norm, and keeps ``_B`` as a plain attribute (not in
    ``state_dict``).  The encoding itself is fused into the first trunk GEMM of the CUDA kernels."""

    def __init__(self, num_input_channels, mapping_size=256, scale=10, exclude=0):
        super().__init__()
        self._num_input_channels = num_input_channels
        self._mapping_size = mapping_size
        self.exclude = exclude
        torch.manual_seed(1)
        B = torch.randn((num_input_channels, mapping_size)) * scale
        B_sort = sorted(B, key=lambda x: torch.norm(x, p=2))
        self._B = torch.stack(B_sort)
        self._B_dev: Dict[torch.device, torch.Tensor] = {}

    def B_on(self, device) -> torch.Tensor:
        device = torch.device(device)
        if device not in self._B_dev:
            self._B_dev[device] = self._B.to(device=device, dtype=torch.float32).contiguous()
        return self._B_dev[device]

    def forward(self, x):
        raise RuntimeError("find_b200 fuses the Fourier encoding into the field kernels; call the field module "
                           "(NeuralDisplacementField.forward) instead of the encoder alone")


class LatentVector(nn.Module):
    """model.py:92-152 — learned per-item latent table."""

    def __init__(self, dataset_size=None, vec_size=512, name="", device="cuda", key=None, labels: list = None,
                 init_values=None):
        super().__init__()
        self.dataset_size = dataset_size
        self.vec_size = vec_size
        if labels is not None:
            dataset_size = len(labels)
        self.labels = labels
        self.key = key
        vec_init = torch.zeros(dataset_size, vec_size)
        if init_values is not None:
            if isinstance(init_values, np.ndarray):
                init_values = torch.from_numpy(init_values).unsqueeze(0).float()
            else:
                raise NotImplementedError
            vec_init[:] = init_values
        self.data = nn.Parameter(vec_init.to(device))
        self.name = name
        self._label_index = None if labels is None else {}
        if labels is not None:
            for i, lab in enumerate(labels):          # first occurrence wins, like list.index
                self._label_index.setdefault(lab, i)

    def __len__(self):
        return self.dataset_size

    def _lookup(self, label):
        if label not in self._label_index:
            raise ValueError(f"{label!r} is not in list")   # what list.index raises in the reference
        return self._label_index[label]

    def __getitem__(self, idx):
        if isinstance(idx, int):
            return self.data[idx]
        if isinstance(idx, torch.Tensor):
            return self.data[idx]
        if isinstance(idx, str):
            assert self.labels is not None, \
                f"Tried to access item {idx} from LatentVector {self.name}, LV does not have labels"
            return self.data[self._lookup(idx)]
        if isinstance(idx, list):
            if isinstance(idx[0], int):
                return self.data[idx]
            if isinstance(idx[0], str):
                assert self.labels is not None, \
                    f"Tried to access item {idx} from LatentVector {self.name}, LV does not have labels"
                return self.data[[self._lookup(o) for o in idx]]
        else:
            raise NotImplementedError(f"Didn't understand indexing of LatentVector {self.name}, type {type(idx)}")


def sample_latent_vectors(batch, latent_vectors):
    """trainer.py:29-46"""
    if latent_vectors is None:
        return {}
    out = {}
    for vec in latent_vectors:
        if vec.labels is not None:
            label_key = vec.key
            assert label_key in batch, \
                f"Trying to sample from latent vector {label_key} using keys, but not found in dataset"
            out[vec.name] = vec[batch[label_key]]
        else:
            out[vec.name] = vec[batch["idx"]]
    return out


class TemplateMeshes:
    """What the hot path needs of ``pytorch3d.structures.Meshes`` (absent on this box): padded verts/faces + textures."""

    def __init__(self, verts: torch.Tensor, faces: torch.Tensor, textures=None):
        self._verts, self._faces, self.textures = verts, faces, textures

    def verts_padded(self):
        return self._verts

    def faces_padded(self):
        return self._faces

    def __len__(self):
        return self._verts.shape[0]

    def to(self, device):
        return TemplateMeshes(self._verts.to(device), self._faces.to(device), self.textures)

    def extend(self, N):
        """pytorch3d_tools.extend_template: stride-0 views, nothing materialised."""
        return TemplateMeshes(self._verts.expand(N, -1, -1), self._faces.expand(N, -1, -1), self.textures)

    def update_padded(self, X):
        return TemplateMeshes(X, self._faces.expand(X.shape[0], -1, -1), self.textures)


def _load_obj_verts_faces(path: str):
    verts, faces = [], []
    with open(path) as fh:
        for line in fh:
            if line.startswith("v "):
                verts.append([float(t) for t in line.split()[1:4]])
            elif line.startswith("f "):
                idx = [int(t.split("/")[0]) - 1 for t in line.split()[1:]]
                for k in range(1, len(idx) - 1):
                    faces.append([idx[0], idx[k], idx[k + 1]])
    return torch.tensor(verts, dtype=torch.float32), torch.tensor(faces, dtype=torch.int64)


class Model(nn.Module):
    """model.py:155-196"""

    def save_model(self, out_dir="models/tmp", fname="model_tmp"):
        data = {"state_dict": self.state_dict()}
        data["params"] = self.params
        os.makedirs(out_dir, exist_ok=True)
        torch.save(data, os.path.join(out_dir, fname + ".pth"))

    def freeze(self):
        for param in self.parameters():
            param.requires_grad = False

    @classmethod
    def load(cls, file, device="cuda", opts=None, **kwargs):
        ext = os.path.splitext(file)[-1]
        assert ext == ".pth", f"Generic models can only load from .pth files - received `{ext}` file."
        data = torch.load(file, map_location=device, weights_only=False)
        dont_load_latents = bool(getattr(opts, "dont_load_latents", False))
        if dont_load_latents:
            data["params"]["train_size"] = kwargs.get("train_size", 1)
            data["params"]["val_size"] = kwargs.get("val_size", 1)
            data["params"]["latent_labels"] = kwargs.get("latent_labels", None)
        state_dict = data["state_dict"]
        model = cls(**data["params"], device=device, opts=opts)
        model.configure_template(state_dict, device=device)
        if dont_load_latents:
            skip = [f + ".data" for f in ["shapevec", "shapevec_val", "texvec", "texvec_val", "posevec", "posevec_val",
                                          "reg", "reg_val"]]
            state_dict = {k: v for k, v in state_dict.items() if k not in skip}
        model.load_state_dict(state_dict, strict=False)
        return model


class NeuralDisplacementField(Model):
    """Same constructor signature as model.py:207-216.  Only the architecture FIND trains (train.py:148-157: gaussian
    positional encoding of xyz, width 256, depth 4, three hidden layers per head) has CUDA kernels; anything else
    raises — there is no eager fallback."""

    def __init__(self, sigma=10, depth=4, width=256, encoding="gaussian", dispdepth=3, coldepth=3, normratio=0.1,
                 clamp=None, normclamp=None, niter=6000, input_dim=3, positional_encoding=True,
                 progressive_encoding=False, exclude=0,
                 use_shapevec=False, train_size=None, val_size=None, shapevec_size=256,
                 use_texvec=False, texvec_size=256, use_posevec=False, posevec_size=256,
                 template_mesh_loc=None, device="cuda",
                 restyle_features_per_vertex=False, restyle_cluster_per_vertex=False,
                 use_avg_colour=False, opts=None, latent_labels: dict = None):
        super().__init__()
        unsupported = []
        if width != 256: unsupported.append(f"width={width} (kernels: 256)")
        if depth != 4: unsupported.append(f"depth={depth} (kernels: 4)")
        if dispdepth != 3 or coldepth != 3: unsupported.append(f"dispdepth/coldepth={dispdepth}/{coldepth} (kernels: 3/3)")
        if encoding != "gaussian" or not positional_encoding: unsupported.append("non-gaussian / no positional encoding")
        if progressive_encoding: unsupported.append("progressive_encoding (reference: 'untested', opts.py:49-50)")
        if input_dim != 3: unsupported.append(f"input_dim={input_dim}")
        if unsupported:
            raise NotImplementedError("find_b200.NeuralDisplacementField has sm_100a kernels for FIND's trained "
                                      "architecture only; unsupported: " + "; ".join(unsupported))

        self.params = dict(sigma=sigma, depth=depth, width=width, encoding=encoding, dispdepth=dispdepth,
                           coldepth=coldepth,
                           progressive_encoding=progressive_encoding, positional_encoding=positional_encoding,
                           train_size=train_size, val_size=val_size,
                           use_shapevec=use_shapevec, shapevec_size=shapevec_size, use_texvec=use_texvec,
                           texvec_size=texvec_size, use_posevec=use_posevec, posevec_size=posevec_size,
                           restyle_features_per_vertex=restyle_features_per_vertex,
                           restyle_cluster_per_vertex=restyle_cluster_per_vertex,
                           use_avg_colour=use_avg_colour, latent_labels=latent_labels)
        self.clamp, self.normclamp, self.normratio = clamp, normclamp, normratio

        # same construction ORDER as the reference: the Fourier transform reseeds the global RNG, every nn.Linear
        # created afterwards is therefore initialised exactly like the reference's
        encoding_layers = [FourierFeatureTransform(input_dim, width, sigma, exclude)]
        layers = [nn.Linear(width * 2 + input_dim, width), nn.ReLU()]
        for _ in range(depth):
            layers += [nn.Linear(width, width), nn.ReLU()]
        self.encoder = nn.ModuleList(encoding_layers)
        self.base = nn.ModuleList(layers)

        if template_mesh_loc is not None:
            if use_avg_colour:
                raise NotImplementedError("use_avg_colour with template_mesh_loc needs PyTorch3D texture sampling "
                                          "(model.py:263-278); load avg_col from a checkpoint instead")
            verts, faces = _load_obj_verts_faces(template_mesh_loc)
            verts = verts - torch.mean(verts, dim=0)          # model.py:274-275
            avg_col = torch.zeros(3, dtype=torch.float32)
        else:
            verts = torch.zeros((1, 3), dtype=torch.float32)
            faces = torch.ones((1, 3), dtype=torch.int)
            avg_col = torch.zeros(3, dtype=torch.float32)
        self.template_verts = nn.Parameter(verts.unsqueeze(0).float().to(device), requires_grad=False)
        self.template_faces = nn.Parameter(faces.unsqueeze(0).to(device), requires_grad=False)
        self.avg_col = nn.Parameter(avg_col.to(device), requires_grad=False)
        self.template_mesh = TemplateMeshes(self.template_verts, self.template_faces)

        self.latent_vectors_train, self.latent_vectors_val = [], []

        def table(flag, attr, size, key):
            setattr(self, attr, None)
            setattr(self, attr + "_val", None)
            if not flag:
                return
            labels = None if latent_labels is None else latent_labels.get(key, None)
            labels_val = None if latent_labels is None else latent_labels.get(key + "_val", None)
            tr = LatentVector(train_size, vec_size=size, name=attr + "_train", key=key, labels=labels, device=device)
            va = LatentVector(val_size, vec_size=size, name=attr + "_val", key=key, labels=labels_val, device=device)
            setattr(self, attr, tr)
            setattr(self, attr + "_val", va)
            self.latent_vectors_train.append(tr)
            self.latent_vectors_val.append(va)

        self.shapevec_size, self.use_shapevec = shapevec_size, use_shapevec
        table(use_shapevec, "shapevec", shapevec_size, "shape")
        self.posevec_size, self.use_posevec = posevec_size, use_posevec
        table(use_posevec, "posevec", shapevec_size, "pose")       # reference quirk: pose tables use shapevec_size (model.py:320-323)
        self.texvec_size, self.use_texvec = texvec_size, use_texvec
        table(use_texvec, "texvec", texvec_size, "tex")

        labels = None if latent_labels is None else latent_labels.get("reg", None)
        labels_val = None if latent_labels is None else latent_labels.get("reg_val", None)
        self.reg = LatentVector(train_size, vec_size=9, name="reg_train", device=device, key="reg", labels=labels,
                                init_values=np.array([0] * 6 + [1] * 3))
        self.reg_val = LatentVector(val_size, vec_size=9, name="reg_val", device=device, key="reg", labels=labels_val,
                                    init_values=np.array([0] * 6 + [1] * 3))
        self.latent_vectors_train += [self.reg]
        self.latent_vectors_val += [self.reg_val]

        # reference quirk kept: the SHAPE block of disp_input_size is switched by use_texvec (model.py:352)
        disp_input_size = width + self.shapevec_size * self.use_texvec + self.posevec_size * self.use_posevec
        disp_layers = []
        for i in range(dispdepth):
            disp_layers += [nn.Linear(width if i > 0 else disp_input_size, width), nn.ReLU()]
        disp_layers.append(nn.Linear(width, 3))
        self.mlp_disp = nn.Sequential(*disp_layers)
        col_input_size = width + self.texvec_size * self.use_texvec
        col_layers = []
        for i in range(coldepth):
            col_layers += [nn.Linear(width if i > 0 else col_input_size, width), nn.ReLU()]
        self.use_avg_colour = use_avg_colour
        col_layers.append(nn.Linear(width, 3))
        self.mlp_col = nn.Sequential(*col_layers)

        self.restyle_features_per_vertex = restyle_features_per_vertex
        self.per_vertex_features = None
        if opts is not None and getattr(opts, "template_features_pth", None) is not None:
            raise NotImplementedError("per-vertex template features (restyle branch) are outside the field+Chamfer path")

        main_params = [self.base, self.mlp_disp, self.mlp_col, self.shapevec, self.texvec, self.posevec]
        self.main_params = make_params_list(*main_params)
        self.templ_params = make_params_list(self.template_verts, self.reg)
        self.val_params = make_params_list(self.shapevec_val, self.texvec_val, self.posevec_val)
        self.reg_params = make_params_list(self.reg, self.reg_val)
        self.latent_params = make_params_list(self.shapevec, self.shapevec_val, self.texvec, self.texvec_val,
                                              self.posevec, self.posevec_val)
        self.to(device)
        self.reset_weights()
        self.onnx_mode = False
        self.frozen_mlp = False       # set True for latent-code fitting (train.py:218-224): skips every dW kernel

    # ---- parameters in the C-ABI order ---------------------------------------------------------------------
    def mlp_modules(self) -> List[nn.Linear]:
        """The 13 Linear layers in PARAM_NAMES order (base.0,2,4,6,8 | mlp_disp.0,2,4,6 | mlp_col.0,2,4,6)."""
        return [self.base[i] for i in (0, 2, 4, 6, 8)] + [self.mlp_disp[i] for i in (0, 2, 4, 6)] + \
               [self.mlp_col[i] for i in (0, 2, 4, 6)]

    def mlp_parameters(self) -> List[nn.Parameter]:
        """weights then biases, the flat order of the data-parallel gradient bucket (find_b200/parallel.py)."""
        mods = self.mlp_modules()
        return [m.weight for m in mods] + [m.bias for m in mods]

    def _field(self, pos, shapevec, texvec, posevec, reg, want_verts):
        if self.onnx_mode:
            raise NotImplementedError("onnx_mode is an export path of the reference; not part of the CUDA hot path")
        if pos.dim() != 3:
            raise RuntimeError(f"pos must be [Batch x Num points x 3], got {tuple(pos.shape)}")
        assert pos.shape[-1] == 3, \
            "Expected input to have {} channels (got {} channels)".format(3, pos.shape[-1])   # fourier_feature_transform.py:37-38
        mods = self.mlp_modules()
        dev = pos.device
        lat = [z for z in (shapevec, posevec, texvec) if z is not None]
        if lat and pos.shape[0] != 1 and pos.shape[0] != lat[0].shape[0]:
            raise RuntimeError(f"pos batch {pos.shape[0]} does not match latent batch {lat[0].shape[0]}")
        if not lat and pos.shape[0] != 1:
            # no latents: every "foot" is its own row block with per-foot points
            pass
        disp, col, verts = field_apply(pos, shapevec, posevec, texvec, reg, self.encoder[0].B_on(dev),
                                       self.avg_col if self.use_avg_colour else None,
                                       [m.weight for m in mods], [m.bias for m in mods],
                                       want_col=True, want_verts=want_verts, frozen_mlp=self.frozen_mlp)
        return disp, col, verts

    def forward(self, pos, shapevec=None, texvec=None, posevec=None):
        """model.py:393-453 -> dict(disp [B,V,3], col [B,V,3])."""
        disp, col, _ = self._field(pos, shapevec, texvec, posevec, None, False)
        return {"disp": disp, "col": col}

    def get_meshes(self, shapevec=None, reg=None, texvec=None, posevec=None, no_displacement=False,
                   include_texture=True):
        """model.py:455-504: field on the shared template + registration transform, ONE fused forward."""
        if shapevec is None:
            raise RuntimeError("get_meshes needs shapevec (the reference builds an empty batch without it)")
        N = shapevec.shape[0]
        meshes = self.template_mesh.extend(N)
        disp, col, X = self._field(self.template_verts, shapevec, texvec, posevec, reg, True)
        if not no_displacement:
            meshes = meshes.update_padded(X)
        meshes.textures = col[..., :3]
        res = {"disp": disp, "col": col}
        return dict(meshes=meshes, offsets=disp, verts=X, **res)

    def get_meshes_from_batch(self, batch, is_train=True, no_displacement=False):
        svec, tvec, pvec, reg = [f"{a}_{['val', 'train'][is_train]}" for a in ["shapevec", "texvec", "posevec", "reg"]]
        return self.get_meshes(shapevec=batch.get(svec, None), reg=batch.get(reg, None), texvec=batch.get(tvec, None),
                               posevec=batch.get(pvec, None), include_texture=True, no_displacement=no_displacement)

    def reset_weights(self):
        self.mlp_disp[-1].weight.data.zero_()
        self.mlp_disp[-1].bias.data.zero_()

    def set_template(self, verts: torch.Tensor, faces: Optional[torch.Tensor] = None):
        """Install a template from tensors ([V,3], [F,3]) — what configure_template does from a state_dict."""
        dev = self.template_verts.device
        if faces is None:
            faces = torch.ones((1, 3), dtype=torch.int64)
        self.template_verts = nn.Parameter(verts.reshape(1, -1, 3).float().to(dev), requires_grad=False)
        self.template_faces = nn.Parameter(faces.reshape(1, -1, 3).to(dev), requires_grad=False)
        self.template_mesh = TemplateMeshes(self.template_verts, self.template_faces)
        self.templ_params = make_params_list(self.template_verts, self.reg)

    def configure_template(self, state_dict, device="cuda"):
        verts = state_dict["template_verts"]
        faces = state_dict["template_faces"]
        self.template_verts = nn.Parameter(verts.float().to(device), requires_grad=False)
        self.template_faces = nn.Parameter(faces.to(device), requires_grad=False)
        self.template_mesh = TemplateMeshes(self.template_verts, self.template_faces)
        if "avg_col" in state_dict:
            self.avg_col = nn.Parameter(state_dict["avg_col"].to(device), requires_grad=False)

    def to(self, device):
        out = super().to(device)
        self.template_mesh = TemplateMeshes(self.template_verts, self.template_faces)
        return out


__all__ = ["NeuralDisplacementField", "LatentVector", "FourierFeatureTransform", "TemplateMeshes",
           "sample_latent_vectors", "make_params_list", "PARAM_NAMES"]
