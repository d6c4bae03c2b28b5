"""find_b200/renderer.py — the silhouette half of FIND's differentiable renderer on the sm_100a rasteriser of csrc/raster.cu
(first slice).

Mirrors /root/reference/src/model/renderer.py:
  * ``FootRenderer(image_size, device)`` (:107-141), ``sample_views`` (:144-153), ``linspace_views`` (:156-174),
    ``view_from`` (:176-205), ``combine_views`` (:207-208) — same names, argument meaning and return shapes;
  * ``FootRenderer.forward(meshes, R, T, return_mask=True, mask_with_grad=True)`` (:247-311): every mesh is rendered from
    every view by ``FoVPerspectiveCameras(R, T, znear=0.02)`` and the mask is the alpha channel of ``SoftSilhouetteShader``
    over ``rasterize_meshes(faces_per_pixel=100, blur_radius=log(1/1e-4 - 1) * 1e-4)``; returns ``{'mask': [meshes, views, S, S]}``.
``return_images`` / ``return_features`` / ``return_depth`` / ``keypoints`` (Phong shading, feature blending, point
rendering) are NOT built yet and raise NotImplementedError — they never fall back to another renderer.

The camera transform (a 3x3 rotation, a translation and the perspective divide per vertex) is done with torch ops so that
autograd carries the silhouette gradient from NDC back to world-space vertices; the rasterisation, the blend and their
backward are libfindb200.so (``findb200_silhouette_forward/backward``).  There is no CPU path (CPU tensors raise).
"""
from __future__ import annotations

import ctypes
import math
import os

import numpy as np
import torch
from torch import nn

from . import _lib
from .mesh import Meshes, _as_i32_faces

SIGMA = 1e-4                                            # BlendParams(sigma=1e-4), renderer.py:121


def look_at_rotation(eye: torch.Tensor, at=((0.0, 0.0, 0.0),), up=((1.0, 0.0, 0.0),)) -> torch.Tensor:
    """pytorch3d.renderer.look_at_rotation: [n,3] camera positions -> R [n,3,3] with X_view = X_world @ R + T."""
    eye = torch.as_tensor(eye, dtype=torch.float32).reshape(-1, 3)
    at_t = torch.as_tensor(at, dtype=torch.float32).reshape(-1, 3).expand_as(eye)
    up_t = torch.as_tensor(up, dtype=torch.float32).reshape(-1, 3).expand_as(eye)
    z = torch.nn.functional.normalize(at_t - eye, eps=1e-5, dim=1)
    x = torch.nn.functional.normalize(torch.cross(up_t, z, dim=1), eps=1e-5, dim=1)
    y = torch.nn.functional.normalize(torch.cross(z, x, dim=1), eps=1e-5, dim=1)
    degenerate = torch.isclose(x, torch.zeros(()), atol=5e-3).all(dim=1, keepdim=True)
    if degenerate.any():                                # up parallel to the view direction: rebuild x from y and z
        x = torch.where(degenerate, torch.nn.functional.normalize(torch.cross(y, z, dim=1), eps=1e-5, dim=1), x)
    return torch.stack([x, y, z], dim=1).transpose(1, 2)


def look_at_view_transform(dist=1.0, elev=0.0, azim=0.0, at=((0.0, 0.0, 0.0),), up=((0.0, 1.0, 0.0),)):
    """pytorch3d.renderer.look_at_view_transform (degrees) -> R [n,3,3], T [n,3]."""
    d, e, a = (torch.as_tensor(np.asarray(v), dtype=torch.float32).reshape(-1) for v in (dist, elev, azim))
    n = max(d.numel(), e.numel(), a.numel())
    d, e, a = d.expand(n), torch.deg2rad(e.expand(n)), torch.deg2rad(a.expand(n))
    eye = torch.stack([d * torch.cos(e) * torch.sin(a), d * torch.sin(e), d * torch.cos(e) * torch.cos(a)], dim=1)
    eye = eye + torch.as_tensor(at, dtype=torch.float32).reshape(-1, 3)
    R = look_at_rotation(eye, at, up)
    T = -torch.bmm(R.transpose(1, 2), eye[:, :, None])[:, :, 0]
    return R, T


class _Silhouette(torch.autograd.Function):
    @staticmethod
    def forward(ctx, verts_ndc, faces_i32, H, W, blur_radius, sigma, K, perspective):
        lib = _lib.load()
        dev = _lib.require_cuda(verts_ndc)
        v = verts_ndc.contiguous().float()
        N, V, F = v.shape[0], v.shape[1], faces_i32.shape[0]
        alpha = torch.empty(N, H, W, dtype=torch.float32, device=dev)
        prod = torch.empty_like(alpha)
        sel = torch.empty(N, H, W, 2, dtype=torch.int32, device=dev)
        p2f = torch.empty(N, H, W, dtype=torch.int32, device=dev)
        zbuf = torch.empty(N, H, W, dtype=torch.float32, device=dev)
        ws = torch.empty(int(lib.findb200_silhouette_workspace_bytes(N, F)), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_silhouette_forward(_lib.ptr(v), _lib.ptr(faces_i32), N, V, F, H, W, float(blur_radius), float(sigma),
                                                 int(K), int(bool(perspective)), _lib.ptr(alpha), _lib.ptr(prod), _lib.ptr(sel),
                                                 _lib.ptr(p2f), _lib.ptr(zbuf), _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_silhouette_forward")
        # K truncation could not be applied exactly on some tile (> 4096 faces on a 16 x 16 tile AND a pixel with more than K
        # candidates): never accept that silently where a sync is affordable — evaluation / no-grad renders, or when asked to
        if not torch.is_grad_enabled() or os.environ.get("FINDB200_RASTER_CHECK", "0") == "1":
            with torch.cuda.device(dev):
                if lib.findb200_silhouette_status(_lib.ptr(ws), N, F, _lib.stream_ptr(dev)) == 1:
                    raise RuntimeError("rasterize_silhouette: faces_per_pixel truncation was inexact on at least one tile "
                                       "(more than 4096 faces on a 16 x 16 pixel tile); lower the mesh density per tile or raise kListCap")
        if ctx.needs_input_grad[0]:                        # the backward reads the face records only: drop the scratch part
            ws = ws[: int(lib.findb200_silhouette_saved_bytes(N, F))].clone()
        ctx.save_for_backward(faces_i32, prod, sel, ws)
        ctx.cfg = (N, V, F, H, W, float(blur_radius), float(sigma), int(K), int(bool(perspective)))
        ctx.mark_non_differentiable(p2f, zbuf)
        return alpha, p2f, zbuf

    @staticmethod
    def backward(ctx, g_alpha, _g_p2f, _g_zbuf):
        lib = _lib.load()
        faces_i32, prod, sel, ws = ctx.saved_tensors
        N, V, F, H, W, blur_radius, sigma, K, perspective = ctx.cfg
        dev = g_alpha.device
        gv = torch.zeros(N, V, 3, dtype=torch.float32, device=dev)
        g = g_alpha.contiguous().float()
        with torch.cuda.device(dev):
            rc = lib.findb200_silhouette_backward(_lib.ptr(g), _lib.ptr(faces_i32), N, V, F, H, W, blur_radius, sigma, K, perspective,
                                                  _lib.ptr(prod), _lib.ptr(sel), _lib.ptr(gv), _lib.ptr(ws), ws.numel(),
                                                  _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_silhouette_backward")
        return gv, None, None, None, None, None, None, None


def rasterize_silhouette(verts_ndc: torch.Tensor, faces: torch.Tensor, image_size, blur_radius: float = None, sigma: float = SIGMA,
                         faces_per_pixel: int = 100, perspective_correct: bool = True, return_fragments: bool = False):
    """verts_ndc [N,V,3] (x, y NDC; z view depth), faces [F,3] shared -> alpha [N,H,W]; differentiable w.r.t. x, y.
    ``return_fragments`` adds (pix_to_face [N,H,W] int32, zbuf [N,H,W]): the nearest candidate face of every pixel (index within
    the mesh, -1 = none) and its depth — the ``[..., 0]`` slice of PyTorch3D's silhouette Fragments."""
    H, W = (image_size, image_size) if isinstance(image_size, int) else image_size
    if blur_radius is None:
        blur_radius = math.log(1.0 / 1e-4 - 1.0) * sigma
    if faces.dim() != 2 or faces.shape[1] != 3:
        raise ValueError("faces must be [F,3] (one topology shared by the batch)")
    f = faces if faces.dtype == torch.int32 else faces.to(torch.int32)
    alpha, p2f, zbuf = _Silhouette.apply(verts_ndc, f.contiguous(), int(H), int(W), blur_radius, sigma, faces_per_pixel, perspective_correct)
    return (alpha, p2f, zbuf) if return_fragments else alpha


def rasterize_nearest(verts_ndc: torch.Tensor, faces: torch.Tensor, image_size, blur_radius: float = 0.0,
                      perspective_correct: bool = True, clip_barycentric_coords: bool = None):
    """``pytorch3d.renderer.mesh.rasterize_meshes(faces_per_pixel=1)`` for a batch sharing one topology: -> (pix_to_face [N,H,W,1]
    int64, index within the mesh, -1 = none; zbuf [N,H,W,1]; bary_coords [N,H,W,1,3]; dists [N,H,W,1]) — the Fragments of the
    image rasteriser (renderer.py:116-120).  Forward only: the outputs carry no autograd history."""
    lib = _lib.load()
    dev = _lib.require_cuda(verts_ndc)
    H, W = (image_size, image_size) if isinstance(image_size, int) else image_size
    if clip_barycentric_coords is None:
        clip_barycentric_coords = blur_radius > 0.0
    v = verts_ndc.detach().contiguous().float()
    f = (faces if faces.dtype == torch.int32 else faces.to(torch.int32)).contiguous()
    N, V, F = v.shape[0], v.shape[1], f.shape[0]
    p2f = torch.empty(N, H, W, dtype=torch.int32, device=dev)
    zbuf = torch.empty(N, H, W, dtype=torch.float32, device=dev)
    bary = torch.empty(N, H, W, 3, dtype=torch.float32, device=dev)
    dists = torch.empty(N, H, W, dtype=torch.float32, device=dev)
    ws = torch.empty(int(lib.findb200_silhouette_workspace_bytes(N, F)), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = lib.findb200_rasterize_nearest(_lib.ptr(v), _lib.ptr(f), N, V, F, int(H), int(W), float(blur_radius), int(bool(perspective_correct)),
                                            int(bool(clip_barycentric_coords)), _lib.ptr(p2f), _lib.ptr(zbuf), _lib.ptr(bary), _lib.ptr(dists),
                                            _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
    _lib.check(rc, "findb200_rasterize_nearest")
    return p2f.long().unsqueeze(-1), zbuf.unsqueeze(-1), bary.unsqueeze(-2), dists.unsqueeze(-1)


def vertex_normals(verts: torch.Tensor, topo) -> torch.Tensor:
    """``Meshes.verts_normals_padded()``: verts [M,V,3] -> area-weighted unit normals [M,V,3] (forward only)."""
    lib = _lib.load()
    dev = _lib.require_cuda(verts)
    v = verts.detach().contiguous().float()
    topo = topo.to(dev)
    out = torch.empty_like(v)
    with torch.cuda.device(dev):
        rc = lib.findb200_vertex_normals(_lib.ptr(v), v.shape[0], v.shape[1], _lib.ptr(topo.faces), _lib.ptr(topo.vf_ptr),
                                         _lib.ptr(topo.vf_face), _lib.ptr(topo.vf_corner), _lib.ptr(out), _lib.stream_ptr(dev))
    _lib.check(rc, "findb200_vertex_normals")
    return out


class _FeatureRender(torch.autograd.Function):
    """``FeatureShader`` of the reference (renderer.py:74-105: TexturesVertex sampling + its own ``softmax_blend``, :23-72) over the
    silhouette fragments, forward and backward (csrc/raster.cu: findb200_silhouette_forward for the fragments, then
    findb200_feature_blend_forward / _backward).  Differentiable w.r.t. the per-vertex features and the NDC vertices."""

    @staticmethod
    def forward(ctx, verts_ndc, features, faces_i32, H, W, blur_radius, sigma, K, perspective, num_views, gamma, znear, zfar):
        lib = _lib.load()
        dev = _lib.require_cuda(verts_ndc)
        v = verts_ndc.detach().contiguous().float()
        f = features.detach().contiguous().float()
        N, V, F, C = v.shape[0], v.shape[1], faces_i32.shape[0], f.shape[-1]
        alpha = torch.empty(N, H, W, dtype=torch.float32, device=dev)
        prod = torch.empty_like(alpha)
        sel = torch.empty(N, H, W, 2, dtype=torch.int32, device=dev)
        p2f = torch.empty(N, H, W, dtype=torch.int32, device=dev)
        zbuf = torch.empty(N, H, W, dtype=torch.float32, device=dev)
        out = torch.empty(N, H, W, C, dtype=torch.float32, device=dev)
        den = torch.empty(N, H, W, dtype=torch.float32, device=dev)
        ws = torch.empty(int(lib.findb200_silhouette_workspace_bytes(N, F)), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            st = _lib.stream_ptr(dev)
            rc = lib.findb200_silhouette_forward(_lib.ptr(v), _lib.ptr(faces_i32), N, V, F, H, W, float(blur_radius), float(sigma), int(K),
                                                 int(bool(perspective)), _lib.ptr(alpha), _lib.ptr(prod), _lib.ptr(sel), _lib.ptr(p2f),
                                                 _lib.ptr(zbuf), _lib.ptr(ws), ws.numel(), st)
            _lib.check(rc, "findb200_silhouette_forward")
            rc = lib.findb200_feature_blend_forward(_lib.ptr(f), C, int(num_views), _lib.ptr(faces_i32), N, V, F, H, W, float(blur_radius),
                                                    float(sigma), int(K), int(bool(perspective)), float(gamma), float(znear), float(zfar),
                                                    None, _lib.ptr(sel), _lib.ptr(zbuf), _lib.ptr(out), _lib.ptr(den), _lib.ptr(ws),
                                                    ws.numel(), st)
            _lib.check(rc, "findb200_feature_blend_forward")
        ws = ws[: int(lib.findb200_silhouette_saved_bytes(N, F))].clone()
        ctx.save_for_backward(f, faces_i32, sel, zbuf, p2f, out, den, ws)
        ctx.cfg = (N, V, F, C, H, W, float(blur_radius), float(sigma), int(K), int(bool(perspective)), int(num_views), float(gamma),
                   float(znear), float(zfar))
        ctx.mark_non_differentiable(p2f)
        return out, p2f

    @staticmethod
    def backward(ctx, g_out, _g_p2f):
        lib = _lib.load()
        f, faces_i32, sel, zbuf, p2f, out, den, ws = ctx.saved_tensors
        N, V, F, C, H, W, blur_radius, sigma, K, perspective, num_views, gamma, znear, zfar = ctx.cfg
        dev = g_out.device
        g = g_out.contiguous().float()
        g_feat = torch.zeros_like(f)
        g_ndc = torch.zeros(N, V, 3, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_feature_blend_backward(_lib.ptr(g), _lib.ptr(f), C, num_views, _lib.ptr(faces_i32), N, V, F, H, W, blur_radius,
                                                     sigma, K, perspective, gamma, znear, zfar, None, _lib.ptr(sel), _lib.ptr(zbuf),
                                                     _lib.ptr(p2f), _lib.ptr(out), _lib.ptr(den), _lib.ptr(g_feat), _lib.ptr(g_ndc),
                                                     _lib.ptr(ws), ws.numel(), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_feature_blend_backward")
        return g_ndc, g_feat, None, None, None, None, None, None, None, None, None, None, None


def render_points(points_ndc: torch.Tensor, features: torch.Tensor, image_size, radius: float = 0.03, points_per_pixel: int = 10):
    """PyTorch3D's ``PointsRenderer(PointsRasterizer(radius, points_per_pixel), AlphaCompositor())`` as FIND sets it up for its
    keypoint overlay (/root/reference/src/model/renderer.py:139-142): points_ndc [N,P,3] (NDC x, y, view depth), features [N,P,C]
    -> image [N,H,W,C], zero where no point lands.  Forward only (a visualisation)."""
    H, W = (image_size, image_size) if isinstance(image_size, int) else image_size
    dev = _lib.require_cuda(points_ndc)
    pts = points_ndc.detach().to(torch.float32).contiguous()
    f = features.detach().to(dev, torch.float32).contiguous()
    N, P, C = f.shape
    if pts.shape != (N, P, 3):
        raise ValueError(f"points_ndc {tuple(pts.shape)} does not match features {tuple(f.shape)}")
    image = torch.empty((N, int(H), int(W), C), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.load().findb200_points_render(_lib.ptr(pts), _lib.ptr(f), N, P, C, int(H), int(W), float(radius), int(points_per_pixel),
                                                _lib.ptr(image), _lib.stream_ptr(dev))
    _lib.check(rc, "findb200_points_render")
    return image


def render_features(verts_ndc: torch.Tensor, faces: torch.Tensor, features: torch.Tensor, image_size, num_views: int = 1,
                    blur_radius: float = None, sigma: float = SIGMA, gamma: float = 1e-4, faces_per_pixel: int = 100,
                    perspective_correct: bool = True, znear: float = 1.0, zfar: float = 100.0):
    """verts_ndc [N,V,3], features [N // num_views, V, C] -> feature image [N,H,W,C] (+ pix_to_face [N,H,W]); background 0."""
    H, W = (image_size, image_size) if isinstance(image_size, int) else image_size
    if blur_radius is None:
        blur_radius = math.log(1.0 / 1e-4 - 1.0) * sigma
    f = (faces if faces.dtype == torch.int32 else faces.to(torch.int32)).contiguous()
    return _FeatureRender.apply(verts_ndc, features, f, int(H), int(W), blur_radius, sigma, faces_per_pixel, perspective_correct, int(num_views),
                                gamma, znear, zfar)


class _PhongImage(torch.autograd.Function):
    """SoftPhongShader image of the hard fragments with its backward (csrc/raster.cu: findb200_rasterize_nearest,
    findb200_vertex_normals, findb200_phong_shade / _backward, findb200_vertex_normals_backward).  The camera projection's
    backward is torch autograd of ``FootRenderer.project`` (a handful of elementwise ops)."""

    @staticmethod
    def forward(ctx, verts, colours, rdr, topo, R, T):
        lib = _lib.load()
        v = verts.detach().contiguous().float()
        dev = v.device
        M, V = v.shape[0], v.shape[1]
        nv, S = R.shape[0], rdr.image_size
        c = colours.detach().contiguous().float()
        ndc = rdr.project(v, R, T).contiguous()
        p2f, zbuf, bary, dists = rasterize_nearest(ndc, topo.faces, S, blur_radius=0.0, perspective_correct=True)
        normals = vertex_normals(v, topo)
        cam = (-torch.einsum("kj,kij->ki", T.to(torch.float32), R.to(torch.float32))).to(dev).contiguous()   # C = -T R^T
        shading = [*rdr.light_location, *([rdr.light_ambient] * 3), *([rdr.light_diffuse] * 3), *([rdr.light_specular] * 3),
                   *rdr.background_color, rdr.shininess, 1e-4, 1e-4, rdr.znear, rdr.zfar]
        sh = (ctypes.c_float * 20)(*shading)
        image = torch.empty(M * nv, S, S, 3, dtype=torch.float32, device=dev)
        p32 = p2f[..., 0].to(torch.int32).contiguous()
        bary, zbuf, dists = bary.contiguous(), zbuf.contiguous(), dists.contiguous()
        with torch.cuda.device(dev):
            rc = lib.findb200_phong_shade(_lib.ptr(p32), _lib.ptr(bary), _lib.ptr(zbuf), _lib.ptr(dists), M * nv, S, S, nv, _lib.ptr(v),
                                          _lib.ptr(normals), _lib.ptr(c), _lib.ptr(topo.faces), V, _lib.ptr(cam),
                                          ctypes.cast(sh, ctypes.c_void_p), _lib.ptr(image), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_phong_shade")
        ctx.save_for_backward(v, c, ndc, p32, bary, zbuf, dists, normals, cam)
        ctx.misc = (rdr, topo, R, T, shading, M, V, nv, S)
        ctx.mark_non_differentiable(p32)
        return image.reshape(M, nv, S, S, 3), p32

    @staticmethod
    def backward(ctx, g_image, _g_p2f):
        lib = _lib.load()
        v, c, ndc, p32, bary, zbuf, dists, normals, cam = ctx.saved_tensors
        rdr, topo, R, T, shading, M, V, nv, S = ctx.misc
        dev = v.device
        g = g_image.contiguous().float()
        g_verts, g_normals, g_colours = torch.zeros_like(v), torch.zeros_like(v), torch.zeros_like(v)
        g_ndc = torch.zeros_like(ndc)
        scratch = torch.empty_like(v)
        sh = (ctypes.c_float * 20)(*shading)
        with torch.cuda.device(dev):
            st = _lib.stream_ptr(dev)
            rc = lib.findb200_phong_shade_backward(_lib.ptr(g), _lib.ptr(p32), _lib.ptr(bary), _lib.ptr(zbuf), _lib.ptr(dists), M * nv, S, S, nv,
                                                   _lib.ptr(ndc), _lib.ptr(v), _lib.ptr(normals), _lib.ptr(c), _lib.ptr(topo.faces), V,
                                                   _lib.ptr(cam), ctypes.cast(sh, ctypes.c_void_p), _lib.ptr(g_verts), _lib.ptr(g_normals),
                                                   _lib.ptr(g_colours), _lib.ptr(g_ndc), st)
            _lib.check(rc, "findb200_phong_shade_backward")
            rc = lib.findb200_vertex_normals_backward(_lib.ptr(v), M, V, _lib.ptr(topo.faces), _lib.ptr(topo.vf_ptr), _lib.ptr(topo.vf_face),
                                                      _lib.ptr(topo.vf_corner), _lib.ptr(g_normals), _lib.ptr(scratch), _lib.ptr(g_verts), st)
            _lib.check(rc, "findb200_vertex_normals_backward")
        with torch.enable_grad():                                     # backward of the camera projection
            vv = v.detach().requires_grad_(True)
            (g_proj,) = torch.autograd.grad(rdr.project(vv, R, T), vv, g_ndc)
        g_verts = g_verts + g_proj
        return (g_verts if ctx.needs_input_grad[0] else None), (g_colours if ctx.needs_input_grad[1] else None), None, None, None, None


class FootRenderer(nn.Module):
    def __init__(self, image_size, device="cuda", background_color=(1.0, 1.0, 1.0), bin_size=None, z_clip_value=None,
                 max_faces_per_bin=None):
        super().__init__()
        self.image_size = image_size
        self.device = torch.device(device)
        self.fov, self.znear, self.zfar = 60.0, 0.02, 100.0          # FoVPerspectiveCameras defaults + renderer.py:258
        self.sigma = SIGMA
        self.blur_radius = float(np.log(1.0 / 1e-4 - 1.0) * self.sigma)   # renderer.py:124
        self.faces_per_pixel = 100                                   # renderer.py:125
        self.background_color = tuple(float(c) for c in background_color)
        self.light_location = (0.0, 0.0, 100.0)                      # PointLights(location=[[0, 0, 100]]), renderer.py:111
        self.light_ambient, self.light_diffuse, self.light_specular, self.shininess = 0.5, 0.3, 0.2, 64.0   # PyTorch3D defaults
        self.points_radius, self.points_per_pixel = 0.03, 10         # PointsRasterizationSettings of the keypoint overlay, renderer.py:139

    # ---- view helpers (host-side numpy / torch, as in the reference) ---------------------------------------------------
    def sample_views(self, nviews=1, dist_mean=0.25, dist_std=0.05, elev_min=-90, elev_max=90, azim_min=0, azim_max=360,
                     seed: int = None):
        if seed:
            np.random.seed(seed)
        distances = np.random.normal(dist_mean, dist_std, (nviews))
        elev = np.random.uniform(elev_min, elev_max, (nviews))
        azim = np.random.uniform(azim_min, azim_max, (nviews))
        return look_at_view_transform(dist=distances, elev=elev, azim=azim, up=((1, 0, 0),))

    def linspace_views(self, nviews=1, dist=0.3, dist_min=None, dist_max=None, elev_min=None, elev_max=None, azim_min=None,
                       azim_max=None, at=((0, 0, 0),)):
        if dist_min is not None:
            dist = np.linspace(dist_min, dist_max, nviews)
        elev = 0 if elev_min is None else np.linspace(elev_min, elev_max, nviews)
        azim = 0 if azim_min is None else np.linspace(azim_min, azim_max, nviews)
        return look_at_view_transform(dist=dist, elev=elev, azim=azim, up=((1, 0, 0),), at=at)

    def view_from(self, view_kw="topdown"):
        table = {"topdown": (0.3, 0, 0, (0, 0, 0)), "side1": (0.35, 90, 0, (0, 0, 0)), "side2": (0.35, -90, 180, (0, 0, 0)),
                 "toes": (0.1, 0, 0, (0.1, 0, 0)), "45": (0.35, -45, 0, (0, 0, 0)), "60": (0.35, -60, 0, (0, 0, 0))}
        if isinstance(view_kw, str):
            view_kw = [view_kw]
        Rs, Ts = [], []
        for v in view_kw:
            assert v in table, f"View description `{view_kw}` not understood"
            dist, elev, azim, point = table[v]
            R, T = look_at_view_transform(dist=dist, elev=elev, azim=azim, up=((1, 0, 0),), at=(point,))
            Rs.append(R[0]); Ts.append(T[0])
        return torch.stack(Rs), torch.stack(Ts)

    def combine_views(self, R1, T1, R2, T2):
        return torch.cat([R1, R2], dim=0), torch.cat([T1, T2], dim=0)

    # ---- camera -------------------------------------------------------------------------------------------------------
    def project(self, verts: torch.Tensor, R: torch.Tensor, T: torch.Tensor) -> torch.Tensor:
        """verts [M,V,3] world, R [views,3,3], T [views,3] -> [M*views,V,3]: NDC x, y and view-space depth, mesh-major like
        ``input_meshes.extend(num_views)`` with ``torch.cat([R]*num_meshes)`` (renderer.py:253-257)."""
        M, V = verts.shape[0], verts.shape[1]
        nv = R.shape[0]
        R = R.to(verts.device, verts.dtype)
        T = T.to(verts.device, verts.dtype)
        view = torch.einsum("mvi,kij->mkvj", verts, R) + T[None, :, None, :]
        s = 1.0 / math.tan(math.radians(self.fov) / 2.0)
        z = view[..., 2]
        ndc = torch.stack([s * view[..., 0] / z, s * view[..., 1] / z, z], dim=-1)
        return ndc.reshape(M * nv, V, 3)

    def forward(self, input_meshes: Meshes, R, T, return_images=True, return_depth=False, return_mask=False, mask_with_grad=True,
                mask_out_faces=False, masked_faces=None, keypoints=None, keypoints_blend=False, lights=None,
                return_mask_out_masks=False, return_features=False, features=None) -> dict:
        if return_depth:
            raise NotImplementedError("find_b200.renderer: return_depth cannot run in the reference either (renderer.py:296-301 reads "
                                      "`fragments.zbuf` of a dict); the nearest depth is rasterize_nearest(...)[1] / rasterize_silhouette(..., return_fragments=True)[2]")
        if keypoints is not None:
            # renderer.py:365-377: the scene without the overlay, then red discs at the keypoints (PointsRenderer, radius 0.03,
            # 10 points per pixel, AlphaCompositor) and, with keypoints_blend, the discs pasted over the image
            out = self.forward(input_meshes, R, T, return_images=return_images, return_mask=return_mask, mask_with_grad=mask_with_grad,
                               mask_out_faces=mask_out_faces, masked_faces=masked_faces, lights=lights,
                               return_mask_out_masks=return_mask_out_masks, return_features=return_features, features=features)
            num_views, num_meshes = R.shape[0], len(input_meshes)
            if keypoints.dim() != 3 or keypoints.shape[0] != num_meshes or keypoints.shape[-1] != 3:
                # the reference's own stacking over the views (renderer.py:367) is computed and discarded, so it only runs with one
                # keypoint cloud per camera; here a cloud per MESH is drawn in every view of that mesh, which is what that line is for
                raise ValueError(f"keypoints must be [num_meshes = {num_meshes}, K, 3], got {tuple(keypoints.shape)}")
            dev = _lib.require_cuda(input_meshes.verts_padded())
            kp = keypoints.detach().to(dev, torch.float32)
            feats = torch.zeros_like(kp)
            feats[..., 0] = 1.0
            pcl = render_points(self.project(kp, R, T), feats.repeat_interleave(num_views, dim=0), self.image_size,
                                radius=self.points_radius, points_per_pixel=self.points_per_pixel)
            S = self.image_size
            pcl = pcl.reshape(num_meshes, num_views, S, S, 3)
            out["keypoints"] = pcl
            if keypoints_blend:
                if not return_images:
                    raise ValueError("keypoints_blend pastes the keypoints over out['image']: it needs return_images=True")
                pcl_mask = torch.any(pcl > 0, dim=-1, keepdim=True)
                out["keypoints_blend"] = (~pcl_mask) * out["image"].clone() + pcl_mask * pcl
            return out
        if return_features and features is None:
            raise AssertionError("Requires features to be passed for rendering")        # renderer.py:278
        if mask_out_faces and masked_faces is None:
            raise NotImplementedError("find_b200.renderer: mask_out_faces needs masked_faces (the TexturesUV u = v = 0 convention of "
                                      "renderer.py:337-345 is not built)")
        """Same defaults as the reference (renderer.py:247-253: ``return_images=True, return_mask=False``), so the call sites that
        rely on them keep their meaning: ``renderer(mesh, R, T)['image']`` (mesh_turntable.py:57, model.py:1074,1089),
        ``return_mask=False`` (eval_2d.py:89-90)."""
        if not return_mask and not return_images and not return_features:
            return {}
        if return_images:
            images, img_p2f = self._render_images(input_meshes, R, T)
            if (not return_mask or not mask_with_grad) and not return_features:
                out = {}
                if return_mask:
                    # renderer.py:311-312: hard mask from the image render — any channel below the white background
                    M_, nv_, S_ = images.shape[0], images.shape[1], images.shape[2]
                    out["mask"] = torch.any(images < 1, dim=-1).reshape(M_, nv_, S_, S_).float()
                if mask_out_faces:
                    images, mpix = self._mask_images(images, img_p2f, masked_faces)
                    if return_mask:
                        for n, m in enumerate(mpix):
                            out["mask"][n, m] = 0.0
                    if return_mask_out_masks:
                        out["mask_out_masks"] = torch.stack(mpix).reshape(images.shape[:4])
                out["image"] = images
                return out
        verts = input_meshes.verts_padded()
        faces, shared = _as_i32_faces(input_meshes)
        if not shared:
            raise NotImplementedError("find_b200.renderer: the meshes of one call must share their topology (FIND's template)")
        num_views, num_meshes = R.shape[0], len(input_meshes)
        S = self.image_size
        ndc = self.project(verts, R, T)
        out = {}
        mask, p2f = None, None
        if return_mask:
            if mask_with_grad or not return_images:
                alpha, p2f, _ = rasterize_silhouette(ndc, faces, S, self.blur_radius, self.sigma, self.faces_per_pixel, True, return_fragments=True)
                mask = alpha.reshape(num_meshes, num_views, S, S)
            else:                                                  # renderer.py:311-312: hard mask from the image render
                mask = torch.any(images < 1, dim=-1).reshape(num_meshes, num_views, S, S).float()
        feat_renders = None
        if return_features:
            # renderer.py:277-284: TexturesVertex(features) on the SILHOUETTE fragments, background 0, default BlendParams
            feats = features if features.dim() == 3 else features.unsqueeze(0).expand(num_meshes, -1, -1)
            fr, fp2f = render_features(ndc, faces, feats.to(ndc.device), S, num_views=num_views, blur_radius=self.blur_radius, sigma=self.sigma,
                                       gamma=1e-4, faces_per_pixel=self.faces_per_pixel)
            feat_renders = fr.reshape(num_meshes, num_views, S, S, -1)
            if p2f is None:
                p2f = fp2f
        if return_images:
            out["image"] = images
        if mask_out_faces:
            # renderer.py:318-365: the closest face of every pixel; pixels whose closest face is in masked_faces are zeroed
            p2f = (img_p2f if return_images else p2f).reshape(num_meshes, num_views, S, S)   # the image fragments when rendered
            masks = []
            if mask is not None:
                mask = mask.clone()
            if feat_renders is not None:
                feat_renders = feat_renders.clone()
            for n in range(num_meshes):
                mf = masked_faces[n] if isinstance(masked_faces, list) else masked_faces
                mpix = torch.isin(p2f[n], mf.to(p2f.device, p2f.dtype))
                masks.append(mpix)
                if mask is not None:
                    mask[n, mpix] = 0.0
                if feat_renders is not None:
                    feat_renders[n, mpix] = 0.0
                if return_images:
                    if out["image"].requires_grad and n == 0:
                        out["image"] = out["image"].clone()
                    out["image"][n, mpix] = 1.0
            if return_mask_out_masks:
                out["mask_out_masks"] = torch.stack(masks).reshape(num_meshes, num_views, S, S)
        if mask is not None:
            out["mask"] = mask
        if feat_renders is not None:
            out["features"] = feat_renders
        return out


    # ---- Phong image -----------------------------------------------------------------------------------
    def _colours(self, meshes: Meshes, required=True):
        tex = meshes.textures
        if tex is None:
            if required:
                raise ValueError("Meshes do not contain textures (TexturesVertex features [N,V,3] expected).")
            return None
        c = tex.verts_features_padded() if hasattr(tex, "verts_features_padded") else tex
        return c

    def _render_images(self, meshes: Meshes, R, T):
        """-> (image [M,views,S,S,3], pix_to_face int32 [M*views,S,S]).  Differentiable w.r.t. the mesh vertices and the vertex
        colours (``_PhongImage``); under ``torch.no_grad()`` only the forward kernels run."""
        from .mesh import _topology_of
        verts = meshes.verts_padded()
        dev = _lib.require_cuda(verts)
        topo = _topology_of(meshes).to(dev)
        M = verts.shape[0]
        colours = self._colours(meshes).to(dev)
        if colours.dim() == 2:
            colours = colours.unsqueeze(0).expand(M, -1, -1)
        image, p32 = _PhongImage.apply(verts.float(), colours.float(), self, topo, R, T)
        return image, p32

    def _mask_images(self, images, p2f, masked_faces):
        if images.requires_grad:
            images = images.clone()                                   # in-place masking below (renderer.py:354-356)
        M, nv, S = images.shape[0], images.shape[1], images.shape[2]
        p2f = p2f.reshape(M, nv, S, S)
        masks = []
        for n in range(M):
            mf = masked_faces[n] if isinstance(masked_faces, list) else masked_faces
            mpix = torch.isin(p2f[n], mf.to(p2f.device, p2f.dtype))
            masks.append(mpix)
            images[n, mpix] = 1.0
        return images, masks
