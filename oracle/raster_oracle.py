"""oracle/raster_oracle.py — CPU restatement of the soft-silhouette half of FIND's differentiable renderer.  TEST
INFRASTRUCTURE ONLY: imported by tests/ and __graft_entry__.smoke(); never by the product path.

Follows
  * /root/reference/src/model/renderer.py:122-128 (silhouette RasterizationSettings: blur_radius = log(1/1e-4 - 1) * sigma,
    sigma = 1e-4, faces_per_pixel = 100), :208-245 (``rasterize`` -> ``rasterize_meshes`` with clip_barycentric_coords =
    blur_radius > 0, perspective_correct = cameras.is_perspective()), :309-311 (``SoftSilhouetteShader`` alpha channel);
  * PyTorch3D @ 1706eb82 (requirements_mac_linux.txt:31, NOT on this box), restated from its published algorithm:
    renderer/mesh/rasterize_meshes.{py,cu} (pixel (r, c) has NDC centre (1 - (2c+1)/W, 1 - (2r+1)/H); a face is a candidate
    of a pixel when it is not degenerate (|area| > 1e-8), not entirely behind the camera, the pixel lies inside its
    bounding box grown by sqrt(blur_radius), the interpolated depth pz (perspective-corrected, clipped barycentrics) is
    >= 0, and the pixel is inside the face or its SQUARED distance to the nearest edge is < blur_radius; the K candidates
    of smallest pz are kept, the earlier face on ties), utils/geometry (edge function, point-segment distance) and
    renderer/blending.py::sigmoid_alpha_blend (alpha = 1 - prod_k (1 - sigmoid(-signed_dist_k / sigma))).
    Face clipping against z_clip_value (faces that straddle the camera plane) is not restated: FIND's cameras look at a
    foot from 0.1-0.35 m (renderer.py:146-205) and no face crosses znear / 2.

PARITY UNPINNED against PyTorch3D itself (absent; the reference ships no rendered vectors); pinned by known-answer tests in
tests/test_oracle_raster.py (one triangle: analytic inside / outside distances, K truncation by depth, fp64 autograd vs
finite differences).
"""
from __future__ import annotations

import math

import torch

Z_CLIP = 0.01
K_EPS = 1e-8


def blur_radius_for(sigma: float = 1e-4) -> float:
    """renderer.py:124 — np.log(1. / 1e-4 - 1.) * sigma."""
    return math.log(1.0 / 1e-4 - 1.0) * sigma


def pixel_centres(H: int, W: int, dtype=torch.float32):
    r = torch.arange(H, dtype=dtype)
    c = torch.arange(W, dtype=dtype)
    y = 1.0 - (2.0 * r + 1.0) / H
    x = 1.0 - (2.0 * c + 1.0) / W
    return x, y


def _edge(px, py, ax, ay, bx, by):
    return (px - ax) * (by - ay) - (py - ay) * (bx - ax)


def _seg_dist2(px, py, ax, ay, bx, by):
    bax, bay = bx - ax, by - ay
    l2 = bax * bax + bay * bay
    safe = torch.where(l2 <= K_EPS, torch.ones_like(l2), l2)
    t = ((bax * (px - ax) + bay * (py - ay)) / safe).clamp(0.0, 1.0)
    qx, qy = ax + t * bax, ay + t * bay
    d_seg = (px - qx) ** 2 + (py - qy) ** 2
    d_pt = (px - bx) ** 2 + (py - by) ** 2
    return torch.where(l2 <= K_EPS, d_pt, d_seg)


def silhouette(verts: torch.Tensor, faces: torch.Tensor, H: int, W: int, blur_radius: float | None = None,
               sigma: float = 1e-4, faces_per_pixel: int = 100, perspective_correct: bool = True, return_aux: bool = False,
               rows: tuple | None = None, clip_barycentric_coords: bool | None = None):
    """verts [V,3] in NDC (x left, y up, z depth; any float dtype, autograd flows to x, y), faces [F,3] int64 -> alpha [H,W].
    Dense [H*W, F] evaluation: small cases only."""
    if blur_radius is None:
        blur_radius = blur_radius_for(sigma)
    dt = verts.dtype
    fv = verts[faces]                                                   # [F,3,3]
    x0, y0, z0 = fv[:, 0, 0], fv[:, 0, 1], fv[:, 0, 2]
    x1, y1, z1 = fv[:, 1, 0], fv[:, 1, 1], fv[:, 1, 2]
    x2, y2, z2 = fv[:, 2, 0], fv[:, 2, 1], fv[:, 2, 2]
    xs, ys = pixel_centres(H, W, dt)
    if rows is not None:                                                # only rows [r0, r1) of the H-row image
        ys = ys[rows[0]:rows[1]]
        H = ys.shape[0]
    px = xs[None, :].expand(H, W).reshape(-1, 1)
    py = ys[:, None].expand(H, W).reshape(-1, 1)
    area = _edge(x2, y2, x0, y0, x1, y1)                                # [F]
    blur = math.sqrt(blur_radius)
    xmin, xmax = torch.minimum(torch.minimum(x0, x1), x2) - blur, torch.maximum(torch.maximum(x0, x1), x2) + blur
    ymin, ymax = torch.minimum(torch.minimum(y0, y1), y2) - blur, torch.maximum(torch.maximum(y0, y1), y2) + blur
    # PyTorch3D clips faces against z_clip = znear / 2 = 0.01 (renderer.py:228-236, 258); restated conservatively as a cull of
    # every face with a vertex nearer than that (no face of a foot seen from 0.1-0.35 m gets there) — same rule as csrc/raster.cu
    zmin = torch.minimum(torch.minimum(z0, z1), z2)
    face_ok = (area.abs() > K_EPS) & (zmin >= Z_CLIP)
    in_box = (px >= xmin) & (px <= xmax) & (py >= ymin) & (py <= ymax)
    den = area + K_EPS
    w0 = _edge(px, py, x1, y1, x2, y2) / den
    w1 = _edge(px, py, x2, y2, x0, y0) / den
    w2 = _edge(px, py, x0, y0, x1, y1) / den
    inside = (w0 > 0) & (w1 > 0) & (w2 > 0)
    if perspective_correct:
        t0, t1, t2 = w0 * z1 * z2, z0 * w1 * z2, z0 * z1 * w2
        dn = (t0 + t1 + t2).clamp(min=K_EPS)
        b0, b1, b2 = t0 / dn, t1 / dn, t2 / dn
    else:
        b0, b1, b2 = w0, w1, w2
    if clip_barycentric_coords is None:                                 # renderer.py:219-221
        clip_barycentric_coords = blur_radius > 0.0
    if clip_barycentric_coords:
        c0, c1, c2 = b0.clamp(0, 1), b1.clamp(0, 1), b2.clamp(0, 1)
        cs = (c0 + c1 + c2).clamp(min=1e-5)
        pz = (c0 * z0 + c1 * z1 + c2 * z2) / cs
        u0, u1, u2 = c0 / cs, c1 / cs, c2 / cs
    else:
        pz = b0 * z0 + b1 * z1 + b2 * z2
        u0, u1, u2 = b0, b1, b2
    d2 = torch.minimum(torch.minimum(_seg_dist2(px, py, x0, y0, x1, y1), _seg_dist2(px, py, x0, y0, x2, y2)),
                       _seg_dist2(px, py, x1, y1, x2, y2))
    cand = face_ok[None, :] & in_box & (pz >= 0) & (inside | (d2 < blur_radius))
    # the K candidates of smallest depth, earlier face on ties (stable sort)
    key = torch.where(cand, pz.detach(), torch.full_like(pz, float("inf")))
    order = torch.sort(key, dim=1, stable=True).indices
    rank = torch.empty_like(order)
    rank.scatter_(1, order, torch.arange(order.shape[1]).expand_as(order))
    keep = cand & (rank < faces_per_pixel)
    sd = torch.where(inside, -d2, d2)
    prob = torch.sigmoid(-sd / sigma) * keep.to(dt)
    prod = torch.prod(1.0 - prob, dim=1)
    alpha = (1.0 - prod).reshape(H, W)
    if return_aux:
        nearest = torch.where(cand.any(1), order[:, 0], torch.full_like(order[:, 0], -1))       # pix_to_face[..., 0]
        znear = torch.where(cand.any(1), key.gather(1, order[:, :1])[:, 0], torch.full_like(key[:, 0], -1.0))
        first = order[:, :1]
        hit = cand.any(1)
        bary = torch.stack([u.gather(1, first)[:, 0] for u in (u0, u1, u2)], -1)
        bary = torch.where(hit[:, None], bary, torch.full_like(bary, -1.0))
        dist = torch.where(hit, sd.gather(1, first)[:, 0], torch.full_like(sd[:, 0], -1.0))
        return alpha, dict(count=cand.sum(1).reshape(H, W), keep=keep, prod=prod.reshape(H, W), pix_to_face=nearest.reshape(H, W),
                           zbuf=znear.reshape(H, W), bary=bary.reshape(H, W, 3), dists=dist.reshape(H, W),
                           pair=dict(prob=prob, pz=pz, u=(u0, u1, u2)))
    return alpha


def rasterize_nearest(verts: torch.Tensor, faces: torch.Tensor, H: int, W: int, blur_radius: float = 0.0,
                      perspective_correct: bool = True, clip_barycentric_coords: bool | None = None):
    """rasterize_meshes(faces_per_pixel=1) Fragments of one mesh: pix_to_face [H,W] (-1 = none), zbuf, bary [H,W,3], dists
    (signed squared distance, negative inside) — the image rasteriser of renderer.py:116-120."""
    _, aux = silhouette(verts, faces, H, W, blur_radius=blur_radius, sigma=1.0, faces_per_pixel=1, return_aux=True,
                        perspective_correct=perspective_correct, clip_barycentric_coords=clip_barycentric_coords)
    return aux["pix_to_face"], aux["zbuf"], aux["bary"], aux["dists"]


def silhouette_blocked(verts: torch.Tensor, faces: torch.Tensor, H: int, W: int, rows_per_block: int = 4, **kw) -> torch.Tensor:
    """The same image evaluated in row blocks, each against only the faces whose grown box reaches the block (dropping the
    others cannot change a pixel: they fail its bounding-box test; the face order, hence the tie rule, is preserved; only the
    rounding order of the fp32 product may differ, by ~1e-7).
    Makes full-size images (256 x 256, 20 480 faces) affordable: seconds instead of a 1.3 G-pair dense evaluation."""
    blur_radius = kw.get("blur_radius")
    if blur_radius is None:
        blur_radius = blur_radius_for(kw.get("sigma", 1e-4))
    blur = math.sqrt(blur_radius)
    _, ys = pixel_centres(H, W, verts.dtype)
    fy = verts.detach()[faces][:, :, 1]
    fmin, fmax = fy.min(1).values - blur, fy.max(1).values + blur
    out = []
    for r0 in range(0, H, rows_per_block):
        r1 = min(r0 + rows_per_block, H)
        keep = (fmax >= ys[r1 - 1]) & (fmin <= ys[r0])                   # y decreases with the row index
        out.append(silhouette(verts, faces[keep], H, W, rows=(r0, r1), **kw))
    return torch.cat(out, 0)


def silhouette_batch(verts: torch.Tensor, faces: torch.Tensor, H: int, W: int, **kw) -> torch.Tensor:
    """verts [N,V,3], shared faces -> [N,H,W]."""
    return torch.stack([silhouette(verts[n], faces, H, W, **kw) for n in range(verts.shape[0])])


# ---- cameras (renderer.py:255-258: FoVPerspectiveCameras(R, T, znear=0.02), PyTorch3D defaults fov = 60 deg, zfar = 100) ---
def look_at_rotation(eye: torch.Tensor, at=(0.0, 0.0, 0.0), up=(1.0, 0.0, 0.0)) -> torch.Tensor:
    """PyTorch3D look_at_rotation: rows x, y, z axes of the camera in world coordinates, returned transposed (world -> view is
    X_view = X_world @ R + T)."""
    eye = torch.as_tensor(eye, dtype=torch.float64).reshape(-1, 3)
    at_t = torch.as_tensor(at, dtype=torch.float64).reshape(-1, 3).expand_as(eye)
    up_t = torch.as_tensor(up, dtype=torch.float64).reshape(-1, 3).expand_as(eye)
    z = torch.nn.functional.normalize(at_t - eye, dim=1)
    x = torch.nn.functional.normalize(torch.cross(up_t, z, dim=1), dim=1)
    y = torch.nn.functional.normalize(torch.cross(z, x, dim=1), dim=1)
    return torch.stack([x, y, z], dim=1).transpose(1, 2)


def look_at_view_transform(dist=1.0, elev=0.0, azim=0.0, at=(0.0, 0.0, 0.0), up=(1.0, 0.0, 0.0)):
    """PyTorch3D look_at_view_transform (degrees): camera position from (dist, elev, azim), R as above, T = -R^T eye."""
    d, e, a = (torch.as_tensor(v, dtype=torch.float64).reshape(-1) for v in (dist, elev, azim))
    n = max(d.numel(), e.numel(), a.numel())
    d, e, a = d.expand(n), torch.deg2rad(e.expand(n)), torch.deg2rad(a.expand(n))
    eye = torch.stack([d * torch.cos(e) * torch.sin(a), d * torch.sin(e), d * torch.cos(e) * torch.cos(a)], dim=1)
    eye = eye + torch.as_tensor(at, dtype=torch.float64).reshape(-1, 3)
    R = look_at_rotation(eye, at, up)
    T = -torch.bmm(R.transpose(1, 2), eye[:, :, None])[:, :, 0]
    return R, T


def fov_project(verts_world: torch.Tensor, R: torch.Tensor, T: torch.Tensor, fov_deg: float = 60.0, znear: float = 0.02,
                zfar: float = 100.0) -> torch.Tensor:
    """FoVPerspectiveCameras.transform_points as MeshRasterizer.transform uses it: x, y in NDC, z = VIEW-space depth.
    verts_world [V,3], R [3,3], T [3]."""
    v = verts_world @ R.to(verts_world.dtype) + T.to(verts_world.dtype)
    s = 1.0 / math.tan(math.radians(fov_deg) / 2.0)
    z = v[:, 2]
    return torch.stack([s * v[:, 0] / z, s * v[:, 1] / z, z], dim=1)


# ---- Phong image of the hard fragments (SoftPhongShader as FootRenderer uses it, renderer.py:110-115, 133, 273-275) -------------
def vertex_normals(verts: torch.Tensor, faces: torch.Tensor) -> torch.Tensor:
    """Meshes.verts_normals_packed(): per-corner cross products accumulated on the vertices, then normalised (eps 1e-6)."""
    vf = verts[faces]
    n = torch.zeros_like(verts)
    n = n.index_add(0, faces[:, 1], torch.cross(vf[:, 2] - vf[:, 1], vf[:, 0] - vf[:, 1], dim=1))
    n = n.index_add(0, faces[:, 2], torch.cross(vf[:, 0] - vf[:, 2], vf[:, 1] - vf[:, 2], dim=1))
    n = n.index_add(0, faces[:, 0], torch.cross(vf[:, 1] - vf[:, 0], vf[:, 2] - vf[:, 0], dim=1))
    return torch.nn.functional.normalize(n, eps=1e-6, dim=1)


def phong_image(p2f, bary, zbuf, dists, verts_world, faces, colours, cam_centre, light=(0.0, 0.0, 100.0), ambient=0.5, diffuse=0.3,
                specular=0.2, shininess=64.0, background=(1.0, 1.0, 1.0), sigma=1e-4, gamma=1e-4, znear=0.02, zfar=100.0):
    """p2f [H,W] (-1 = none), bary [H,W,3], zbuf / dists [H,W]; verts_world / colours [V,3] -> rgb [H,W,3].
    phong_shading: colour = (ambient + diffuse * relu(n.l)) * texel + specular * relu(v.r)^shininess [n.l > 0], with the pixel's
    position / normal / texel interpolated by the barycentrics; softmax_rgb_blend with one face per pixel."""
    dt = verts_world.dtype
    hit = p2f >= 0
    f = faces[p2f.clamp(min=0)]                                           # [H,W,3]
    vn = vertex_normals(verts_world, faces)
    interp = lambda a: (bary[..., None] * a[f]).sum(-2)                    # noqa: E731
    p, n, tex = interp(verts_world), interp(vn), interp(colours)
    n = torch.nn.functional.normalize(n, eps=1e-6, dim=-1)
    l = torch.nn.functional.normalize(torch.as_tensor(light, dtype=dt) - p, eps=1e-6, dim=-1)
    cosang = (n * l).sum(-1)
    v = torch.nn.functional.normalize(torch.as_tensor(cam_centre, dtype=dt) - p, eps=1e-6, dim=-1)
    r = -l + 2 * cosang[..., None] * n
    al = torch.relu((v * r).sum(-1)) * (cosang > 0).to(dt)
    colour = (ambient + diffuse * torch.relu(cosang))[..., None] * tex + specular * torch.pow(al, shininess)[..., None]
    eps = 1e-10
    mask = hit.to(dt)
    prob = torch.sigmoid(-dists / sigma) * mask
    z_inv = (zfar - zbuf) / (zfar - znear) * mask
    z_max = z_inv.clamp(min=eps)
    wnum = prob * torch.exp((z_inv - z_max) / gamma)
    delta = torch.exp((eps - z_max) / gamma).clamp(min=eps)
    bg = torch.as_tensor(background, dtype=dt)
    return (wnum[..., None] * colour + delta[..., None] * bg) / (wnum + delta)[..., None]


def render_image(verts_world: torch.Tensor, faces: torch.Tensor, colours: torch.Tensor, R: torch.Tensor, T: torch.Tensor, H: int, W: int,
                 **shading) -> torch.Tensor:
    """One view of FootRenderer.forward(return_images=True) end to end, differentiable by autograd w.r.t. the world vertices and
    the vertex colours exactly where PyTorch3D is: through the barycentrics (position, normal, texel), the vertex normals and the
    signed distance of the blend; pix_to_face and the depth ordering carry no gradient.  The reference for the image BACKWARD
    that the CUDA path does not have yet."""
    ndc = fov_project(verts_world, R, T)
    _, aux = silhouette(ndc, faces, H, W, blur_radius=0.0, sigma=1.0, faces_per_pixel=1, return_aux=True, perspective_correct=True,
                        clip_barycentric_coords=False)
    cam = -(T.to(verts_world.dtype) @ R.to(verts_world.dtype).T)
    return phong_image(aux["pix_to_face"], aux["bary"], aux["zbuf"], aux["dists"], verts_world, faces, colours, cam, **shading)


def feature_blend(verts_ndc: torch.Tensor, faces: torch.Tensor, features: torch.Tensor, H: int, W: int, sigma: float = 1e-4,
                  gamma: float = 1e-4, background=None, znear: float = 1.0, zfar: float = 100.0, faces_per_pixel: int = 100,
                  blur_radius: float | None = None) -> torch.Tensor:
    """FeatureShader (renderer.py:74-105) = the reference's own ``softmax_blend`` (renderer.py:23-72, called with its default
    znear = 1, zfar = 100) over the silhouette Fragments: features [V,C] interpolated with the clipped barycentrics of each of the K
    kept faces, weights prob_k * exp((z_inv_k - max z_inv) / gamma), background weight exp((eps - max z_inv) / gamma) clamped at
    eps.  -> [H,W,C].  Dense: small cases only.  Reference for the feature rendering of the next round."""
    _, aux = silhouette(verts_ndc, faces, H, W, blur_radius=blur_radius, sigma=sigma, faces_per_pixel=faces_per_pixel, return_aux=True)
    dt = verts_ndc.dtype
    keep = aux["keep"].to(dt)
    prob, pz, (u0, u1, u2) = aux["pair"]["prob"], aux["pair"]["pz"], aux["pair"]["u"]
    C = features.shape[1]
    bg = torch.zeros(C, dtype=dt) if background is None else torch.as_tensor(background, dtype=dt)
    eps = 1e-10
    z_inv = (zfar - pz) / (zfar - znear) * keep
    z_max = z_inv.max(dim=1, keepdim=True).values.clamp(min=eps)
    wnum = prob * torch.exp((z_inv - z_max) / gamma)                       # prob already carries the keep mask
    delta = torch.exp((eps - z_max) / gamma).clamp(min=eps)
    denom = wnum.sum(1, keepdim=True) + delta
    ff = features[faces]                                                    # [F,3,C]
    texel = u0[..., None] * ff[None, :, 0] + u1[..., None] * ff[None, :, 1] + u2[..., None] * ff[None, :, 2]   # [P,F,C]
    out = ((wnum[..., None] * texel).sum(1) + delta * bg) / denom
    return out.reshape(H, W, C)


def render_points(points_ndc: torch.Tensor, features: torch.Tensor, H: int, W: int, radius: float = 0.03,
                  points_per_pixel: int = 10) -> torch.Tensor:
    """The keypoint overlay of /root/reference/src/model/renderer.py:139-142, 365-373: PyTorch3D's ``PointsRenderer`` with
    ``PointsRasterizationSettings(radius=0.03, points_per_pixel=10)`` and ``AlphaCompositor`` (no background colour).
    Restated from PyTorch3D @ 1706eb82's published algorithm (renderer/points/rasterize_points.{py,cu}, renderer.py,
    compositing.py; the package is not on this box, so this function is UNPINNED like the rest of the file):

      * a point (x, y, z) — NDC x, y and view-space depth — is a candidate of the pixel with NDC centre (cx, cy) when z >= 0 and
        d2 = (cx - x)^2 + (cy - y)^2 < radius^2;
      * the ``points_per_pixel`` candidates of smallest z are kept, sorted by z (the earlier point on ties);
      * weight w = 1 - d2 / radius^2; front to back: out += T * w * feature, T *= (1 - w), T starting at 1.

    points_ndc [P, 3], features [P, C] -> [H, W, C] (zero where no point lands)."""
    P, C = features.shape
    xs, ys = pixel_centres(H, W, points_ndc.dtype)
    cx = xs[None, :].expand(H, W)
    cy = ys[:, None].expand(H, W)
    out = torch.zeros(H, W, C, dtype=points_ndc.dtype)
    if P == 0:
        return out
    dx = cx[..., None] - points_ndc[:, 0]
    dy = cy[..., None] - points_ndc[:, 1]
    d2 = dx * dx + dy * dy                                            # [H, W, P]
    z = points_ndc[:, 2].expand(H, W, P)
    cand = (z >= 0) & (d2 < radius * radius)
    zkey = torch.where(cand, z, torch.full_like(z, float("inf")))
    order = torch.sort(zkey, dim=-1, stable=True).indices             # stable: the earlier point on ties
    K = min(points_per_pixel, P)
    trans = torch.ones(H, W, dtype=points_ndc.dtype)
    for k in range(K):
        idx = order[..., k]
        ok = torch.gather(cand, -1, idx[..., None])[..., 0]
        w = 1.0 - torch.gather(d2, -1, idx[..., None])[..., 0] / (radius * radius)
        w = torch.where(ok, w, torch.zeros_like(w))
        out = out + (trans * w)[..., None] * features[idx]
        trans = trans * (1.0 - w)
    return out
