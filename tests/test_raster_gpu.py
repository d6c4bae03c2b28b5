"""Soft-silhouette rasteriser (csrc/raster.cu through the C ABI) against oracle/raster_oracle.py, plus the host-side camera
helpers of find_b200.renderer (CPU).  Tolerances: alpha is fp32 with sigmoid / product rounding -> 2e-5 absolute (a candidate that
flips at the blur boundary carries p = 1e-4, hence the 3e-4 ceiling on the maximum); gradients rel-L2 1e-3 (fp32 atomics)."""
import math

import pytest
import torch

from find_b200 import synthetic as syn
from find_b200.mesh import Meshes
from find_b200.renderer import FootRenderer, look_at_view_transform, rasterize_silhouette
from oracle import raster_oracle as ro


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def test_camera_helpers_match_the_oracle_cpu():
    r = FootRenderer(64, device="cpu")
    for kw in (dict(dist=0.3, elev=0.0, azim=0.0), dict(dist=0.35, elev=-45.0, azim=0.0), dict(dist=[0.2, 0.3], elev=[10.0, 80.0], azim=[5.0, 200.0])):
        R, T = look_at_view_transform(up=((1, 0, 0),), **kw)
        Ro, To = ro.look_at_view_transform(up=(1.0, 0.0, 0.0), **kw)
        assert torch.allclose(R.double(), Ro, atol=1e-6) and torch.allclose(T.double(), To, atol=1e-6)
    R, T = r.view_from(["topdown", "side1", "side2", "toes", "45", "60"])
    assert R.shape == (6, 3, 3) and T.shape == (6, 3)
    R2, T2 = r.combine_views(R, T, *r.linspace_views(3, dist=0.3, azim_min=0, azim_max=90))
    assert R2.shape == (9, 3, 3) and T2.shape == (9, 3)
    Rs, Ts = r.sample_views(4, seed=3)
    assert Rs.shape == (4, 3, 3) and torch.allclose(torch.linalg.det(Rs), torch.ones(4), atol=1e-5)
    tv, _ = syn.make_template(162)
    ndc = r.project(tv[None], R[:2], T[:2])
    for k in range(2):
        assert torch.allclose(ndc[k], ro.fov_project(tv, R[k], T[k]), atol=1e-6)
    m = Meshes(tv[None], torch.zeros(1, 3, dtype=torch.int64))
    with pytest.raises(AssertionError):          # renderer.py:278: features must come with return_features
        r(m, R, T, return_features=True)
    with pytest.raises(NotImplementedError):     # renderer.py:296-301 cannot run in the reference either; refused with the reason
        r(m, R, T, return_depth=True)


def _scene(V, views, B=1, seed=0):
    dev = torch.device("cuda")
    tv, tf = syn.make_template(V)
    g = torch.Generator().manual_seed(seed)
    verts = tv[None] + 0.002 * torch.randn(B, tv.shape[0], 3, generator=g)
    r = FootRenderer(64, device=dev)
    R, T = r.view_from(views)
    return dev, verts, tf, r, R, T


@pytest.mark.gpu
@pytest.mark.parametrize("V,size", [(162, (64, 64)), (642, (96, 96)), (642, (50, 70))])
def test_silhouette_matches_oracle(V, size):
    dev, verts, tf, r, R, T = _scene(V, ["topdown", "45"])
    ndc = r.project(verts, R, T)                                          # [2, V, 3] on CPU
    H, W = size
    got = rasterize_silhouette(ndc.to(dev), tf.to(dev), (H, W))
    want = ro.silhouette_batch(ndc, tf, H, W)
    err = (got.cpu() - want).abs()
    assert float(err.max()) < 3e-4 and float(err.mean()) < 2e-6
    assert float(want.max()) > 0.99 and float(want.min()) == 0.0 and 0.02 < float(want.mean()) < 0.9


@pytest.mark.gpu
def test_silhouette_gradient_matches_oracle_autograd():
    dev, verts, tf, r, R, T = _scene(162, ["topdown", "side1"], seed=1)
    H = W = 48
    w = torch.rand(2, H, W, generator=torch.Generator().manual_seed(5))
    ndc = r.project(verts, R, T)
    a = ndc.clone().requires_grad_(True)
    (ro.silhouette_batch(a, tf, H, W) * w).sum().backward()
    b = ndc.to(dev).requires_grad_(True)
    (rasterize_silhouette(b, tf.to(dev), (H, W)) * w.to(dev)).sum().backward()
    assert float(a.grad.abs().max()) > 0
    assert rel_l2(b.grad.cpu(), a.grad) < 1e-3
    assert float(b.grad[..., 2].abs().max()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("K", [1, 5, 100])
def test_k_truncation_by_depth_with_ties(K):
    """150 overlapping triangles at 50 distinct depths (three share each depth): more candidates than K on most pixels."""
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(2)
    base = torch.tensor([[-0.6, -0.5], [0.6, -0.4], [-0.1, 0.7]])
    vs, fs = [], []
    for i in range(150):
        xy = base + 0.05 * torch.randn(3, 2, generator=g)
        z = torch.full((3, 1), 0.5 + 0.01 * (i % 50))
        vs.append(torch.cat([xy, z], 1)); fs.append([3 * i, 3 * i + 1, 3 * i + 2])
    v = torch.cat(vs)[None]; f = torch.tensor(fs)
    kw = dict(blur_radius=0.01, sigma=0.3, faces_per_pixel=K)
    want, aux = ro.silhouette(v[0], f, 32, 32, return_aux=True, **kw)
    assert int(aux["count"].max()) > 100
    vg = v.to(dev).requires_grad_(True)
    got = rasterize_silhouette(vg, f.to(dev), 32, **kw)
    assert float((got[0].detach().cpu() - want).abs().max()) < 2e-5
    w = torch.rand(32, 32, generator=g)
    (got[0] * w.to(dev)).sum().backward()
    vo = v[0].clone().requires_grad_(True)
    (ro.silhouette(vo, f, 32, 32, **kw) * w).sum().backward()
    assert rel_l2(vg.grad[0].cpu(), vo.grad) < 1e-3


@pytest.mark.gpu
def test_empty_and_degenerate_inputs():
    dev = torch.device("cuda")
    v = torch.zeros(2, 5, 3, device=dev)
    a = rasterize_silhouette(v, torch.zeros(0, 3, dtype=torch.int64, device=dev), 20)
    assert a.shape == (2, 20, 20) and float(a.abs().max()) == 0.0
    a = rasterize_silhouette(v, torch.tensor([[0, 1, 2]], device=dev), 20)               # zero-area face
    assert float(a.abs().max()) == 0.0
    v = torch.tensor([[[-0.5, -0.5, -1.0], [0.5, -0.5, -1.0], [0.0, 0.5, -1.0]]], device=dev)  # behind the camera
    assert float(rasterize_silhouette(v, torch.tensor([[0, 1, 2]], device=dev), 20).abs().max()) == 0.0
    with pytest.raises(RuntimeError):
        rasterize_silhouette(torch.zeros(1, 3, 3), torch.tensor([[0, 1, 2]]), 8)           # CPU tensors: no fallback


@pytest.mark.gpu
def test_foot_renderer_mask_full_size_properties_and_world_gradient():
    """FootRenderer.forward at the reference's scale (10 242-vertex template, 256 x 256, 3 views x 2 feet): shapes, range,
    agreement of the soft mask with the hard coverage away from the boundary, and a gradient that reaches world vertices."""
    dev, verts, tf, _, _, _ = _scene(10242, ["topdown"], B=2, seed=3)
    r = FootRenderer(256, device=dev)
    R, T = r.view_from(["topdown", "side1", "45"])
    vw = verts.to(dev).requires_grad_(True)
    out = r(Meshes(vw, tf.to(dev)), R, T, return_images=False, return_mask=True)
    m = out["mask"]; md = m.detach()
    assert m.shape == (2, 3, 256, 256) and float(md.min()) >= 0.0 and float(md.max()) <= 1.0
    frac = (m > 0.5).float().mean(dim=(2, 3))
    assert float(frac.min()) > 0.03 and float(frac.max()) < 0.8
    soft = ((m > 1e-3) & (m < 1 - 1e-3)).float().mean()
    assert 0.0 < float(soft) < 0.1                                        # only a thin boundary band is soft
    target = torch.roll(m.detach(), shifts=4, dims=3)
    loss = ((m - target) ** 2).mean()
    loss.backward()
    g = vw.grad
    assert torch.isfinite(g).all() and float(g.abs().max()) > 0
    # one full-size view against the oracle, evaluated in row blocks
    ndc = r.project(verts[:1], R[2:3], T[2:3])
    got = rasterize_silhouette(ndc.to(dev), tf.to(dev), 256)[0].cpu()
    want = ro.silhouette_blocked(ndc[0], tf, 256, 256)
    err = (got - want).abs()
    assert float(err.max()) < 3e-4 and float(err.mean()) < 2e-6


@pytest.mark.gpu
def test_k_truncation_re_evaluation_path(monkeypatch):
    """FINDB200_SIL_CAP below the candidate count forces the block-wide search that re-evaluates the tile's faces."""
    monkeypatch.setenv("FINDB200_SIL_CAP", "40")
    test_k_truncation_by_depth_with_ties(100)
    test_k_truncation_by_depth_with_ties(5)


@pytest.mark.gpu
def test_nearest_face_fragments_and_mask_out_faces():
    dev, verts, tf, r, R, T = _scene(642, ["topdown", "side1"], B=2, seed=4)
    S = 64
    ndc = r.project(verts, R, T)                                          # [4, V, 3]
    alpha, p2f, zbuf = rasterize_silhouette(ndc.to(dev), tf.to(dev), S, return_fragments=True)
    for k in range(ndc.shape[0]):
        _, aux = ro.silhouette(ndc[k], tf, S, S, return_aux=True)
        same = (p2f[k].cpu().long() == aux["pix_to_face"])
        assert float(same.float().mean()) > 0.999                        # a depth tie between two faces may resolve differently
        assert torch.equal(p2f[k].cpu() < 0, aux["pix_to_face"] < 0)
        assert float((zbuf[k].cpu() - aux["zbuf"].float())[same].abs().max()) < 1e-6
    # FootRenderer: pixels whose closest face is masked are zeroed, the rest is untouched, gradients still flow
    masked = torch.arange(0, tf.shape[0], 3, device=dev)
    vw = verts.to(dev).requires_grad_(True)
    mesh = Meshes(vw, tf.to(dev))
    mk = dict(return_images=False, return_mask=True)     # reference defaults are images on / mask off (renderer.py:247-249)
    plain = r(mesh, R, T, **mk)["mask"]
    out = r(mesh, R, T, mask_out_faces=True, masked_faces=masked, return_mask_out_masks=True, **mk)
    mm = out["mask_out_masks"]
    assert mm.shape == (2, 2, S, S) and mm.dtype == torch.bool and 0.05 < float(mm.float().mean()) < 0.5
    assert float(out["mask"][mm].abs().max()) == 0.0 and torch.equal(out["mask"][~mm], plain[~mm])
    expect = torch.isin(p2f.reshape(2, 2, S, S), masked.to(torch.int32))
    assert torch.equal(mm, expect)
    out["mask"].sum().backward()
    assert torch.isfinite(vw.grad).all() and float(vw.grad.abs().max()) > 0
    lst = r(mesh, R, T, mask_out_faces=True, masked_faces=[masked, masked[:5]], return_mask_out_masks=True, **mk)
    assert torch.equal(lst["mask_out_masks"][0], mm[0]) and int(lst["mask_out_masks"][1].sum()) < int(mm[1].sum())


@pytest.mark.gpu
@pytest.mark.parametrize("perspective", [True, False])
def test_hard_rasterisation_fragments_match_oracle(perspective):
    """faces_per_pixel = 1, blur 0 (the image rasteriser): nearest containing face, depth, barycentrics, signed distance."""
    from find_b200.renderer import rasterize_nearest
    dev, verts, tf, r, R, T = _scene(642, ["topdown", "60"], seed=6)
    ndc = r.project(verts, R, T)
    p2f, zbuf, bary, dists = rasterize_nearest(ndc.to(dev), tf.to(dev), (72, 56), perspective_correct=perspective)
    assert p2f.shape == (2, 72, 56, 1) and bary.shape == (2, 72, 56, 1, 3) and p2f.dtype == torch.int64
    for k in range(2):
        wp, wz, wb, wd = ro.rasterize_nearest(ndc[k], tf, 72, 56, perspective_correct=perspective)
        gp = p2f[k, ..., 0].cpu()
        same = gp == wp
        assert float(same.float().mean()) > 0.999 and torch.equal(gp < 0, wp < 0)
        hit = same & (wp >= 0)
        assert 0.1 < float(hit.float().mean()) < 0.9
        assert float((zbuf[k, ..., 0].cpu() - wz)[same].abs().max()) < 1e-6
        assert float((bary[k, ..., 0, :].cpu() - wb)[same].abs().max()) < 2e-5
        assert float((dists[k, ..., 0].cpu() - wd)[same].abs().max()) < 1e-7
        b = bary[k, ..., 0, :].cpu()[hit]
        assert float((b.sum(-1) - 1).abs().max()) < 1e-4 and float(b.min()) > 0
        assert float(dists[k, ..., 0].cpu()[hit].max()) <= 0.0               # inside: negative squared distance


@pytest.mark.gpu
def test_phong_image_forward_matches_oracle():
    """FootRenderer.forward(return_images=True) under no_grad (ground-truth / evaluation renders): vertex normals, Phong shading
    and softmax_rgb_blend on the hard fragments, against the oracle evaluated on the SAME fragments; 1e-4 absolute (powf(., 64))."""
    from find_b200.mesh import MeshTopology
    from find_b200.renderer import rasterize_nearest, vertex_normals
    dev, verts, tf, _, _, _ = _scene(642, ["topdown"], B=2, seed=8)
    S = 80
    r = FootRenderer(S, device=dev)
    R, T = r.view_from(["topdown", "side1", "45"])
    g = torch.Generator().manual_seed(9)
    colours = torch.rand(2, verts.shape[1], 3, generator=g)
    topo = MeshTopology(tf.to(dev), verts.shape[1])
    vn = vertex_normals(verts.to(dev), topo).cpu()
    for m in range(2):
        assert float((vn[m] - ro.vertex_normals(verts[m], tf)).abs().max()) < 1e-5
    mesh = Meshes(verts.to(dev), tf.to(dev), textures=colours.to(dev))
    with torch.no_grad():
        out = r(mesh, R, T, return_images=True, return_mask=True)
    img = out["image"].cpu()
    assert img.shape == (2, 3, S, S, 3) and out["mask"].shape == (2, 3, S, S)
    ndc = r.project(verts, R, T)
    p2f, zbuf, bary, dists = (t.cpu() for t in rasterize_nearest(ndc.to(dev), tf.to(dev), S))
    cam = -torch.einsum("kj,kij->ki", T, R)
    for m in range(2):
        for k in range(3):
            i = m * 3 + k
            want = ro.phong_image(p2f[i, ..., 0], bary[i, ..., 0, :], zbuf[i, ..., 0], dists[i, ..., 0], verts[m], tf, colours[m], cam[k])
            err = (img[m, k] - want).abs()
            assert float(err.max()) < 1e-4
            empty = p2f[i, ..., 0] < 0
            assert float((img[m, k][empty] - 1.0).abs().max()) == 0.0 and 0.1 < float((~empty).float().mean()) < 0.9
            assert float(img[m, k][~empty].std()) > 0.02                  # shaded, not flat
    # images only; masked faces paint the image white; gradients are refused (forward only)
    with torch.no_grad():
        only = r(mesh, R, T, return_images=True, return_mask=False)
        masked = r(mesh, R, T, return_images=True, return_mask=True, mask_out_faces=True, masked_faces=torch.arange(0, tf.shape[0], 2, device=dev),
                   return_mask_out_masks=True)
    assert torch.equal(only["image"].cpu(), img)
    # reference defaults: renderer(mesh, R, T) returns the IMAGE only (mesh_turntable.py:57); mask_with_grad=False with images
    # gives the hard mask any(image < 1) (renderer.py:311-312)
    with torch.no_grad():
        dflt = r(mesh, R, T)
        hard = r(mesh, R, T, return_mask=True, mask_with_grad=False)
    assert set(dflt) == {"image"} and torch.equal(dflt["image"].cpu(), img)
    assert torch.equal(hard["mask"].cpu(), torch.any(img < 1, dim=-1).float()) and torch.equal(hard["image"].cpu(), img)
    mm = masked["mask_out_masks"]
    assert float((masked["image"][mm] - 1.0).abs().max()) == 0.0 and float(masked["mask"][mm].abs().max()) == 0.0


@pytest.mark.gpu
def test_phong_image_backward_matches_oracle_autograd():
    """FootRenderer.forward(return_images=True) WITH gradients (the predicted renders of the pixel / perceptual losses,
    model.py:1089-1107): d/d(world vertices) and d/d(vertex colours) of a weighted sum of the image against torch autograd of
    oracle/raster_oracle.py::render_image (same rules as PyTorch3D: through the barycentrics of position / normal / texel, the
    vertex normals, the signed edge distance of the blend; pix_to_face and the depth order carry none).  Tolerance rel-L2 2e-3:
    fp32 atomics, and the sigmoid(-dist / 1e-4) factor that makes edge pixels 1e4 times more sensitive than interior ones."""
    dev, verts, tf, _, _, _ = _scene(162, ["topdown"], B=2, seed=11)
    S = 40
    r = FootRenderer(S, device=dev)
    R, T = r.view_from(["topdown", "45"])
    g = torch.Generator().manual_seed(12)
    colours = torch.rand(2, verts.shape[1], 3, generator=g)
    wimg = torch.rand(2, 2, S, S, 3, generator=g)
    # oracle (CPU, fp32 like the kernels so that the same pixels hit the same faces)
    vo = verts.clone().requires_grad_(True)
    co = colours.clone().requires_grad_(True)
    tot = 0.0
    imgs_o = []
    for m in range(2):
        for k in range(2):
            im = ro.render_image(vo[m], tf, co[m], R[k], T[k], S, S)
            imgs_o.append(im.detach())
            tot = tot + (im * wimg[m, k]).sum()
    tot.backward()
    # CUDA
    vg = verts.to(dev).requires_grad_(True)
    cg = colours.to(dev).requires_grad_(True)
    out = r(Meshes(vg, tf.to(dev), textures=cg), R, T, return_images=True)
    img = out["image"]
    assert img.requires_grad
    (img * wimg.to(dev)).sum().backward()
    same = torch.stack([(img[m, k].detach().cpu() - imgs_o[m * 2 + k]).abs().max() for m in range(2) for k in range(2)])
    assert float(same.max()) < 2e-3                                         # same fragments on both sides
    assert float(vo.grad.abs().max()) > 0 and float(co.grad.abs().max()) > 0
    assert rel_l2(cg.grad.cpu(), co.grad) < 2e-3, rel_l2(cg.grad.cpu(), co.grad)
    assert rel_l2(vg.grad.cpu(), vo.grad) < 2e-3, rel_l2(vg.grad.cpu(), vo.grad)
    # the image term composes with the silhouette term in one backward (model.py:1089: both requested together)
    vg2 = verts.to(dev).requires_grad_(True)
    both = r(Meshes(vg2, tf.to(dev), textures=colours.to(dev)), R, T, return_images=True, return_mask=True)
    ((both["image"] * wimg.to(dev)).sum() + both["mask"].sum()).backward()
    assert torch.isfinite(vg2.grad).all() and float((vg2.grad - vg.grad).abs().max()) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("K", [100, 3])
def test_feature_render_matches_oracle_forward_and_backward(K):
    """FootRenderer(return_features=True) = FeatureShader over the silhouette fragments (renderer.py:74-105, 277-284): forward and
    the gradients w.r.t. the per-vertex features and the NDC vertices against oracle/raster_oracle.py::feature_blend under autograd
    (K = 3 exercises the faces-per-pixel truncation).  Tolerances: 2e-4 absolute forward, rel-L2 2e-3 on the gradients."""
    from find_b200.renderer import render_features
    dev, verts, tf, r, R, T = _scene(162, ["topdown", "side1"], seed=21)
    H = W = 40
    C = 5
    g = torch.Generator().manual_seed(22)
    feats = torch.randn(1, verts.shape[1], C, generator=g)
    wimg = torch.rand(2, H, W, C, generator=g)
    ndc = r.project(verts, R, T)                                          # [2, V, 3]
    a = ndc.clone().requires_grad_(True)
    fo = feats.clone().requires_grad_(True)
    want = torch.stack([ro.feature_blend(a[i], tf, fo[0], H, W, faces_per_pixel=K) for i in range(2)])
    (want * wimg).sum().backward()
    b = ndc.to(dev).requires_grad_(True)
    fg = feats.to(dev).requires_grad_(True)
    got, p2f = render_features(b, tf.to(dev), fg, (H, W), num_views=2, faces_per_pixel=K)
    assert got.shape == (2, H, W, C)
    assert float((got.detach().cpu() - want.detach()).abs().max()) < 2e-4
    (got * wimg.to(dev)).sum().backward()
    assert float(fo.grad.abs().max()) > 0 and float(a.grad.abs().max()) > 0
    assert rel_l2(fg.grad.cpu(), fo.grad) < 2e-3, rel_l2(fg.grad.cpu(), fo.grad)
    assert rel_l2(b.grad.cpu(), a.grad) < 2e-3, rel_l2(b.grad.cpu(), a.grad)
    # through FootRenderer: world vertices, masked faces zero the feature image, shapes as the reference's
    vg = verts.to(dev).requires_grad_(True)
    out = r(Meshes(vg, tf.to(dev)), R, T, return_images=False, return_mask=True, return_features=True, features=fg.detach()[0],
            mask_out_faces=True, masked_faces=torch.arange(0, tf.shape[0], 2, device=dev), return_mask_out_masks=True)
    assert out["features"].shape == (1, 2, 64, 64, C) and out["mask"].shape == (1, 2, 64, 64)
    assert float(out["features"][out["mask_out_masks"]].abs().max()) == 0.0
    out["features"].sum().backward()
    assert torch.isfinite(vg.grad).all() and float(vg.grad.abs().max()) > 0


@pytest.mark.gpu
def test_points_render_matches_oracle():
    """findb200_points_render (the keypoint overlay's PointsRenderer) against oracle/raster_oracle.py::render_points: clouds with
    overlapping discs, depth ties, points behind the camera and more than K points over one pixel."""
    from find_b200.renderer import render_points
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(3)
    for (N, P, C, S, radius, K) in [(3, 40, 3, 64, 0.15, 10), (2, 300, 5, 48, 0.3, 4), (1, 1, 3, 32, 0.03, 10), (2, 17, 1, (40, 24), 0.2, 16)]:
        pts = torch.rand(N, P, 3, generator=g) * torch.tensor([2.0, 2.0, 0.5]) - torch.tensor([1.0, 1.0, 0.1])   # some z < 0
        pts[:, ::5, 2] = 0.25                                           # depth ties
        if P > 8:
            pts[:, 1] = pts[:, 0]                                       # a coincident pair: same disc, same depth
        feats = torch.rand(N, P, C, generator=g)
        img = render_points(pts.to(dev), feats.to(dev), S, radius=radius, points_per_pixel=K).cpu()
        H, W = (S, S) if isinstance(S, int) else S
        assert img.shape == (N, H, W, C)
        for n in range(N):
            ref = ro.render_points(pts[n], feats[n], H, W, radius=radius, points_per_pixel=K)
            assert float((img[n] - ref).abs().max()) < 2e-6, (N, P, C, S, n)
        assert float(img.abs().max()) > 0
    # nothing to draw
    assert float(render_points(torch.zeros(2, 0, 3, device=dev), torch.zeros(2, 0, 3, device=dev), 16).abs().sum()) == 0


@pytest.mark.gpu
def test_foot_renderer_keypoint_overlay():
    """FootRenderer.forward(keypoints=..., keypoints_blend=True) (renderer.py:365-377, call site eval_3d.py:97): red discs at the
    projected keypoints, pasted over the image where they land; the overlay equals the oracle's point render of the projected
    keypoints per (mesh, view)."""
    dev = torch.device("cuda")
    tv, tf = syn.make_template(642)
    B, S = 2, 64
    verts = torch.stack([tv, tv * 1.05]).to(dev)
    cols = torch.rand(B, tv.shape[0], 3, generator=torch.Generator().manual_seed(1)).to(dev)
    r = FootRenderer(S, device=dev)
    R, T = r.view_from(["topdown", "side1"])
    kps = verts[:, ::80].contiguous()                                   # 9 keypoints per mesh, on the surface
    out = r(Meshes(verts, tf.to(dev), textures=cols), R, T, return_images=True, keypoints=kps, keypoints_blend=True)
    assert out["keypoints"].shape == (B, 2, S, S, 3) and out["keypoints_blend"].shape == (B, 2, S, S, 3)
    ndc = r.project(kps, R, T).cpu()
    feats = torch.zeros(kps.shape[1], 3)
    feats[:, 0] = 1.0
    for m in range(B):
        for v in range(2):
            ref = ro.render_points(ndc[m * 2 + v], feats, S, S, radius=0.03, points_per_pixel=10)
            assert float((out["keypoints"][m, v].cpu() - ref).abs().max()) < 2e-6
    kp = out["keypoints"]
    on = (kp > 0).any(-1)
    assert on.any() and not on.all()
    assert torch.equal(out["keypoints_blend"][on], kp[on]) and torch.equal(out["keypoints_blend"][~on], out["image"][~on])
    assert float(kp[..., 1:].abs().max()) == 0.0                         # red only
    with pytest.raises(ValueError):
        r(Meshes(verts, tf.to(dev), textures=cols), R, T, keypoints=kps[:1])
