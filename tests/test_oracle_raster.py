"""Known-answer tests that pin oracle/raster_oracle.py (the soft-silhouette restatement; PyTorch3D itself is absent)."""
import math

import torch

from oracle import raster_oracle as ro

TRI = torch.tensor([[-0.5, -0.5, 1.0], [0.5, -0.5, 1.0], [-0.5, 0.5, 1.0]], dtype=torch.float64)
FACE = torch.tensor([[0, 1, 2]])


def _pix(H, W, x, y):
    xs, ys = ro.pixel_centres(H, W, torch.float64)
    c = int(torch.argmin((xs - x).abs())); r = int(torch.argmin((ys - y).abs()))
    return r, c


def test_pixel_centres_follow_pytorch3d_ndc():
    xs, ys = ro.pixel_centres(4, 4, torch.float64)
    assert torch.allclose(xs, torch.tensor([0.75, 0.25, -0.25, -0.75], dtype=torch.float64))   # +x is LEFT, +y is UP
    assert torch.allclose(ys, xs)


def test_one_triangle_known_alphas():
    sigma, blur = 0.1, 1.0
    for tri in (TRI, TRI[[0, 2, 1]]):                                   # both windings (no back-face culling)
        a = ro.silhouette(tri, FACE, 4, 4, blur_radius=blur, sigma=sigma)
        r, c = _pix(4, 4, -0.25, -0.25)                                 # inside: nearest edges at 0.25
        assert abs(float(a[r, c]) - 1 / (1 + math.exp(-0.0625 / sigma))) < 1e-12
        r, c = _pix(4, 4, 0.25, 0.25)                                   # outside: hypotenuse x + y = 0 at 0.5 / sqrt 2
        assert abs(float(a[r, c]) - 1 / (1 + math.exp(0.125 / sigma))) < 1e-12
        r, c = _pix(4, 4, 0.75, 0.75)                                   # squared distance 1.125 >= blur_radius: not a candidate
        assert float(a[r, c]) == 0.0
        r, c = _pix(4, 4, 0.75, -0.75)                                  # nearest point is the vertex (0.5, -0.5)
        assert abs(float(a[r, c]) - 1 / (1 + math.exp(0.125 / sigma))) < 1e-12


def test_blur_radius_is_compared_with_the_squared_distance():
    # distance 0.25 to the left edge: squared 0.0625; candidate iff blur_radius > 0.0625 (NOT 0.25)
    r, c = _pix(8, 8, -0.625, -0.125)
    xs, _ = ro.pixel_centres(8, 8, torch.float64)
    assert abs(float(xs[c]) + 0.625) < 1e-12
    tri = TRI.clone(); tri[:, 0] += 0.125                               # left edge at x = -0.375: distance 0.25
    a_in = ro.silhouette(tri, FACE, 8, 8, blur_radius=0.07, sigma=0.1)
    a_out = ro.silhouette(tri, FACE, 8, 8, blur_radius=0.06, sigma=0.1)
    assert float(a_in[r, c]) > 0 and float(a_out[r, c]) == 0.0


def test_behind_camera_and_degenerate_faces_are_skipped():
    behind = TRI.clone(); behind[:, 2] = -1.0
    assert float(ro.silhouette(behind, FACE, 4, 4, blur_radius=1.0, sigma=0.1).abs().max()) == 0.0
    line = torch.tensor([[-0.5, -0.5, 1.0], [0.0, 0.0, 1.0], [0.5, 0.5, 1.0]], dtype=torch.float64)
    assert float(ro.silhouette(line, FACE, 4, 4, blur_radius=1.0, sigma=0.1).abs().max()) == 0.0


def test_k_truncation_keeps_the_nearest_faces_and_the_earlier_one_on_ties():
    near, far = TRI.clone(), TRI.clone() * torch.tensor([0.5, 0.5, 2.0])
    v = torch.cat([far, near, near + torch.tensor([0.01, 0.0, 0.0])])
    f = torch.tensor([[0, 1, 2], [3, 4, 5], [6, 7, 8]])
    kw = dict(blur_radius=1.0, sigma=0.1)
    a1 = ro.silhouette(v, f, 8, 8, faces_per_pixel=1, **kw)
    only_near = ro.silhouette(near, FACE, 8, 8, **kw)                    # face 1 wins the depth tie with face 2 (earlier)
    assert torch.equal(a1, only_near)
    a2, aux = ro.silhouette(v, f, 8, 8, faces_per_pixel=2, return_aux=True, **kw)
    two_near = ro.silhouette(v[3:], torch.tensor([[0, 1, 2], [3, 4, 5]]), 8, 8, **kw)
    assert torch.allclose(a2, two_near, atol=1e-15) and int(aux["count"].max()) == 3
    a3 = ro.silhouette(v, f, 8, 8, faces_per_pixel=3, **kw)
    assert float((a3 - a2).abs().max()) > 1e-3


def test_gradient_matches_finite_differences_fp64():
    torch.manual_seed(0)
    v = torch.rand(12, 3, dtype=torch.float64) * torch.tensor([1.6, 1.6, 1.0]) - torch.tensor([0.8, 0.8, -0.5])
    f = torch.tensor([[0, 1, 2], [3, 4, 5], [6, 7, 8], [9, 10, 11], [0, 4, 8], [1, 5, 9]])
    w = torch.rand(10, 10, dtype=torch.float64)
    v.requires_grad_(True)
    kw = dict(blur_radius=0.05, sigma=0.02)
    (ro.silhouette(v, f, 10, 10, **kw) * w).sum().backward()
    g = v.grad.clone()
    assert float(g[:, 2].abs().max()) == 0.0                             # depth only orders the candidates
    with torch.no_grad():
        for (i, d) in [(0, 0), (4, 1), (7, 0), (10, 1), (5, 0)]:
            e = torch.zeros_like(v); e[i, d] = 1e-6
            num = ((ro.silhouette(v + e, f, 10, 10, **kw) * w).sum() - (ro.silhouette(v - e, f, 10, 10, **kw) * w).sum()) / 2e-6
            assert abs(float(num) - float(g[i, d])) < 1e-5 * max(1.0, abs(float(num)))


def test_look_at_view_transform_known_answers():
    R, T = ro.look_at_view_transform(dist=1.0, elev=0.0, azim=0.0, up=(0.0, 1.0, 0.0))
    assert torch.allclose(R[0], torch.diag(torch.tensor([-1.0, 1.0, -1.0], dtype=torch.float64)), atol=1e-12)
    assert torch.allclose(T[0], torch.tensor([0.0, 0.0, 1.0], dtype=torch.float64), atol=1e-12)
    # the camera sits at `dist` from the target and looks at it: the target maps to (0, 0, dist) in view space
    R, T = ro.look_at_view_transform(dist=0.35, elev=-45.0, azim=30.0, at=(0.1, 0.0, 0.0), up=(1.0, 0.0, 0.0))
    at_view = torch.tensor([0.1, 0.0, 0.0], dtype=torch.float64) @ R[0] + T[0]
    assert torch.allclose(at_view, torch.tensor([0.0, 0.0, 0.35], dtype=torch.float64), atol=1e-12)
    assert torch.allclose(R[0].T @ R[0], torch.eye(3, dtype=torch.float64), atol=1e-12)
    p = ro.fov_project(torch.tensor([[0.1, 0.0, 0.0]], dtype=torch.float64), R[0], T[0])
    assert torch.allclose(p[0], torch.tensor([0.0, 0.0, 0.35], dtype=torch.float64), atol=1e-12)


def test_blocked_evaluation_equals_dense():
    torch.manual_seed(3)
    v = torch.rand(60, 3) * torch.tensor([1.8, 1.8, 1.0]) - torch.tensor([0.9, 0.9, -0.5])
    f = torch.randint(0, 60, (80, 3))
    kw = dict(blur_radius=0.01, sigma=0.05, faces_per_pixel=6)
    dense = ro.silhouette(v, f, 40, 24, **kw)
    # equal up to the rounding order of the fp32 product (torch.prod reduces rows of different length differently)
    assert float((ro.silhouette_blocked(v, f, 40, 24, rows_per_block=7, **kw) - dense).abs().max()) < 3e-7
    assert torch.equal(ro.silhouette(v, f, 40, 24, rows=(8, 19), **kw), dense[8:19])


def test_hard_rasterisation_known_barycentrics_and_depth():
    # pixel (x, y) = (-0.25, -0.25) in TRI: plain barycentrics w0 = 0.5 (vertex (-.5,-.5)), w1 = w2 = 0.25
    r, c = _pix(4, 4, -0.25, -0.25)
    p2f, z, bary, d = ro.rasterize_nearest(TRI, FACE, 4, 4, perspective_correct=False)
    assert int(p2f[r, c]) == 0 and torch.allclose(bary[r, c], torch.tensor([0.5, 0.25, 0.25], dtype=torch.float64), atol=1e-7)
    assert abs(float(z[r, c]) - 1.0) < 1e-7 and abs(float(d[r, c]) + 0.0625) < 1e-12          # inside: minus the squared distance
    rr, cc = _pix(4, 4, 0.25, 0.25)
    assert int(p2f[rr, cc]) == -1 and float(z[rr, cc]) == -1.0 and float(bary[rr, cc, 0]) == -1.0   # blur 0: outside pixels are empty
    # perspective correction with depths (1, 2, 4): b_i = (w_i / z_i) / sum_j (w_j / z_j), depth = 1 / sum_j (w_j / z_j)
    tri = TRI.clone(); tri[:, 2] = torch.tensor([1.0, 2.0, 4.0], dtype=torch.float64)
    p2f, z, bary, d = ro.rasterize_nearest(tri, FACE, 4, 4, perspective_correct=True)
    s = 0.5 / 1 + 0.25 / 2 + 0.25 / 4
    want = torch.tensor([0.5 / 1 / s, 0.25 / 2 / s, 0.25 / 4 / s], dtype=torch.float64)
    assert torch.allclose(bary[r, c], want, atol=1e-6) and abs(float(z[r, c]) - 1 / s) < 1e-6
    # the nearer of two stacked faces wins
    v = torch.cat([tri + torch.tensor([0.0, 0.0, 5.0]), TRI]); f = torch.tensor([[0, 1, 2], [3, 4, 5]])
    p2f, z, _, _ = ro.rasterize_nearest(v, f, 4, 4)
    assert int(p2f[r, c]) == 1 and abs(float(z[r, c]) - 1.0) < 1e-7


def test_phong_known_answer_and_background():
    v = torch.tensor([[0.0, 0.0, 0.0], [1.0, 0.0, 0.0], [0.0, 1.0, 0.0]], dtype=torch.float64)
    f = torch.tensor([[0, 1, 2]])
    assert torch.allclose(ro.vertex_normals(v, f), torch.tensor([[0.0, 0.0, 1.0]] * 3, dtype=torch.float64))
    p2f = torch.tensor([[0, -1]]); bary = torch.full((1, 2, 3), 1 / 3, dtype=torch.float64)
    zbuf = torch.tensor([[0.3, -1.0]], dtype=torch.float64); dists = torch.tensor([[-1.0, -1.0]], dtype=torch.float64)
    col = torch.full((3, 3), 0.5, dtype=torch.float64)
    img = ro.phong_image(p2f, bary, zbuf, dists, v, f, col, cam_centre=(1 / 3, 1 / 3, 5.0))
    ln = math.sqrt(2 / 9 + 100.0 ** 2)
    cosang = 100.0 / ln                                                    # n = (0,0,1), l = normalize((-1/3, -1/3, 100))
    spec = (2 * cosang * 1.0 - cosang) ** 64                                # v = (0,0,1): v.r = -l_z + 2 cos = cos
    want = (0.5 + 0.3 * cosang) * 0.5 + 0.2 * spec
    assert torch.allclose(img[0, 0], torch.full((3,), want, dtype=torch.float64), atol=1e-9)
    assert torch.equal(img[0, 1], torch.ones(3, dtype=torch.float64))       # empty pixel: the background colour
    # a face turned away from the light gets ambient only
    img2 = ro.phong_image(p2f, bary, zbuf, dists, v, f[:, [0, 2, 1]], col, cam_centre=(1 / 3, 1 / 3, 5.0))
    assert torch.allclose(img2[0, 0], torch.full((3,), 0.25, dtype=torch.float64), atol=1e-9)


def test_image_oracle_is_differentiable_end_to_end():
    """render_image (projection -> hard fragments -> Phong -> blend) in fp64: autograd vs central differences for vertices and
    colours; pixels that change face under the perturbation are avoided by a small step on a coarse image."""
    torch.manual_seed(4)
    v = torch.tensor([[-0.05, -0.04, 0.0], [0.06, -0.05, 0.01], [0.0, 0.06, -0.01], [0.07, 0.05, 0.02]], dtype=torch.float64)
    f = torch.tensor([[0, 1, 2], [1, 3, 2]])
    col = torch.rand(4, 3, dtype=torch.float64)
    R, T = ro.look_at_view_transform(dist=0.3, elev=20.0, azim=10.0)
    w = torch.rand(12, 12, 3, dtype=torch.float64)
    fn = lambda vv, cc: (ro.render_image(vv, f, cc, R[0], T[0], 12, 12) * w).sum()      # noqa: E731
    vg, cg = v.clone().requires_grad_(True), col.clone().requires_grad_(True)
    fn(vg, cg).backward()
    assert float(vg.grad.abs().max()) > 0 and float(cg.grad.abs().max()) > 0
    h = 1e-7
    for (i, d) in [(0, 0), (1, 1), (2, 2), (3, 0)]:
        e = torch.zeros_like(v); e[i, d] = h
        num = (fn(v + e, col) - fn(v - e, col)) / (2 * h)
        assert abs(float(num) - float(vg.grad[i, d])) < 1e-4 * max(1.0, abs(float(num)))
    for (i, d) in [(0, 1), (3, 2)]:
        e = torch.zeros_like(col); e[i, d] = 1e-6
        num = (fn(v, col + e) - fn(v, col - e)) / 2e-6
        assert abs(float(num) - float(cg.grad[i, d])) < 1e-6 * max(1.0, abs(float(num)))


def test_feature_blend_known_answers():
    feats = torch.tensor([[1.0, 0.0], [0.0, 1.0], [0.0, 0.0]], dtype=torch.float64)
    r, c = _pix(4, 4, -0.25, -0.25)                                       # inside TRI: barycentrics (0.5, 0.25, 0.25), depth 1
    out = ro.feature_blend(TRI, FACE, feats, 4, 4, sigma=0.1, gamma=0.5, blur_radius=1.0)
    prob = 1 / (1 + math.exp(-0.0625 / 0.1))
    z_inv = (100.0 - 1.0) / 99.0                                          # = 1: the face sits on the near plane of the blend
    delta = math.exp((1e-10 - z_inv) / 0.5)
    want = prob * torch.tensor([0.5, 0.25], dtype=torch.float64) / (prob + delta)
    assert torch.allclose(out[r, c], want, atol=1e-12)
    rr, cc = _pix(4, 4, 0.75, 0.75)                                       # no candidate: background (zeros) over delta = 1
    assert torch.equal(out[rr, cc], torch.zeros(2, dtype=torch.float64))
    # the nearer of two stacked faces dominates as gamma -> 0
    far = TRI.clone(); far[:, 2] = 3.0
    v = torch.cat([far, TRI]); f = torch.tensor([[0, 1, 2], [3, 4, 5]])
    fe = torch.cat([torch.zeros(3, 1, dtype=torch.float64), torch.ones(3, 1, dtype=torch.float64)])
    sharp = ro.feature_blend(v, f, fe, 4, 4, sigma=0.1, gamma=1e-4, blur_radius=1.0)
    assert abs(float(sharp[r, c, 0]) - 1.0) < 1e-6


def test_render_points_known_answers():
    """renderer.py:139-142, 365-373 (PointsRenderer + AlphaCompositor): disc weights 1 - d^2 / r^2, front-to-back compositing by
    depth, points behind the camera dropped, K truncation keeps the nearest, the earlier point wins a depth tie."""
    H = W = 8
    r = 0.3
    red, green, blue = torch.eye(3)
    # pixel (3, 3) has its centre at (0.125, 0.125): d^2 = 2 * 0.125^2 to a point at the origin
    w = 1.0 - 2 * 0.125 ** 2 / r ** 2
    pts = torch.tensor([[0.0, 0.0, 0.3], [0.0, 0.0, 0.2], [0.0, 0.0, -0.1]])
    img = ro.render_points(pts, torch.stack([red, green, blue]), H, W, radius=r)
    assert torch.allclose(img[3, 3], torch.tensor([(1 - w) * w, w, 0.0]), atol=1e-6)      # green (nearer) first, blue is behind
    assert float(img[0, 0].abs().sum()) == 0.0                                          # outside every disc: zeros
    # K = 1 keeps only the nearest point
    img1 = ro.render_points(pts, torch.stack([red, green, blue]), H, W, radius=r, points_per_pixel=1)
    assert torch.allclose(img1[3, 3], torch.tensor([0.0, w, 0.0]), atol=1e-6)
    # equal depths: the earlier point is composited first
    tie = torch.tensor([[0.0, 0.0, 0.2], [0.0, 0.0, 0.2]])
    imgt = ro.render_points(tie, torch.stack([red, green]), H, W, radius=r)
    assert torch.allclose(imgt[3, 3], torch.tensor([w, (1 - w) * w, 0.0]), atol=1e-6)
    assert ro.render_points(torch.zeros(0, 3), torch.zeros(0, 3), 4, 4).abs().sum() == 0
