"""Pins oracle/mesh_oracle.py (CPU restatement of PyTorch3D's sampler / edge loss / cot-Laplacian as FIND calls them,
/root/reference/src/model/losses.py:63-67,93-99) with known-answer tests.  PyTorch3D is not on this box and the reference
ships no vectors, so these KATs are what anchors the restatement ("parity unpinned" against PyTorch3D itself)."""
import numpy as np
import torch

from oracle import mesh_oracle as mo
from find_b200 import synthetic as syn


def test_philox_published_vectors():
    # Random123 kat_vectors, philox4x32-10
    def h(t):
        return [int(x[0]) for x in t]
    assert h(mo.philox4x32_10([0], [0], [0], [0], 0, 0)) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    f = 0xFFFFFFFF
    assert h(mo.philox4x32_10([f], [f], [f], [f], f, f)) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert h(mo.philox4x32_10([0x243F6A88], [0x85A308D3], [0x13198A2E], [0x03707344], 0xA4093822, 0x299F31D0)) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]


def test_face_areas_known():
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 2]], dtype=np.float32)
    f = np.array([[0, 1, 2], [0, 1, 3], [1, 1, 2]])
    a = mo.face_areas(v, f)
    assert a.tolist() == [0.5, 1.0, 0.0]


def test_barycentric_identities():
    rng = np.random.default_rng(0)
    u, v = rng.random(1000, dtype=np.float32), rng.random(1000, dtype=np.float32)
    w0, w1, w2 = mo.barycentric(u, v)
    assert np.all(w0 >= 0) and np.all(w1 >= 0) and np.all(w2 >= 0)
    assert np.allclose(w0 + w1 + w2, 1.0, atol=2e-7)
    w = mo.barycentric(np.array([0.0, 1.0, 1.0], np.float32), np.array([0.3, 0.0, 1.0], np.float32))
    assert [x.tolist() for x in w] == [[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]


def test_draw_is_area_weighted_chi_square():
    # two triangles of area ratio 1:3 plus a degenerate face that must never be drawn
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [3, 0, 0], [0, 0, 0]], dtype=np.float32)[None]
    f = np.array([[0, 1, 2], [0, 3, 2], [0, 0, 1]])
    out = mo.sample_points(v, f, 40000, seed=123)
    cnt = np.bincount(out["face_idx"][0], minlength=3)
    assert cnt[2] == 0
    exp = 40000 * np.array([0.25, 0.75])
    chi2 = float(((cnt[:2] - exp) ** 2 / exp).sum())
    assert chi2 < 10.8                                    # 1 dof, p = 0.001
    # samples lie in the plane z = 0 and inside the union of the two triangles
    s = out["samples"][0]
    assert np.all(s[:, 2] == 0) and np.all(s[:, 0] >= 0) and np.all(s[:, 1] >= 0)
    assert np.all(s[:, 0] / 3.0 + s[:, 1] <= 1.0 + 1e-6)


def test_draw_uniform_inside_triangle():
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=np.float32)[None]
    out = mo.sample_points(v, np.array([[0, 1, 2]]), 60000, seed=7)
    s = out["samples"][0]
    # uniform on the unit right triangle: E[x] = E[y] = 1/3, P(x + y < 0.5) = 0.25
    assert abs(s[:, 0].mean() - 1 / 3) < 5e-3 and abs(s[:, 1].mean() - 1 / 3) < 5e-3
    assert abs((s[:, 0] + s[:, 1] < 0.5).mean() - 0.25) < 8e-3


def test_injected_draw_and_backward():
    tv, tf = syn.make_template(162)
    verts = tv.numpy()[None].repeat(2, 0)
    S = 50
    rng = np.random.default_rng(1)
    fi = rng.integers(0, tf.shape[0], (2, S))
    uv = rng.random((2, 2, S), dtype=np.float32)
    out = mo.sample_points(verts, tf.numpy(), S, face_idx=fi, uv=uv)
    assert np.array_equal(out["face_idx"], fi)
    # autograd cross-check of the restated backward
    vt = torch.tensor(verts, dtype=torch.float64, requires_grad=True)
    f = torch.as_tensor(tf.numpy().astype(np.int64))
    w = torch.tensor(out["bary"], dtype=torch.float64)
    smp = sum(w[:, :, k:k + 1] * torch.stack([vt[n][f[torch.as_tensor(fi[n]), k]] for n in range(2)]) for k in range(3))
    g = torch.randn(2, S, 3, dtype=torch.float64)
    (smp * g).sum().backward()
    mine = mo.sample_points_backward(g.numpy(), tf.numpy(), fi, out["bary"], verts.shape[1])
    assert np.allclose(mine, vt.grad.numpy(), atol=1e-12)


def test_edge_loss_known():
    # unit square split in two triangles: 5 unique edges, lengths 1,1,1,1,sqrt2 -> mean of squares = 6/5
    v = torch.tensor([[[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]]], dtype=torch.float32)
    f = np.array([[0, 1, 2], [0, 2, 3]])
    assert mo.unique_edges(f).shape == (5, 2)
    assert abs(float(mo.mesh_edge_loss(v, f)) - 1.2) < 1e-6
    assert abs(float(mo.mesh_edge_loss(v, f, target_length=1.0)) - (np.sqrt(2) - 1) ** 2 / 5) < 1e-6


def test_cot_laplacian_known():
    # regular hexagon fan around a centre vertex: all angles 60 deg, cot = 1/sqrt3; every spoke is shared by two faces
    ang = np.arange(6) * np.pi / 3
    ring = np.stack([np.cos(ang), np.sin(ang), np.zeros(6)], 1)
    v = torch.tensor(np.concatenate([[[0, 0, 0]], ring])[None], dtype=torch.float64)
    f = np.array([[0, 1 + i, 1 + (i + 1) % 6] for i in range(6)])
    L = mo.cot_laplacian_dense(v[0], torch.as_tensor(f))
    c = 1 / np.sqrt(3)
    assert torch.allclose(L, L.t())
    # (B2 + C2 - A2) / area = 4 cot, so after the /4 an entry is the plain sum of the opposite cotangents
    assert np.allclose(L[0, 1:].numpy(), 2 * c)              # spoke: two opposite angles of 60 deg
    assert np.allclose(L[1, 2].item(), c)                    # rim edge: one face
    # centre vertex sits at the weighted mean of its ring -> its laplacian residual is 0; rim residuals are not
    loss = mo.mesh_laplacian_smoothing_cot(v, f)
    Lv = (L @ v[0]) / L.sum(1, keepdim=True) - v[0]
    assert abs(float(Lv[0].norm())) < 1e-12
    assert abs(float(loss) - float(Lv.norm(dim=1).mean())) < 1e-12


def test_smoothness_gradient_flows_only_through_explicit_verts():
    tv, tf = syn.make_template(162)
    torch.manual_seed(0)
    v = (tv[None] + 0.002 * torch.randn(2, tv.shape[0], 3)).requires_grad_(True)
    total, e, l = mo.mesh_smoothness_loss(v, tf.numpy())
    total.backward()
    assert torch.isfinite(v.grad).all() and float(v.grad.norm()) > 0
    assert abs(float(total) - (0.1 * float(l) + 10 * float(e))) < 1e-7
