"""GPU parity of csrc/mesh.cu (surface sampling, MeshSmoothnessLoss) against oracle/mesh_oracle.py, through the C ABI
(find_b200.mesh wraps it with ctypes).  Bars: face ids, barycentrics and samples BIT-EXACT (the draw is integer / exactly
rounded fp32 by construction); sampling gradient rel-L2 <= 1e-6 (fp32 atomics, order only); smoothness losses and
gradients rel-L2 <= 2e-4 vs the fp32 oracle and <= 2e-5 vs the same oracle run in fp64 is NOT required (fp32 cotangents)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from find_b200 import synthetic as syn
from oracle import mesh_oracle as mo


def _dev():
    return torch.device("cuda:0")


def _rel(a, b):
    a, b = np.asarray(a, np.float64).ravel(), np.asarray(b, np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _deformed(N, V, seed=0, amp=0.003):
    tv, tf = syn.make_template(V)
    g = torch.Generator().manual_seed(seed)
    verts = tv[None] + amp * torch.randn(N, tv.shape[0], 3, generator=g)
    return verts.contiguous(), tf


@pytest.mark.parametrize("N,V,S", [(1, 162, 257), (3, 642, 5000), (8, 10242, 5000)])
def test_sampling_draw_bit_exact(N, V, S):
    from find_b200.mesh import Meshes, sample_points_from_meshes
    verts, tf = _deformed(N, V, seed=N)
    m = Meshes(verts.to(_dev()), tf.to(_dev()))
    smp, fidx, bary = sample_points_from_meshes(m, S, seed=0x1234_5678_9ABC, return_draw=True)
    ref = mo.sample_points(verts.numpy(), tf.numpy(), S, seed=0x1234_5678_9ABC)
    assert np.array_equal(fidx.cpu().numpy(), ref["face_idx"])
    assert np.array_equal(bary.cpu().numpy(), ref["bary"])
    assert np.array_equal(smp.cpu().numpy(), ref["samples"])


def test_sampling_shared_template_and_seed_semantics():
    from find_b200.mesh import Meshes, sample_points_from_meshes
    tv, tf = syn.make_template(642)
    m = Meshes(tv[None].to(_dev()).expand(4, -1, -1), tf.to(_dev()))          # stride-0 verts (extend_template)
    a = sample_points_from_meshes(m, 1000, seed=5)
    b = sample_points_from_meshes(m, 1000, seed=5)
    c = sample_points_from_meshes(m, 1000, seed=6)
    assert torch.equal(a, b) and not torch.equal(a, c)
    assert not torch.equal(a[0], a[1])                       # meshes draw from different counters
    ref = mo.sample_points(tv.numpy()[None].repeat(4, 0), tf.numpy(), 1000, seed=5)
    assert np.array_equal(a.cpu().numpy(), ref["samples"])
    torch.manual_seed(3); d = sample_points_from_meshes(m, 100)
    torch.manual_seed(3); e = sample_points_from_meshes(m, 100)
    assert torch.equal(d, e)


def test_sampling_ragged_meshes_and_textures():
    from find_b200.mesh import Meshes, sample_points_from_meshes
    (v0, f0), (v1, f1) = syn.make_template(162), syn.make_template(642)
    cols = [torch.rand(v0.shape[0], 3), torch.rand(v1.shape[0], 3)]
    m = Meshes([v0.to(_dev()), v1.to(_dev())], [f0.to(_dev()), f1.to(_dev())])
    colp = torch.zeros(2, v1.shape[0], 3)
    colp[0, : v0.shape[0]], colp[1] = cols[0], cols[1]
    m.textures = colp.to(_dev())
    smp, tex, fidx, bary = sample_points_from_meshes(m, 3000, return_textures=True, seed=11, return_draw=True)
    assert int(fidx[0].max()) < f0.shape[0]                  # padding faces are never drawn
    vp = m.verts_padded().cpu().numpy()
    fp = m.faces_padded().cpu().numpy()
    ref = mo.sample_points(vp, fp, 3000, seed=11, num_faces=[f0.shape[0], f1.shape[0]], attr=colp.numpy())
    assert np.array_equal(fidx.cpu().numpy(), ref["face_idx"])
    assert np.array_equal(smp.cpu().numpy(), ref["samples"])
    assert np.array_equal(tex.cpu().numpy(), ref["attr"])


def test_sampling_injected_draw_and_gradient():
    from find_b200.mesh import Meshes, sample_points_from_meshes
    N, S = 3, 4000
    verts, tf = _deformed(N, 642, seed=2)
    rng = np.random.default_rng(4)
    fi = rng.integers(0, tf.shape[0], (N, S))
    uv = rng.random((2, N, S), dtype=np.float32)
    vd = verts.to(_dev()).requires_grad_(True)
    m = Meshes(vd, tf.to(_dev()))
    smp = sample_points_from_meshes(m, S, face_idx=torch.as_tensor(fi).to(_dev()), uv=torch.as_tensor(uv).to(_dev()))
    ref = mo.sample_points(verts.numpy(), tf.numpy(), S, face_idx=fi, uv=uv)
    assert np.array_equal(smp.detach().cpu().numpy(), ref["samples"])
    g = torch.randn(N, S, 3, generator=torch.Generator().manual_seed(0))
    (smp * g.to(_dev())).sum().backward()
    gref = mo.sample_points_backward(g.numpy(), tf.numpy(), fi, ref["bary"], verts.shape[1])
    assert _rel(vd.grad.cpu().numpy(), gref) <= 1e-6


def test_sampling_area_distribution_full_size():
    """size-independent property at the full cfg2 size: face counts follow the areas (chi-square), samples lie on faces."""
    from find_b200.mesh import Meshes, sample_points_from_meshes
    verts, tf = _deformed(2, 10242, seed=9)
    m = Meshes(verts.to(_dev()), tf.to(_dev()))
    S = 400000
    smp, fidx, bary = sample_points_from_meshes(m, S, seed=99, return_draw=True)
    area = mo.face_areas(verts[0].numpy(), tf.numpy().astype(np.int64)).astype(np.float64)
    # bin faces into 64 groups of equal expected mass to keep expected counts large
    order = np.argsort(area)
    cum = np.cumsum(area[order]) / area.sum()
    grp = np.minimum((cum * 64).astype(int), 63)
    gof = np.zeros(len(area), int); gof[order] = grp
    cnt = np.bincount(gof[fidx[0].cpu().numpy()], minlength=64).astype(np.float64)
    exp = np.bincount(gof, weights=area, minlength=64) / area.sum() * S
    chi2 = float(((cnt - exp) ** 2 / exp).sum())
    assert chi2 < 120.0                                       # 63 dof, p ~ 1e-5
    w = bary.cpu().numpy()
    assert np.all(w >= 0) and np.allclose(w.sum(-1), 1.0, atol=3e-7)


@pytest.mark.parametrize("N,V", [(1, 162), (4, 2562), (8, 10242)])
def test_mesh_smoothness_loss_and_grad(N, V):
    from find_b200.mesh import Meshes, MeshSmoothnessLoss, mesh_edge_loss, mesh_laplacian_smoothing
    verts, tf = _deformed(N, V, seed=V, amp=0.002)
    vd = verts.to(_dev()).requires_grad_(True)
    m = Meshes(vd, tf.to(_dev()))
    loss = MeshSmoothnessLoss()(m)
    loss.backward()
    vr = verts.clone().requires_grad_(True)
    total, e, l = mo.mesh_smoothness_loss(vr, tf.numpy())
    total.backward()
    assert abs(float(loss) - float(total)) <= 2e-4 * abs(float(total))
    assert _rel(vd.grad.cpu().numpy(), vr.grad.numpy()) <= 2e-4
    with torch.no_grad():
        assert abs(float(mesh_edge_loss(m)) - float(e)) <= 1e-5 * abs(float(e))
        assert abs(float(mesh_laplacian_smoothing(m, method="cot")) - float(l)) <= 2e-4 * abs(float(l))


def test_mesh_edge_loss_target_length_and_upstream_scale():
    from find_b200.mesh import Meshes, mesh_edge_loss
    verts, tf = _deformed(2, 642, seed=5)
    vd = verts.to(_dev()).requires_grad_(True)
    (3.0 * mesh_edge_loss(Meshes(vd, tf.to(_dev())), target_length=0.01)).backward()
    vr = verts.clone().requires_grad_(True)
    (3.0 * mo.mesh_edge_loss(vr, tf.numpy(), target_length=0.01)).backward()
    assert _rel(vd.grad.cpu().numpy(), vr.grad.numpy()) <= 1e-4


def test_displacement_loss_with_surface_sampling():
    """losses.py:59-90 with num_samples: Chamfer between 5000 surface samples of both meshes (reference default)."""
    from find_b200 import chamfer_distance
    from find_b200.losses import DisplacementLoss
    from find_b200.mesh import Meshes, sample_points_from_meshes
    verts, tf = _deformed(2, 2562, seed=1)
    gtv, _ = _deformed(2, 2562, seed=2, amp=0.004)
    vd = verts.to(_dev()).requires_grad_(True)
    pred, gt = Meshes(vd, tf.to(_dev())), Meshes(gtv.to(_dev()), tf.to(_dev()))
    out = DisplacementLoss()(None, {"meshes": pred, "verts": vd}, {"mesh": gt}, 0, num_samples=5000, seed=21)["loss"]
    out.backward()
    sp = sample_points_from_meshes(pred, 5000, seed=43)
    sg = sample_points_from_meshes(gt, 5000, seed=42)
    ref, _ = chamfer_distance(sp, sg)
    assert float(out) == float(ref)
    assert torch.isfinite(vd.grad).all() and float(vd.grad.abs().sum()) > 0


def test_cpu_tensors_refused():
    from find_b200.mesh import Meshes, MeshSmoothnessLoss, sample_points_from_meshes
    tv, tf = syn.make_template(162)
    m = Meshes(tv[None], tf)
    with pytest.raises(RuntimeError):
        sample_points_from_meshes(m, 10, seed=0)
    with pytest.raises(RuntimeError):
        MeshSmoothnessLoss()(m)
