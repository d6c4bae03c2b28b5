"""GPU: parity of csrc/chamfer.cu (through the C ABI via find_b200.chamfer) against oracle/chamfer_oracle.c
in FMA-chain mode.  Bar: NN indices and distances BIT-EXACT; loss/grads within fp32 summation-order tolerance."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["cull_kq1", "cull_kq2", "cull_kq4", "brute", "cta_kq1", "cta_kq2", "cta_kq4", "cull_kq2_sort1", "grid"], autouse=True)
def knn_mode(request):
    """Every test runs on the culled kernel (1, 2 and 4 queries per lane; one warp per group and the CTA-cooperative variant) and
    on the brute-force tile kernel."""
    from find_b200 import _lib
    mode, kq = {"cull_kq1": (0, 1), "cull_kq2": (0, 2), "cull_kq4": (0, 4), "brute": (1, 0),
                "cta_kq1": (2, 1), "cta_kq2": (2, 2), "cta_kq4": (2, 4), "cull_kq2_sort1": (0, 2), "grid": (3, 0)}[request.param]
    assert _lib.load().findb200_set_knn_mode(mode, kq) == 0
    # "sort1": clouds below 32768 points are sorted by one cluster per cloud in a single launch instead of six kernels
    _lib.load().findb200_set_sort_cta_max(32768 if request.param.endswith("sort1") else -1)
    yield request.param
    _lib.load().findb200_set_knn_mode(0, 0)
    _lib.load().findb200_set_sort_cta_max(-1)


def _surface(seed, n, p, radii=(0.125, 0.05, 0.05), jitter=2e-4):
    """Points on an ellipsoid shell: the spatial structure the training clouds have (culling is exercised hard)."""
    g = torch.Generator().manual_seed(seed)
    u = torch.randn(n, p, 3, generator=g)
    u = u / u.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    return u * torch.tensor(radii) * (1 + 0.05 * torch.randn(n, 1, 1, generator=g)) + jitter * torch.randn(n, p, 3, generator=g)


def _cloud(seed, n, p, scale=0.1):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, p, 3, generator=g) * scale


def _check_knn(p1, p2, l1=None, l2=None):
    from find_b200 import knn_points
    from oracle import chamfer_oracle as co
    dev = "cuda:0"
    r = knn_points(p1.to(dev), p2.to(dev), None if l1 is None else l1.to(dev), None if l2 is None else l2.to(dev))
    d, i = co.knn1(p1.numpy(), p2.numpy(), None if l1 is None else l1.numpy(), None if l2 is None else l2.numpy(),
                   threads=co.max_threads())
    assert r.dists.shape == (p1.shape[0], p1.shape[1], 1) and r.idx.dtype == torch.int64
    np.testing.assert_array_equal(r.idx[..., 0].cpu().numpy(), i)
    np.testing.assert_array_equal(r.dists[..., 0].cpu().numpy().view(np.uint32), d.view(np.uint32))


@pytest.mark.parametrize("n,p1,p2", [(1, 1, 1), (1, 33, 7), (2, 512, 513), (3, 1000, 2049), (2, 5000, 3000),
                                     (1, 10242, 10000)])
def test_knn_bit_exact(n, p1, p2):
    _check_knn(_cloud(1, n, p1), _cloud(2, n, p2))


def test_knn_ragged_empty_and_ties():
    p1, p2 = _cloud(3, 4, 700), _cloud(4, 4, 900)
    _check_knn(p1, p2, torch.tensor([700, 1, 0, 333]), torch.tensor([900, 450, 17, 0]))
    # exact ties: duplicated targets -> lowest index wins (SURVEY §8(c) KAT 4)
    t = _cloud(5, 1, 300)
    p2 = torch.cat([t, t, t], 1)
    _check_knn(_cloud(6, 1, 400), p2)
    # lattice queries vs lattice targets: massive exact ties
    g = torch.stack(torch.meshgrid(*[torch.arange(8.0)] * 3, indexing="ij"), -1).reshape(1, -1, 3)
    _check_knn(g + 0.5, g)


@pytest.mark.parametrize("n,p1,p2", [(2, 10242, 10000), (1, 20000, 50000), (3, 777, 31)])
def test_knn_surface_clouds_bit_exact(n, p1, p2):
    _check_knn(_surface(20, n, p1), _surface(21, n, p2))


def test_knn_cross_chunk_ties_and_degenerate_clouds():
    # the same target set repeated far apart in index order: equal distances in different chunks -> lowest ORIGINAL index
    t = _surface(22, 1, 2000)
    _check_knn(_surface(23, 1, 3000), torch.cat([t, t.flip(1), t], 1))
    # every target identical; all points on a line; a single cell
    _check_knn(_cloud(24, 1, 200), torch.ones(1, 500, 3) * 0.3)
    line = torch.linspace(-1, 1, 4000).view(1, -1, 1) * torch.tensor([1.0, 0.0, 0.0])
    _check_knn(line[:, ::3] + 1e-3, line)
    _check_knn(torch.zeros(2, 65, 3), torch.zeros(2, 97, 3))


def test_knn_split_target_path():
    # few queries, many targets -> the scheduler splits the target range and combines with atomicMin
    _check_knn(_cloud(7, 1, 100), _cloud(8, 1, 50000))


def test_knn_lattice_shift_kat():
    from find_b200 import chamfer_distance
    g = torch.stack(torch.meshgrid(*[torch.arange(6.0)] * 3, indexing="ij"), -1).reshape(1, -1, 3)
    delta = torch.tensor([0.125, -0.25, 0.0625])
    loss, none = chamfer_distance(g.cuda(), (g + delta).cuda())
    assert none is None
    assert loss.item() == pytest.approx(2 * float((delta.double() ** 2).sum()), rel=1e-6)


@pytest.mark.parametrize("n,p1,p2,ragged", [(2, 300, 200, False), (3, 2000, 1500, True), (8, 10242, 10000, False)])
def test_chamfer_loss_and_grads(n, p1, p2, ragged):
    from find_b200 import chamfer_distance
    from oracle import chamfer_oracle as co
    x, y = _cloud(10, n, p1), _cloud(11, n, p2)
    xl = yl = None
    if ragged:
        xl = torch.tensor([p1, p1 // 2, 5][:n])
        yl = torch.tensor([p2 // 3, p2, 1][:n])
    xg = x.cuda().requires_grad_(True)
    yg = y.cuda().requires_grad_(True)
    loss, _ = chamfer_distance(xg, yg, x_lengths=None if xl is None else xl.cuda(),
                               y_lengths=None if yl is None else yl.cuda())
    (loss * 3.0).backward()
    r = co.chamfer(x.numpy(), y.numpy(), None if xl is None else xl.numpy(), None if yl is None else yl.numpy(),
                   threads=co.max_threads())
    assert loss.item() == pytest.approx(r["loss"], rel=2e-6)
    np.testing.assert_allclose(xg.grad.cpu().numpy(), 3.0 * r["grad_x"], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(yg.grad.cpu().numpy(), 3.0 * r["grad_y"], rtol=1e-5, atol=1e-9)


def test_knn_backward_matches_oracle():
    from find_b200 import knn_points
    from oracle import chamfer_oracle as co
    p1, p2 = _cloud(12, 2, 500), _cloud(13, 2, 400)
    a = p1.cuda().requires_grad_(True)
    b = p2.cuda().requires_grad_(True)
    r = knn_points(a, b)
    w = torch.rand(2, 500, 1, generator=torch.Generator().manual_seed(1))
    (r.dists * w.cuda()).sum().backward()
    _, i = co.knn1(p1.numpy(), p2.numpy())
    g1, g2 = co.knn1_backward(p1.numpy(), p2.numpy(), i, w[..., 0].numpy())
    np.testing.assert_allclose(a.grad.cpu().numpy(), g1, rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(b.grad.cpu().numpy(), g2, rtol=1e-5, atol=1e-8)


def test_other_reductions_and_pointclouds():
    from find_b200 import Pointclouds, chamfer_distance
    x, y = _cloud(14, 2, 50).cuda(), _cloud(15, 2, 70).cuda()
    full, _ = chamfer_distance(x, y)
    s, _ = chamfer_distance(x, y, batch_reduction="sum", point_reduction="mean")
    assert s.item() == pytest.approx(2 * full.item(), rel=1e-5)
    (cx, cy), _ = chamfer_distance(x, y, batch_reduction=None, point_reduction=None)
    assert cx.shape == (2, 50) and cy.shape == (2, 70)
    assert (cx.mean(1) + cy.mean(1)).mean().item() == pytest.approx(full.item(), rel=1e-5)
    # ragged Pointclouds path used by the z-cutoff branches (losses.py:69-85)
    pc = Pointclouds([x[0, :20], x[1, :50]])
    a, _ = chamfer_distance(pc, y)
    b, _ = chamfer_distance(x, y, x_lengths=torch.tensor([20, 50]).cuda())
    assert a.item() == pytest.approx(b.item(), rel=1e-6)
    with pytest.raises(ValueError):
        chamfer_distance(x, y, point_reduction="max")


def test_property_symmetry_and_permutation():
    from find_b200 import chamfer_distance, knn_points
    x, y = _cloud(16, 2, 800).cuda(), _cloud(17, 2, 900).cuda()
    a, _ = chamfer_distance(x, y)
    b, _ = chamfer_distance(y, x)
    assert a.item() == pytest.approx(b.item(), rel=1e-6)
    perm = torch.randperm(900, generator=torch.Generator().manual_seed(0)).cuda()
    r0, r1 = knn_points(x, y), knn_points(x, y[:, perm])
    assert torch.equal(r0.dists, r1.dists)                       # distances are permutation invariant, bit for bit
    assert torch.equal(perm[r1.idx[..., 0]], r0.idx[..., 0])      # indices map back (no exact ties in random data)
