"""CPU: known-answer tests that pin oracle/chamfer_oracle.c (SURVEY.md §4 list, §8(c) items 3-5).
PyTorch3D is absent, so these KATs are the only pin for the Chamfer/kNN oracle ("parity unpinned")."""
import numpy as np
import pytest

from oracle import chamfer_oracle as co


def brute_f64(p1, p2):
    d = ((p1[:, :, None, :].astype(np.float64) - p2[:, None, :, :].astype(np.float64)) ** 2).sum(-1)
    return d.min(-1), d.argmin(-1)


def test_identical_clouds_zero():
    rng = np.random.RandomState(0)
    x = rng.randn(2, 50, 3).astype(np.float32)
    r = co.chamfer(x, x.copy())
    assert r["loss"] == 0.0
    assert (r["idx_x"] == np.arange(50)[None]).all() and (r["dist_x"] == 0).all()
    assert np.abs(r["grad_x"]).max() == 0 and np.abs(r["grad_y"]).max() == 0


def test_lattice_shift():
    # lattice pitch 1, shift delta with 2|delta| < pitch  =>  every NN is the shifted twin, loss = 2|delta|^2
    g = np.stack(np.meshgrid(*[np.arange(4.0)] * 3, indexing="ij"), -1).reshape(1, -1, 3).astype(np.float32)
    delta = np.array([0.125, -0.25, 0.0625], np.float32)
    r = co.chamfer(g, g + delta)
    assert (r["idx_x"] == np.arange(64)).all() and (r["idx_y"] == np.arange(64)).all()
    assert r["loss"] == pytest.approx(2 * float((delta.astype(np.float64) ** 2).sum()), rel=1e-7)
    # d/dx of mean_i |x_i - y_i|^2 + mean_j |y_j - x_j|^2  = 2 * 2 (x - y) / P
    np.testing.assert_allclose(r["grad_x"][0], np.tile(-4 * delta / 64, (64, 1)), rtol=1e-6)


def test_ties_lowest_index():
    p1 = np.zeros((1, 3, 3), np.float32)
    p2 = np.array([[[1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 0, 0], [0, 0, 0.5], [0, 0, 0.5]]], np.float32)
    d, i = co.knn1(p1, p2)
    assert (i == 4).all() and (d == 0.25).all()
    d, i = co.knn1(p1, p2[:, :4])
    assert (i == 0).all() and (d == 1.0).all()


def test_ragged_and_empty():
    rng = np.random.RandomState(1)
    p1, p2 = rng.randn(3, 7, 3).astype(np.float32), rng.randn(3, 9, 3).astype(np.float32)
    l1, l2 = np.array([7, 3, 0]), np.array([9, 0, 4])
    d, i = co.knn1(p1, p2, l1, l2)
    assert (d[1] == 0).all() and (i[1] == 0).all()          # no targets -> zero-initialised outputs
    assert (d[2] == 0).all() and (i[2] == 0).all()          # no queries
    dd, ii = brute_f64(p1[:1], p2[:1])
    assert (i[0] == ii[0]).all()
    r = co.chamfer(p1, p2, l1, l2)
    want = (d[0].astype(np.float64).sum() / 7 + co.knn1(p2, p1, l2, l1)[0][0].astype(np.float64).sum() / 9) / 3
    assert r["loss"] == pytest.approx(want, rel=1e-12)
    assert np.abs(r["grad_x"][1:]).max() == 0 and np.abs(r["grad_y"][1:]).max() == 0
    assert np.abs(r["grad_x"][0, :]).max() > 0


def test_single_points_and_p1_ne_p2():
    p1 = np.array([[[0.0, 0, 0]]], np.float32)
    p2 = np.array([[[3.0, 4, 0], [1, 0, 0]]], np.float32)
    r = co.chamfer(p1, p2)
    assert r["loss"] == pytest.approx(1.0 + (25.0 + 1.0) / 2)
    assert r["idx_x"][0, 0] == 1


def test_vs_float64_bruteforce():
    rng = np.random.RandomState(2)
    p1, p2 = rng.randn(2, 600, 3).astype(np.float32), rng.randn(2, 700, 3).astype(np.float32)
    d, i = co.knn1(p1, p2)
    dd, ii = brute_f64(p1, p2)
    np.testing.assert_allclose(d, dd, rtol=2e-6)
    assert (i != ii).mean() < 0.002                           # only fp32 near-ties may differ
    d_t, i_t = co.knn1(p1, p2, threads=4)
    assert (d_t == d).all() and (i_t == i).all()


def test_rounding_modes_differ_in_last_ulp():
    # KAT: a triple where fmaf-chain and separately-rounded sums differ -> proves the two modes are distinct
    rng = np.random.RandomState(3)
    p1, p2 = rng.randn(1, 2000, 3).astype(np.float32), rng.randn(1, 2000, 3).astype(np.float32)
    d0, _ = co.knn1(p1, p2, mode=co.MODE_FMA_CHAIN)
    d1, _ = co.knn1(p1, p2, mode=co.MODE_SEPARATE)
    assert (d0 != d1).any()
    np.testing.assert_allclose(d0, d1, rtol=3e-7)
    # explicit check of the contract formula for query 0
    q = p1[0, 0].astype(np.float32)
    t = p2[0]
    dx, dy, dz = [(q[k] - t[:, k]).astype(np.float32) for k in range(3)]
    fma = lambda a, b, c: np.float32(np.float64(a) * np.float64(b) + np.float64(c))   # exact product, one rounding
    cand = np.array([fma(dz[j], dz[j], fma(dy[j], dy[j], np.float32(dx[j] * dx[j]))) for j in range(2000)], np.float32)
    assert d0[0, 0] == cand.min()


def test_backward_matches_finite_difference():
    rng = np.random.RandomState(4)
    x, y = rng.randn(1, 40, 3).astype(np.float32), rng.randn(1, 30, 3).astype(np.float32)
    r = co.chamfer(x, y)
    eps = 1e-3
    for (i, d) in [(0, 0), (5, 1), (17, 2)]:
        xp, xm = x.copy(), x.copy()
        xp[0, i, d] += eps
        xm[0, i, d] -= eps
        fd = (co.chamfer(xp, y, with_grad=False)["loss"] - co.chamfer(xm, y, with_grad=False)["loss"]) / (2 * eps)
        assert r["grad_x"][0, i, d] == pytest.approx(fd, rel=2e-2, abs=1e-4)
    g1, g2 = co.knn1_backward(x, y, r["idx_x"], np.full((1, 40), 1.0 / 40, np.float32))
    assert np.abs(g1.sum(1) + g2.sum(1)).max() < 1e-5         # scatter is the negative of the direct term
