"""Property tests (hypothesis) of the Chamfer / kNN contract — SURVEY.md §4 item 2.

  * permuting the target cloud permutes the returned indices consistently (distances unchanged bit for bit; with exact ties
    the winner is the tied target with the lowest index in the PERMUTED order),
  * Chamfer is symmetric under swapping its arguments,
  * translating both clouds leaves the loss unchanged up to fp32 rounding,
  * a cloud against itself gives zero, a ragged cloud equals its unpadded prefix.

The CPU half runs the C oracle (oracle/chamfer_oracle.c); the GPU half (``-m gpu``) runs the same properties through the C ABI
(find_b200.knn_points / chamfer_distance) and, per example, compares indices and distances bit for bit with the oracle.
Shapes are drawn small (the oracle is O(P1 P2)); coordinates come from a coarse lattice with probability 1/2, so exact ties
are common.
"""
import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import chamfer_oracle as co


@st.composite
def clouds(draw, max_n=3, max_p=96):
    n = draw(st.integers(1, max_n))
    p1, p2 = draw(st.integers(1, max_p)), draw(st.integers(1, max_p))
    seed = draw(st.integers(0, 2**31 - 1))
    lattice = draw(st.booleans())
    rng = np.random.default_rng(seed)
    if lattice:      # few distinct coordinates: many exact ties
        x = rng.integers(-3, 4, size=(n, p1, 3)).astype(np.float32) * np.float32(0.25)
        y = rng.integers(-3, 4, size=(n, p2, 3)).astype(np.float32) * np.float32(0.25)
    else:
        x = (rng.standard_normal((n, p1, 3)) * 0.1).astype(np.float32)
        y = (rng.standard_normal((n, p2, 3)) * 0.1).astype(np.float32)
    return x, y, seed


COMMON = dict(deadline=None, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large])


def _check_permutation(knn, x, y, seed):
    """knn(x, y) -> (dists [n,p1], idx [n,p1]).  perm maps permuted position -> original index."""
    rng = np.random.default_rng(seed ^ 0x5EED)
    perm = rng.permutation(y.shape[1])
    d0, i0 = knn(x, y)
    d1, i1 = knn(x, np.ascontiguousarray(y[:, perm]))
    # distances are a property of the sets: identical bit for bit
    np.testing.assert_array_equal(d0.view(np.uint32), d1.view(np.uint32))
    # the permuted call returns, for every query, the LOWEST permuted position among the targets at the minimum distance
    for n in range(x.shape[0]):
        # candidates = targets whose fp32 FMA-chain distance equals the reported minimum, recomputed the oracle's way
        # (d = q - t rounded to fp32; products are exact in fp64, so one rounding per fma)
        dx, dy, dz = [(x[n][:, None, k] - y[n][None, perm, k]).astype(np.float32) for k in range(3)]
        dist = (dx.astype(np.float64) * dx.astype(np.float64)).astype(np.float32)
        dist = (dy.astype(np.float64) * dy.astype(np.float64) + dist.astype(np.float64)).astype(np.float32)
        dist = (dz.astype(np.float64) * dz.astype(np.float64) + dist.astype(np.float64)).astype(np.float32)
        ties = dist == d1[n][:, None]
        assert ties.any(axis=1).all()
        np.testing.assert_array_equal(i1[n], ties.argmax(axis=1))
        # and it names the same POINT as the unpermuted call whenever the minimum is unique
        unique = ties.sum(axis=1) == 1
        np.testing.assert_array_equal(perm[i1[n]][unique], i0[n][unique])


def _oracle_knn(x, y):
    d, i = co.knn1(x, y)
    return d, i


@settings(max_examples=40, **COMMON)
@given(clouds())
def test_oracle_permutation_of_targets(c):
    x, y, seed = c
    _check_permutation(_oracle_knn, x, y, seed)


@settings(max_examples=40, **COMMON)
@given(clouds())
def test_oracle_chamfer_symmetric_and_self_zero(c):
    x, y, _ = c
    a = co.chamfer(x, y, with_grad=False)["loss"]
    b = co.chamfer(y, x, with_grad=False)["loss"]
    assert a == pytest.approx(b, rel=1e-6, abs=1e-12)
    assert co.chamfer(x, x, with_grad=False)["loss"] == 0.0


@settings(max_examples=40, **COMMON)
@given(clouds(), st.floats(-0.5, 0.5), st.floats(-0.5, 0.5), st.floats(-0.5, 0.5))
def test_oracle_translation_invariance(c, tx, ty, tz):
    x, y, _ = c
    t = np.array([tx, ty, tz], dtype=np.float32)
    a = co.chamfer(x, y, with_grad=False)["loss"]
    b = co.chamfer(x + t, y + t, with_grad=False)["loss"]
    # the subtraction q - t is where a translation costs bits: |q| grows to ~0.5 + extent, so d = q - t is exact only to an
    # ulp of that magnitude; the bound is a few ulps of the largest coordinate, squared terms dominated by 2 d eps
    scale = float(np.abs(np.concatenate([x.ravel(), y.ravel()])).max() + np.abs(t).max())
    assert abs(a - b) <= 64 * scale * 2.0 ** -23 * (2 * np.sqrt(max(a, b)) + scale * 2.0 ** -23) + 1e-12


@settings(max_examples=30, **COMMON)
@given(clouds(max_n=2, max_p=64), st.data())
def test_oracle_ragged_equals_prefix(c, data):
    x, y, _ = c
    l1 = np.array([data.draw(st.integers(1, x.shape[1])) for _ in range(x.shape[0])], dtype=np.int64)
    l2 = np.array([data.draw(st.integers(1, y.shape[1])) for _ in range(x.shape[0])], dtype=np.int64)
    d, i = co.knn1(x, y, l1, l2)
    for n in range(x.shape[0]):
        dn, in_ = co.knn1(x[n:n + 1, : l1[n]], y[n:n + 1, : l2[n]])
        np.testing.assert_array_equal(d[n, : l1[n]].view(np.uint32), dn[0].view(np.uint32))
        np.testing.assert_array_equal(i[n, : l1[n]], in_[0])
        assert (d[n, l1[n]:] == 0).all() and (i[n, l1[n]:] == 0).all()      # padded rows are zeroed (knn.py convention)


# ---------------------------------------------------------------------------------------------------------------- GPU
def _gpu_knn(x, y):
    from find_b200 import knn_points
    r = knn_points(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda())
    return r.dists[..., 0].cpu().numpy(), r.idx[..., 0].cpu().numpy()


@pytest.mark.gpu
@settings(max_examples=60, **COMMON)
@given(clouds(max_n=3, max_p=200))
def test_gpu_knn_matches_oracle_and_permutes(c):
    x, y, seed = c
    d, i = _gpu_knn(x, y)
    do, io = co.knn1(x, y)
    np.testing.assert_array_equal(i, io)
    np.testing.assert_array_equal(d.view(np.uint32), do.view(np.uint32))
    _check_permutation(_gpu_knn, x, y, seed)


@pytest.mark.gpu
@settings(max_examples=60, **COMMON)
@given(clouds(max_n=3, max_p=200))
def test_gpu_chamfer_symmetric_self_zero_and_matches_oracle(c):
    from find_b200 import chamfer_distance
    x, y, _ = c
    xt, yt = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    a = float(chamfer_distance(xt, yt)[0])
    b = float(chamfer_distance(yt, xt)[0])
    assert a == pytest.approx(b, rel=2e-6, abs=1e-12)
    assert float(chamfer_distance(xt, xt)[0]) == 0.0
    assert a == pytest.approx(co.chamfer(x, y, with_grad=False)["loss"], rel=2e-6, abs=1e-12)
