"""CPU: pins oracle/field_oracle.py against tests/golden/field_golden.npz, which oracle/make_golden.py
produced by running the reference's OWN NeuralDisplacementField (src/model/model.py:206-453) under stubs."""
import hashlib

import numpy as np
import pytest
import torch

from oracle import field_oracle as fo


def _params(golden, dtype=torch.float32):
    return {k[2:]: torch.from_numpy(golden[k]).to(dtype) for k in golden.files if k.startswith("w/")}


def test_fourier_B_known_answers(golden):
    B = fo.fourier_B()
    assert hashlib.sha256(B.numpy().tobytes()).hexdigest() == str(golden["B_sha256"])
    # SURVEY.md §3.2 known-answer values
    assert str(golden["B_sha256"]) == "56ea7d6792aa67255f40794c4135159fbb9a9d7f6f0aa95ef5cd6b72abe17d52"
    np.testing.assert_allclose(B[0, :4].numpy(), [-12.5601, 8.9556, 1.6748, 7.5142], atol=1e-4)
    np.testing.assert_array_equal(B.numpy(), golden["B"])


@pytest.mark.parametrize("tag", ["A", "B"])
def test_forward_and_grads_match_reference(golden, tag):
    torch.set_num_threads(1)
    p = {k: v.requires_grad_(True) for k, v in _params(golden).items()}
    B = torch.from_numpy(golden["B"])
    z = {n: torch.from_numpy(golden[f"{tag}/{n}"]).requires_grad_(True) for n in ("shapevec", "texvec", "posevec")}
    out = fo.field_forward(p, B, torch.from_numpy(golden[f"{tag}/pos"]), **z)
    np.testing.assert_allclose(out["disp"].detach().numpy(), golden[f"{tag}/disp"], rtol=1e-5, atol=1e-7)
    np.testing.assert_allclose(out["col"].detach().numpy(), golden[f"{tag}/col"], rtol=1e-5, atol=1e-7)
    loss = (out["disp"] * torch.from_numpy(golden[f"{tag}/gd"])).sum() + \
           (out["col"] * torch.from_numpy(golden[f"{tag}/gc"])).sum()
    loss.backward()
    for n in z:
        np.testing.assert_allclose(z[n].grad.numpy(), golden[f"{tag}/grad_{n}"], rtol=2e-4, atol=1e-6)
    for k in golden.files:
        if k.startswith(f"{tag}/grad/"):
            np.testing.assert_allclose(p[k.split("/", 2)[2]].grad.numpy(), golden[k], rtol=2e-4, atol=2e-5)
        elif k.startswith(f"{tag}/gradhead/"):
            g = p[k.split("/", 2)[2]].grad
            np.testing.assert_allclose(g[:4].numpy(), golden[k], rtol=2e-4, atol=2e-5)
            s = golden[k.replace("gradhead", "gradsum")]
            assert g.double().norm().item() == pytest.approx(s[1], rel=1e-4)


def test_fp64_oracle_close_to_fp32(golden):
    p64 = _params(golden, torch.float64)
    B = torch.from_numpy(golden["B"]).double()
    z = {n: torch.from_numpy(golden[f"A/{n}"]).double() for n in ("shapevec", "texvec", "posevec")}
    out = fo.field_forward(p64, B, torch.from_numpy(golden["A/pos"]).double(), **z)
    rel = np.linalg.norm(out["disp"].numpy() - golden["A/disp"]) / np.linalg.norm(golden["A/disp"])
    assert rel < 1e-4      # fp32 reference vs fp64 restatement (the fp32 sin/cos argument rounding dominates)


def test_reg_transform_identity_and_known_rotation():
    pts = torch.randn(2, 5, 3)
    reg = torch.tensor([[0, 0, 0, 0, 0, 0, 1, 1, 1.0]]).repeat(2, 1)   # reference init, model.py:343-346
    assert torch.equal(fo.reg_transform(pts, reg), pts)
    # 90 deg about X (row-vector convention p' = p @ Rx): y-axis -> (0,0,-1)... check against explicit matrix
    e = torch.tensor([[np.pi / 2, 0.0, 0.0]])
    R = fo.euler_xyz_to_matrix(e)[0]
    np.testing.assert_allclose(R.numpy(), [[1, 0, 0], [0, 0, -1], [0, 1, 0]], atol=1e-7)
    reg = torch.tensor([[1.0, 2.0, 3.0, np.pi / 2, 0, 0, 2.0, 2.0, 2.0]])
    X = fo.reg_transform(torch.tensor([[[0.0, 1.0, 0.0]]]), reg)
    np.testing.assert_allclose(X.numpy(), [[[1.0, 2.0, 1.0]]], atol=1e-6)   # (0,2,0)@R = (0,0,-2) -> +t
    # R is orthonormal for random angles, composition order Rx Ry Rz
    e = torch.randn(4, 3)
    R = fo.euler_xyz_to_matrix(e)
    np.testing.assert_allclose((R @ R.transpose(1, 2)).numpy(), np.tile(np.eye(3), (4, 1, 1)), atol=1e-6)


def test_chamfer_torch_matches_c_oracle():
    from oracle import chamfer_oracle as co
    g = torch.Generator().manual_seed(0)
    x, y = torch.randn(2, 60, 3, generator=g), torch.randn(2, 50, 3, generator=g)
    xl, yl = torch.tensor([60, 41]), torch.tensor([50, 17])
    xr = x.clone().requires_grad_(True)
    loss = fo.chamfer_torch(xr, y, xl, yl)
    loss.backward()
    r = co.chamfer(x.numpy(), y.numpy(), xl.numpy(), yl.numpy())
    assert loss.item() == pytest.approx(r["loss"], rel=1e-5)
    np.testing.assert_allclose(xr.grad.numpy(), r["grad_x"], rtol=1e-4, atol=1e-7)
