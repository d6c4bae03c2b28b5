"""GPU parity: the sm_100a displacement-field kernels (through the C ABI / field_apply) against
  * the golden fixture produced by the reference's own NeuralDisplacementField (tests/golden/field_golden.npz), and
  * the fp32 CPU oracle (oracle/field_oracle.py) with torch autograd on seeded inputs, incl. the registration transform.

Stated tolerances (fp32 reference vs fp16-operand / fp32-accumulate tensor-core compute; DESIGN.md "Numerics"):
  * outputs (disp, col, verts):                       rel-L2 <= 1e-3   (measured ~1e-4)
  * kink-free parameters (oracle/field_oracle.kinkfree_params: no ReLU pre-activation near 0, so the masks of the
    fp16-operand forward equal the fp32 forward's): gradients rel-L2 <= 2e-3 vs the fp32 reference (measured
    <= 1.2e-3) — this isolates the backward kernels from mask flips.  Those parameters put a large common mode on
    every activation (mean/std = 6), which amplifies fp16 operand rounding on the OUTPUTS, so outputs are held to
    1e-2 there (measured 6e-3..8e-3; the rounding-matched oracle shows the same 6e-3..8e-3 against fp32)
  * gradients with ordinary weights:                   rel-L2 <= 5e-2 vs the fp32 reference, AND no further from it
    than 1.5 x the distance of the rounding-matched oracle (CPU fp32 autograd whose forward rounds GEMM operands
    to fp16) + 2e-3.  Reason: any reduced-precision forward (fp16, tf32 and bf16 alike) flips the sign of ~1e-4
    of the pre-activations, and every flip changes that unit's gradient by 100 %: the deviation is a property of
    the operand format (bf16 would be 3e-2..1e-1), not of the backward kernels.
"""
import numpy as np
import pytest
import torch

from oracle import field_oracle as fo

pytestmark = pytest.mark.gpu

TOL_OUT = 1e-3
TOL_GRAD = 2e-3          # kink-free parameters: gradients vs the fp32 reference
TOL_KINKFREE_OUT = 1e-2  # kink-free parameters: outputs vs the fp32 reference (ill-conditioned on purpose)
TOL_GRAD_FP32 = 5e-2     # ordinary weights, vs the fp32 reference


def rel_l2(a, b):
    a = torch.as_tensor(a).double().cpu().reshape(-1)
    b = torch.as_tensor(b).double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _params_from(p, dev, requires_grad=True):
    from find_b200.field_ops import PARAM_NAMES
    ws = [p[n + ".weight"].to(dev).clone().requires_grad_(requires_grad) for n in PARAM_NAMES]
    bs = [p[n + ".bias"].to(dev).clone().requires_grad_(requires_grad) for n in PARAM_NAMES]
    return ws, bs


def _golden_params(golden):
    return {k[2:]: torch.from_numpy(golden[k]) for k in golden.files if k.startswith("w/")}


@pytest.mark.parametrize("tag", ["A", "B"])
def test_field_matches_reference_golden(golden, tag):
    from find_b200.field_ops import PARAM_NAMES, field_apply
    dev = torch.device("cuda:0")
    p = _golden_params(golden)
    ws, bs = _params_from(p, dev)
    Bm = torch.from_numpy(golden["B"]).to(dev)
    t = lambda k: torch.from_numpy(golden[f"{tag}/{k}"]).to(dev)
    pos = t("pos")
    zs, zt, zp = (t(k).requires_grad_(True) for k in ("shapevec", "texvec", "posevec"))
    disp, col, _ = field_apply(pos, zs, zp, zt, None, Bm, None, ws, bs, want_col=True, want_verts=False)
    assert rel_l2(disp, golden[f"{tag}/disp"]) < TOL_OUT
    assert rel_l2(col, golden[f"{tag}/col"]) < TOL_OUT
    ((disp * t("gd")).sum() + (col * t("gc")).sum()).backward()
    for n, z in (("shapevec", zs), ("texvec", zt), ("posevec", zp)):
        assert rel_l2(z.grad, golden[f"{tag}/grad_{n}"]) < TOL_GRAD_FP32, n
    for i, name in enumerate(PARAM_NAMES):
        gb = golden[f"{tag}/grad/{name}.bias"]
        assert rel_l2(bs[i].grad, gb) < TOL_GRAD_FP32, name + ".bias"
        key = f"{tag}/grad/{name}.weight"
        if key in golden.files:
            assert rel_l2(ws[i].grad, golden[key]) < TOL_GRAD_FP32, name
        else:
            s = golden[f"{tag}/gradsum/{name}.weight"]
            g64 = ws[i].grad.double()
            assert abs(g64.norm().item() - s[1]) <= TOL_GRAD_FP32 * s[1], name
            assert rel_l2(ws[i].grad[:4], golden[f"{tag}/gradhead/{name}.weight"]) < TOL_GRAD_FP32, name


def _oracle_grads(p, Bm, pos, zs, zp, zt, reg, avg, gd, gc, gv, operand_dtype=None):
    from find_b200.field_ops import PARAM_NAMES
    B = zs.shape[0]
    po = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    zo = [z.clone().requires_grad_(True) for z in (zs, zp, zt)]
    rego = reg.clone().requires_grad_(True) if reg is not None else None
    res = fo.field_forward(po, Bm, pos, shapevec=zo[0], posevec=zo[1], texvec=zo[2], avg_col=avg,
                           operand_dtype=operand_dtype)
    verts = fo.reg_transform(pos.expand(B, -1, -1) + res["disp"], rego)
    ((res["disp"] * gd).sum() + (res["col"] * gc).sum() + (verts * gv).sum()).backward()
    out = {"disp": res["disp"].detach(), "col": res["col"].detach(), "verts": verts.detach(),
           "g_shape": zo[0].grad, "g_pose": zo[1].grad, "g_tex": zo[2].grad}
    if reg is not None:
        out["g_reg"] = rego.grad
    for n in PARAM_NAMES:
        out["gW " + n] = po[n + ".weight"].grad
        out["gb " + n] = po[n + ".bias"].grad
    return out


def _oracle_case(B, V, shared, seed, with_reg=True, frozen=False, kinkfree=False):
    from find_b200.field_ops import PARAM_NAMES, field_apply
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(seed)
    Bm = fo.fourier_B()
    pos = torch.randn(1 if shared else B, V, 3, generator=g) * 0.05
    zs, zp, zt = (torch.randn(B, 100, generator=g) * 0.1 for _ in range(3))
    reg = torch.cat([torch.randn(B, 3, generator=g) * 0.005, torch.randn(B, 3, generator=g) * 0.05,
                     1 + torch.randn(B, 3, generator=g) * 0.02], 1) if with_reg else None
    avg = torch.rand(3, generator=g) * 0.1
    gd, gc, gv = (torch.randn(B, V, 3, generator=g) for _ in range(3))
    p = fo.kinkfree_params(seed, Bm, pos, zs, zp, zt) if kinkfree else fo.random_params(seed)

    ref = _oracle_grads(p, Bm, pos, zs, zp, zt, reg, avg, gd, gc, gv)
    matched = _oracle_grads(p, Bm, pos, zs, zp, zt, reg, avg, gd, gc, gv, torch.float16)

    ws, bs = _params_from(p, dev, requires_grad=not frozen)
    zg = [z.to(dev).requires_grad_(True) for z in (zs, zp, zt)]
    regg = reg.to(dev).requires_grad_(True) if with_reg else None
    disp, col, verts = field_apply(pos.to(dev), zg[0], zg[1], zg[2], regg, Bm.to(dev), avg.to(dev), ws, bs,
                                   frozen_mlp=frozen)
    ((disp * gd.to(dev)).sum() + (col * gc.to(dev)).sum() + (verts * gv.to(dev)).sum()).backward()
    mine = {"disp": disp, "col": col, "verts": verts, "g_shape": zg[0].grad, "g_pose": zg[1].grad, "g_tex": zg[2].grad}
    if with_reg:
        mine["g_reg"] = regg.grad
    if not frozen:
        for i, n in enumerate(PARAM_NAMES):
            mine["gW " + n], mine["gb " + n] = ws[i].grad, bs[i].grad
    else:
        assert all(w.grad is None for w in ws)
    for k, v in mine.items():
        e = rel_l2(v, ref[k])
        if kinkfree:
            assert e < (TOL_KINKFREE_OUT if k in ("disp", "col", "verts") else TOL_GRAD), (k, e)
        elif k in ("disp", "col", "verts"):
            assert e < TOL_OUT, (k, e)
        else:
            em = rel_l2(matched[k], ref[k])
            assert e < TOL_GRAD_FP32, (k, e)
            assert e < 1.5 * em + 2e-3, (k, e, em)


def test_field_gradients_kinkfree():
    _oracle_case(B=3, V=333, shared=True, seed=21, kinkfree=True)


def test_field_gradients_kinkfree_per_foot_points():
    _oracle_case(B=2, V=257, shared=False, seed=22, kinkfree=True)


def test_field_vs_oracle_shared_template():
    _oracle_case(B=3, V=333, shared=True, seed=11)


def test_field_vs_oracle_per_foot_points():
    _oracle_case(B=2, V=257, shared=False, seed=12)


def test_field_vs_oracle_no_reg_single_tile():
    _oracle_case(B=1, V=5, shared=True, seed=13, with_reg=False)


def test_field_frozen_mlp_latent_fitting():
    _oracle_case(B=4, V=300, shared=True, seed=14, frozen=True)


def test_field_many_tiles_multi_wave():
    # more items than SMs: exercises the persistent loops, ring phases and barrier parities across items
    _oracle_case(B=24, V=1500, shared=True, seed=15)


def test_field_many_trunk_tiles_two_per_cta():
    # per-foot points with more trunk tiles than 2 x SMs: the trunk kernels run TWO tiles per CTA (pair), which changes the
    # order in which base.0's nine weight chunks travel through the ring
    _oracle_case(B=8, V=4800, shared=False, seed=16)


def test_field_no_cpu_fallback():
    from find_b200.field_ops import field_apply
    p = fo.random_params(0)
    from find_b200.field_ops import PARAM_NAMES
    ws = [p[n + ".weight"] for n in PARAM_NAMES]
    bs = [p[n + ".bias"] for n in PARAM_NAMES]
    with pytest.raises(RuntimeError):
        field_apply(torch.zeros(1, 4, 3), torch.zeros(1, 100), torch.zeros(1, 100), torch.zeros(1, 100), None,
                    fo.fourier_B(), None, ws, bs)
