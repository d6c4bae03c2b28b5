"""tools/field_diag.py — prints every output/gradient error of the field kernels vs the fp32 oracle and vs the
rounding-matched (fp16-operand) oracle, for a few shapes."""
import os, sys, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import field_oracle as fo  # noqa: E402
from find_b200.field_ops import PARAM_NAMES, field_apply  # noqa: E402


def rel(a, b):
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def case(B, V, shared, seed, with_reg=True, kinkfree=False):
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
    refs = {}
    for tag, od in (("fp32", None), ("matched", torch.float16)):
        po = {k: v.clone().requires_grad_(True) for k, v in p.items()}
        zo = [z.clone().requires_grad_(True) for z in (zs, zp, zt)]
        rego = reg.clone().requires_grad_(True) if with_reg else None
        res = fo.field_forward(po, Bm, pos, shapevec=zo[0], posevec=zo[1], texvec=zo[2], avg_col=avg, operand_dtype=od)
        verts = fo.reg_transform(pos.expand(B, -1, -1) + res["disp"], rego)
        ((res["disp"] * gd).sum() + (res["col"] * gc).sum() + (verts * gv).sum()).backward()
        r = {"disp": res["disp"], "col": res["col"], "verts": verts, "g_shape": zo[0].grad, "g_pose": zo[1].grad,
             "g_tex": zo[2].grad}
        if with_reg:
            r["g_reg"] = rego.grad
        for n in PARAM_NAMES:
            r["gW " + n] = po[n + ".weight"].grad
            r["gb " + n] = po[n + ".bias"].grad
        refs[tag] = r
    ws = [p[n + ".weight"].to(dev).requires_grad_(True) for n in PARAM_NAMES]
    bs = [p[n + ".bias"].to(dev).requires_grad_(True) for n in PARAM_NAMES]
    zg = [z.to(dev).requires_grad_(True) for z in (zs, zp, zt)]
    regg = reg.to(dev).requires_grad_(True) if with_reg else None
    disp, col, verts = field_apply(pos.to(dev), zg[0], zg[1], zg[2], regg, Bm.to(dev), avg.to(dev), ws, bs)
    ((disp * gd.to(dev)).sum() + (col * gc.to(dev)).sum() + (verts * gv.to(dev)).sum()).backward()
    torch.cuda.synchronize()
    mine = {"disp": disp, "col": col, "verts": verts, "g_shape": zg[0].grad, "g_pose": zg[1].grad, "g_tex": zg[2].grad}
    if with_reg:
        mine["g_reg"] = regg.grad
    for i, n in enumerate(PARAM_NAMES):
        mine["gW " + n] = ws[i].grad
        mine["gb " + n] = bs[i].grad
    print(f"--- B={B} V={V} shared={shared} kinkfree={kinkfree}")
    for k in mine:
        print(f"{k:24s} vs fp32 {rel(mine[k], refs['fp32'][k]):.2e}   vs matched {rel(mine[k], refs['matched'][k]):.2e}"
              f"   (matched vs fp32 {rel(refs['matched'][k], refs['fp32'][k]):.2e})")


if __name__ == "__main__":
    if "--kinkfree" in sys.argv:
        case(3, 333, True, 21, kinkfree=True)
        case(2, 257, False, 22, kinkfree=True)
        sys.exit(0)
    case(3, 333, True, 11)
    case(2, 257, False, 12)
    case(24, 1500, True, 15)
