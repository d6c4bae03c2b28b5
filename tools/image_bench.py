"""tools/image_bench.py — Phong image forward / forward+backward of FootRenderer at the reference's size (32 images of 256 x 256
on the 10 242-vertex template) and the gradient error on a small case vs the oracle's autograd."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import synthetic as syn
from find_b200.mesh import Meshes
from find_b200.renderer import FootRenderer
from oracle import raster_oracle as ro

dev = torch.device("cuda")
out = {}
# ---- small case: errors
tv, tf = syn.make_template(162)
g = torch.Generator().manual_seed(11)
verts = tv[None] + 0.002 * torch.randn(2, tv.shape[0], 3, generator=g)
S = 40
r = FootRenderer(S, device=dev)
R, T = r.view_from(["topdown", "45"])
colours = torch.rand(2, verts.shape[1], 3, generator=g)
wimg = torch.rand(2, 2, S, S, 3, generator=g)
vo, co = verts.clone().requires_grad_(True), colours.clone().requires_grad_(True)
tot = sum((ro.render_image(vo[m], tf, co[m], R[k], T[k], S, S) * wimg[m, k]).sum() for m in range(2) for k in range(2))
tot.backward()
vg, cg = verts.to(dev).requires_grad_(True), colours.to(dev).requires_grad_(True)
(r(Meshes(vg, tf.to(dev), textures=cg), R, T)["image"] * wimg.to(dev)).sum().backward()
rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
out["grad_rel_l2_vs_oracle"] = {"verts": rel(vg.grad.cpu(), vo.grad), "colours": rel(cg.grad.cpu(), co.grad)}
# ---- full size: timing
tv, tf = syn.make_template(10242)
B, nv, S = 8, 4, 256
verts = (tv[None] + 0.001 * torch.randn(B, tv.shape[0], 3, generator=g)).to(dev)
colours = torch.rand(B, tv.shape[0], 3, generator=g).to(dev)
r = FootRenderer(S, device=dev)
R, T = r.sample_views(nviews=nv, dist_mean=0.3, dist_std=0, elev_min=-90, elev_max=90, azim_min=-90, azim_max=90, seed=3)
def fwd():
    with torch.no_grad():
        return r(Meshes(verts, tf.to(dev), textures=colours), R, T)["image"]
def fwd_bwd():
    v = verts.clone().requires_grad_(True)
    c = colours.clone().requires_grad_(True)
    im = r(Meshes(v, tf.to(dev), textures=c), R, T)["image"]
    im.square().mean().backward()
    return v.grad
for name, fn in (("forward_ms", fwd), ("forward_backward_ms", fwd_bwd)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out[name] = sorted(ts)[len(ts) // 2]
out["images"] = f"{B * nv} x {S} x {S}, {tv.shape[0]} vertices / {tf.shape[0]} faces"
print(json.dumps(out, indent=1))

# ---- feature rendering (FeatureShader over the 100-face silhouette fragments), C = 21 as the reference's per-vertex clusters
from find_b200.renderer import render_features
C = 21
feats = torch.randn(B, tv.shape[0], C, generator=g).to(dev)
ndc = r.project(verts, R, T)
def f_fwd():
    with torch.no_grad():
        return render_features(ndc, tf.to(dev), feats, S, num_views=nv)[0]
def f_fwd_bwd():
    a = ndc.clone().requires_grad_(True)
    f = feats.clone().requires_grad_(True)
    o = render_features(a, tf.to(dev), f, S, num_views=nv)[0]
    o.square().mean().backward()
    return a.grad
res = {}
for name, fn in (("features_forward_ms", f_fwd), ("features_forward_backward_ms", f_fwd_bwd)):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res[name] = sorted(ts)[len(ts) // 2]
res["channels"] = C
print(json.dumps(res, indent=1))
