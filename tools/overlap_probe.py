"""tools/overlap_probe.py — do the field's head kernels (HBM-store-bound) and the kNN (issue-bound) overlap when they run on two
streams?  Times the field forward (B feet) and the Chamfer forward (sort + kNN) alone and together.  GPU only."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from find_b200 import _lib, synthetic as syn  # noqa: E402
from find_b200.field_ops import field_apply  # noqa: E402
from oracle import field_oracle as fo  # noqa: E402  (parameters only; a probe, not the product)

lib = _lib.load()
dev = torch.device("cuda:0")
B, V, P2 = int(os.environ.get("OP_B", 8)), 10242, 10000
g = torch.Generator().manual_seed(1)
Bm = fo.fourier_B().to(dev)
tv, _ = syn.make_template(V)
V = tv.shape[0]
pos = tv[None].contiguous().to(dev)
zs, zp, zt = (torch.randn(B, 100, generator=g).to(dev) * 0.1 for _ in range(3))
avg = torch.rand(3, generator=g).to(dev) * 0.1
p = fo.random_params(3)
ws_ = [w.to(dev) for w in p["W"]] if isinstance(p, dict) and "W" in p else None
if ws_ is None:
    from test_field_gpu import _params_from
    ws_, bs_ = _params_from(p, dev, requires_grad=False)
else:
    bs_ = [b.to(dev) for b in p["b"]]

x = (tv[None] * (1 + 0.02 * torch.randn(B, 1, 3))).contiguous().to(dev)
y = syn.make_targets(B, P2, 0).to(dev)
nb = lib.findb200_knn_workspace_bytes(B, V, P2)
kws = torch.empty(nb, dtype=torch.uint8, device=dev)
loss = torch.empty((), device=dev)
ix = torch.empty((B, V), dtype=torch.int64, device=dev)
iy = torch.empty((B, P2), dtype=torch.int64, device=dev)
sA, sB = torch.cuda.Stream(), torch.cuda.Stream()


def field():
    with torch.no_grad(), torch.cuda.stream(sA):
        field_apply(pos, zs, zp, zt, None, Bm, avg, ws_, bs_)


def knn():
    with torch.cuda.stream(sB):
        rc = lib.findb200_chamfer_forward(_lib.ptr(x), _lib.ptr(y), None, None, B, V, P2, _lib.ptr(loss), None, _lib.ptr(ix),
                                          None, _lib.ptr(iy), _lib.ptr(kws), nb, _lib.stream_ptr(dev))
        _lib.check(rc, "chamfer_forward")


def timed(fns, iters=30):
    ts = []
    for _ in range(iters):
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True)
        eA, eB = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sA.wait_event(e0); sB.wait_event(e0)
        for f in fns:
            f()
        eA.record(sA); eB.record(sB)
        torch.cuda.synchronize()
        ts.append(max(e0.elapsed_time(eA), e0.elapsed_time(eB)))
    ts.sort()
    return ts[len(ts) // 2]


for _ in range(3):
    field(); knn()
torch.cuda.synchronize()
print("B", B, "field alone %.4f ms, chamfer fwd alone %.4f ms" % (timed([field]), timed([knn])))
print("field then knn: %.4f ms;  knn then field: %.4f ms" % (timed([field, knn]), timed([knn, field])))


def graph_of(fns):
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        cur = torch.cuda.current_stream()
        sA.wait_stream(cur); sB.wait_stream(cur)
        for f in fns:
            f()
        cur.wait_stream(sA); cur.wait_stream(sB)
    return gr


def timed_graph(gr, iters=50):
    for _ in range(3):
        gr.replay()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(); gr.replay(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


# ---- the field BACKWARD (head_bwd, sum_feet, trunk_bwd, dW: store / HBM-bound) beside the kNN
ws_g = [w.clone().requires_grad_(True) for w in ws_]
bs_g = [b.clone().requires_grad_(True) for b in bs_]
zg = [z.clone().requires_grad_(True) for z in (zs, zp, zt)]
with torch.cuda.stream(sA):
    outs = field_apply(pos, zg[0], zg[1], zg[2], None, Bm, avg, ws_g, bs_g)
    gouts = [torch.randn_like(o) * 1e-4 for o in outs]
torch.cuda.synchronize()


def fbwd():
    with torch.cuda.stream(sA):
        torch.autograd.backward(list(outs), gouts, retain_graph=True)


for _ in range(3):
    fbwd()
torch.cuda.synchronize()
print("backward alone %.4f ms; backward then knn %.4f; knn then backward %.4f ms" % (timed([fbwd]), timed([fbwd, knn]), timed([knn, fbwd])))
gb, gbk, gkb = graph_of([fbwd]), graph_of([fbwd, knn]), graph_of([knn, fbwd])
print("graphs: backward %.4f, backward+knn %.4f, knn+backward %.4f ms" % (timed_graph(gb), timed_graph(gbk), timed_graph(gkb)))

gf, gk, gfk, gkf = graph_of([field]), graph_of([knn]), graph_of([field, knn]), graph_of([knn, field])
print("graphs: field %.4f, chamfer fwd %.4f, field+knn %.4f, knn+field %.4f ms" % (timed_graph(gf), timed_graph(gk), timed_graph(gfk), timed_graph(gkf)))
