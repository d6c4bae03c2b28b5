"""GPU parity: the fused training step (findb200_train_step through find_b200.step.TrainStep) and the module-level
path (NeuralDisplacementField.get_meshes_from_batch + DisplacementLoss + autograd) against the CPU oracle of the
whole step (oracle/step_oracle.py) on the same seeded inputs.

Tolerances: loss / verts rel <= 1e-3 (fp16-operand forward, measured ~1e-4); Chamfer NN indices of the fused step are
bit-exact against the oracle evaluated ON THE GPU's own vertices (the index contract is a property of the Chamfer
kernel; vertices differ from the fp32 oracle's by forward rounding); gradients <= 5e-2 vs the fp32 oracle (ReLU-mask
flips of any reduced-precision forward, see tests/test_field_gpu.py) and <= 1e-4 between the two GPU paths.
"""
import numpy as np
import pytest
import torch

from oracle import chamfer_oracle as co
from oracle import field_oracle as fo
from oracle import step_oracle as so

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    a = torch.as_tensor(a).detach().double().cpu().reshape(-1)
    b = torch.as_tensor(b).detach().double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _setup(B, V, P2, n_rows, seed, frozen=False):
    from find_b200 import synthetic as syn
    from find_b200.field import NeuralDisplacementField
    dev = torch.device("cuda:0")
    model = NeuralDisplacementField(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=n_rows, val_size=1,
                                    shapevec_size=100, texvec_size=100, posevec_size=100, device=dev)
    p = fo.random_params(seed)
    sd = model.state_dict()
    with torch.no_grad():
        for k, v in p.items():
            sd[k].copy_(v)
    tv, tf = syn.make_template(V)
    model.set_template(tv, tf)
    lat = syn.make_latents(n_rows, 0)
    with torch.no_grad():
        for k in ("shapevec", "posevec", "texvec", "reg"):
            getattr(model, k).data.copy_(lat[k])
    g = torch.Generator().manual_seed(seed)
    idx = torch.randperm(n_rows, generator=g)[:B]
    target = syn.make_targets(B, P2, 0)
    colt = syn.make_col_target(B, V, 0)
    tables = {k: lat[k] for k in ("shapevec", "posevec", "texvec", "reg")}
    return dev, model, p, tv, tables, idx, target, colt


def _check_against_oracle(step, model, ref, p, idx, n_rows, frozen):
    total, chamf, tex = step.loss3.tolist()
    assert abs(chamf - ref["chamf"]) <= 1e-3 * ref["chamf"]
    assert abs(tex - ref["tex"]) <= 1e-3 * ref["tex"]
    assert abs(total - ref["loss"]) <= 1e-3 * ref["loss"]
    assert rel_l2(step.verts, ref["verts"]) < 1e-3
    assert rel_l2(step.col, ref["col"]) < 1e-3
    step.populate_grads()
    torch.cuda.synchronize()
    for k in ("shapevec", "posevec", "texvec", "reg"):
        g = getattr(model, k).data.grad
        assert rel_l2(g, ref["table_grads"][k]) < 5e-2, k
        untouched = torch.ones(n_rows, dtype=torch.bool)
        untouched[idx] = False
        if untouched.any():
            assert float(g[untouched.to(g.device)].abs().max()) == 0.0
    if not frozen:
        for name, prm in model.named_parameters():
            if name in p:
                assert rel_l2(prm.grad, ref["grads"][name]) < 5e-2, name


@pytest.mark.parametrize("use_graph", [False, True])
def test_train_step_matches_cpu_oracle(use_graph):
    from find_b200.step import TrainStep
    B, V, P2, n_rows = 3, 700, 600, 5
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=31)
    ref = so.train_step(p, fo.fourier_B(), tv, tables, idx, target, colt)
    step = TrainStep(model, B, P2, with_colour=True, use_graph=use_graph)
    step.load_inputs(idx, target, colt)
    step.run()
    if use_graph:
        step.run()                                   # a replay must give the same answer as the capture run
    torch.cuda.synchronize()
    _check_against_oracle(step, model, ref, p, idx, n_rows, frozen=False)
    # Chamfer index contract on the GPU's own vertices: bit-exact vs the C oracle (FMA-chain mode)
    ch = co.chamfer(step.verts.cpu().numpy(), target.numpy(), with_grad=False)
    ws = step.ws
    from find_b200 import knn_points
    k1 = knn_points(step.verts, target.to(dev))
    k2 = knn_points(target.to(dev), step.verts)
    assert np.array_equal(k1.idx[..., 0].cpu().numpy(), ch["idx_x"])
    assert np.array_equal(k2.idx[..., 0].cpu().numpy(), ch["idx_y"])
    assert np.array_equal(k1.dists[..., 0].cpu().numpy(), ch["dist_x"])


def test_train_step_frozen_mlp_latent_fitting():
    from find_b200.step import TrainStep
    B, V, P2, n_rows = 4, 300, 256, 6
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=32)
    ref = so.train_step(p, fo.fourier_B(), tv, tables, idx, target, colt, need_param_grads=False)
    step = TrainStep(model, B, P2, with_colour=True, frozen_mlp=True, use_graph=False)
    step.load_inputs(idx, target, colt)
    step.run()
    torch.cuda.synchronize()
    _check_against_oracle(step, model, ref, p, idx, n_rows, frozen=True)
    assert all(prm.grad is None for n, prm in model.named_parameters() if n in p)


def test_module_path_equals_fused_step():
    """model.get_meshes_from_batch + DisplacementLoss + autograd (the reference-shaped API) == findb200_train_step."""
    from find_b200.field import sample_latent_vectors
    from find_b200.losses import FieldChamferObjective
    from find_b200.step import TrainStep
    B, V, P2, n_rows = 3, 515, 400, 6
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=33)
    step = TrainStep(model, B, P2, with_colour=True, use_graph=False)
    step.load_inputs(idx, target, colt)
    step.run()
    step.populate_grads()
    torch.cuda.synchronize()
    fused = {n: prm.grad.clone() for n, prm in model.named_parameters() if prm.grad is not None}
    fused_loss = step.loss3.tolist()
    model.zero_grad(set_to_none=True)

    batch = {"idx": idx.to(dev), "points": target.to(dev), "col_target": colt.to(dev)}
    batch.update(sample_latent_vectors(batch, model.latent_vectors_train))
    loss, losses = FieldChamferObjective()(model, batch, is_train=True)
    loss.backward()
    assert abs(float(loss) - fused_loss[0]) <= 1e-5 * abs(fused_loss[0])
    assert abs(float(losses["chamf"]) - fused_loss[1]) <= 1e-6 * abs(fused_loss[1])
    for n, prm in model.named_parameters():
        if n in fused:
            assert prm.grad is not None, n
            assert rel_l2(prm.grad, fused[n]) < 1e-4, n


def test_displacement_loss_z_cutoff_variants_run_ragged():
    from find_b200.losses import DisplacementLoss
    from find_b200 import Pointclouds, chamfer_distance
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(5)
    verts = (torch.randn(2, 300, 3, generator=g) * 0.05).to(dev)
    pts = (torch.randn(2, 250, 3, generator=g) * 0.05).to(dev)
    L = DisplacementLoss()
    full = L(None, {"verts": verts}, {"points": pts}, 0, num_samples=None)["loss"]      # explicit opt-out of the 5000-sample default
    ref = co.chamfer(verts.cpu().numpy(), pts.cpu().numpy(), with_grad=False)["loss"]
    assert abs(float(full) - ref) <= 1e-6 * ref
    cut = L(None, {"verts": verts}, {"points": pts}, 0, num_samples=None, gt_z_cutoff=0.0)["loss"]
    lens = (pts[..., 2] <= 0).sum(1)
    padded = torch.zeros_like(pts)
    for b in range(2):
        padded[b, : int(lens[b])] = pts[b][pts[b, :, 2] <= 0]
    ref2 = co.chamfer(verts.cpu().numpy(), padded.cpu().numpy(), y_lengths=lens.cpu().numpy(), with_grad=False)["loss"]
    assert abs(float(cut) - ref2) <= 1e-6 * ref2
    both = L(None, {"verts": verts}, {"points": pts}, 0, num_samples=None, z_cutoff=0.0)["loss"]
    assert torch.isfinite(both)                    # vs the oracle: tests/test_parity_full_gpu.py::test_z_cutoff_both_sides_matches_oracle
    # the reference default (losses.py:61) is 5000 surface samples: without meshes in res / batch it must not silently fall back
    import inspect
    assert inspect.signature(L.forward).parameters["num_samples"].default == 5000


@pytest.mark.parametrize("use_graph", [False, True])
def test_train_step_mesh_terms_match_cpu_oracle(use_graph):
    """The shipped 3-D configuration (cfgs/train_3d.yaml: chamf on surface samples + smooth + tex) through
    findb200_train_step_mesh vs the CPU oracle of the same step.  The oracle makes its draw from the GPU's own vertices
    (draw parity itself is bit-exact, tests/test_mesh_gpu.py); the second run checks that the device draw counter advanced."""
    from find_b200 import synthetic as syn
    from find_b200.step import TrainStep
    B, V, P2, n_rows, S = 3, 642, 600, 5, 500
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=41)
    _, tf = syn.make_template(V)
    step = TrainStep(model, B, P2, with_colour=True, use_graph=use_graph, num_samples=S, weight_smooth=1000.0, seed=77)
    step.load_inputs(idx, target, colt)
    for it in range(2):
        step.run()
        torch.cuda.synchronize()
        # the draw that produced the current outputs (the first graph-mode run() executes a warm-up enqueue and a replay)
        counter = int(step.draw_counter.item()) - 1
        assert counter == (it + 1 if use_graph else it)
        ref = so.train_step(p, fo.fourier_B(), tv, tables, idx, target, colt, faces=tf.numpy(), num_samples=S, seed=77,
                            counter=counter, w_smooth=1000.0, draw_from_verts=step.verts.cpu().numpy())
        total, chamf, tex = step.loss3.tolist()
        edge, lap = step.loss_extra.tolist()
        assert abs(chamf - ref["chamf"]) <= 1e-3 * ref["chamf"], counter
        assert abs(edge - ref["smooth_edge"]) <= 1e-3 * ref["smooth_edge"]
        assert abs(lap - ref["smooth_lap"]) <= 1e-3 * ref["smooth_lap"]
        assert abs(total - ref["loss"]) <= 1e-3 * ref["loss"]
    model.zero_grad(set_to_none=True)
    step.populate_grads()
    torch.cuda.synchronize()
    for k in ("shapevec", "posevec", "texvec", "reg"):
        assert rel_l2(getattr(model, k).data.grad, ref["table_grads"][k]) < 5e-2, k
    for name, prm in model.named_parameters():
        if name in p:
            assert rel_l2(prm.grad, ref["grads"][name]) < 5e-2, name


def test_train_step_smooth_only_and_sampling_only():
    from find_b200 import synthetic as syn
    from find_b200.step import TrainStep
    B, V, P2, n_rows = 2, 642, 500, 4
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=42)
    _, tf = syn.make_template(V)
    base = TrainStep(model, B, P2, with_colour=True, use_graph=False)
    base.load_inputs(idx, target, colt); base.run()
    sm = TrainStep(model, B, P2, with_colour=True, use_graph=False, weight_smooth=1000.0)
    sm.load_inputs(idx, target, colt); sm.run()
    torch.cuda.synchronize()
    b, s = base.loss3.tolist(), sm.loss3.tolist()
    e, l = sm.loss_extra.tolist()
    assert s[1] == b[1] and s[2] == b[2]                       # same Chamfer / colour terms, bit for bit
    assert abs(s[0] - (b[0] + 1000.0 * (0.1 * l + 10.0 * e))) <= 1e-5 * abs(s[0])
    ref = so.train_step(p, fo.fourier_B(), tv, tables, idx, target, colt, faces=tf.numpy(), w_smooth=1000.0, forward_only=True)
    assert abs(e - ref["smooth_edge"]) <= 1e-3 * ref["smooth_edge"] and abs(l - ref["smooth_lap"]) <= 1e-3 * ref["smooth_lap"]
    sp = TrainStep(model, B, P2, with_colour=False, use_graph=False, num_samples=2000, seed=5)
    sp.load_inputs(idx, target); sp.run()
    torch.cuda.synchronize()
    ref2 = so.train_step(p, fo.fourier_B(), tv, tables, idx, target, None, faces=tf.numpy(), num_samples=2000, seed=5,
                         forward_only=True, draw_from_verts=sp.verts.cpu().numpy())
    assert abs(sp.loss3[1].item() - ref2["chamf"]) <= 1e-3 * ref2["chamf"]


def test_cfg1_full_size_forward_chamfer_matches_cpu_oracle():
    """BASELINE.json configs[0]: field forward + Chamfer, 1 synthetic foot, 10 242-vertex template vs 10 000 target points,
    at full size against the CPU oracle (the reference's own CPU-runnable case)."""
    from find_b200.step import TrainStep
    B, V, P2, n_rows = 1, 10242, 10000, 2
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=51)
    ref = so.train_step(p, fo.fourier_B(), tv, tables, idx, target, colt, forward_only=True, threads=8)
    step = TrainStep(model, B, P2, with_colour=True, use_graph=False)
    step.load_inputs(idx, target, colt)
    step.run()
    torch.cuda.synchronize()
    total, chamf, tex = step.loss3.tolist()
    assert abs(chamf - ref["chamf"]) <= 1e-3 * ref["chamf"]
    assert abs(tex - ref["tex"]) <= 1e-3 * ref["tex"]
    assert rel_l2(step.verts, ref["verts"]) < 1e-3 and rel_l2(step.disp, ref["disp"]) < 1e-3
    ch = co.chamfer(step.verts.cpu().numpy(), target.numpy(), with_grad=False, threads=8)
    from find_b200 import knn_points
    k1 = knn_points(step.verts, target.to(dev))
    assert np.array_equal(k1.idx[..., 0].cpu().numpy(), ch["idx_x"])
    assert np.array_equal(k1.dists[..., 0].cpu().numpy(), ch["dist_x"])


def test_cfg2_full_size_batch_additivity():
    """Size-independent property at BASELINE.json configs[1] (8 feet x 10 242 vertices x 10 000 targets): the loss is a mean
    over feet, so the MLP-gradient bucket of the 8-foot step equals the mean of the buckets of its two 4-foot halves, and the
    per-foot latent gradients of a half are those of the full step times 2 (fp32 summation order only)."""
    from find_b200.step import TrainStep
    B, V, P2 = 8, 10242, 10000
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, B, seed=52)
    idx = torch.arange(B)
    full = TrainStep(model, B, P2, with_colour=True, use_graph=False)
    full.load_inputs(idx, target, colt)
    full.run()
    half = TrainStep(model, 4, P2, with_colour=True, use_graph=False)
    acc, rows = torch.zeros_like(full.flat_grad), []
    for h in range(2):
        sl = slice(4 * h, 4 * h + 4)
        half.load_inputs(idx[sl], target[sl], colt[sl])
        half.run()
        acc += half.flat_grad
        rows.append(half.grows["shapevec"].clone())
        l_half = half.loss3.clone()
        assert torch.isfinite(l_half).all()
    torch.cuda.synchronize()
    assert rel_l2(acc / 2, full.flat_grad) < 2e-3          # fp16 gradient scale differs between the runs (global max)
    assert rel_l2(torch.cat(rows) / 2, full.grows["shapevec"]) < 2e-3
    assert torch.isfinite(full.loss3).all() and float(full.loss3[1]) > 0


def test_workspace_caches_match_recompute_and_follow_template_changes():
    """FINDB200_FIELD_ENC_CACHED / TRUNK_CACHED: the second and later enqueues reuse the encoded template tiles (and, with a
    frozen MLP, the trunk output) left in the workspace — results must be bit-identical to the first (recomputing) enqueue,
    and an in-place change of the template must be picked up (version counter) instead of served from the cache."""
    from find_b200.step import TrainStep
    for frozen in (False, True):
        B, V, P2, n_rows = 3, 700, 600, 5
        dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=61)
        step = TrainStep(model, B, P2, with_colour=True, use_graph=False, frozen_mlp=frozen)
        step.load_inputs(idx, target, colt)
        step.run(); torch.cuda.synchronize()
        first = (step.loss3.clone(), step.verts.clone(), step.col.clone(), step.flat_grad.clone(), step.grows["shapevec"].clone())
        assert step._cparams.flags == 0
        step.run(); torch.cuda.synchronize()
        assert step._cparams.flags == (3 if frozen else 1)
        second = (step.loss3, step.verts, step.col, step.flat_grad, step.grows["shapevec"])
        for a, b in zip(first[:3], second[:3]):
            assert torch.equal(a, b)
        for a, b in zip(first[3:], second[3:]):                  # gradients: fp32 atomics in the column sums
            assert rel_l2(b, a) < 1e-5
        with torch.no_grad():
            model.template_verts.mul_(1.01)                      # in-place edit bumps the version counter
        step.run(); torch.cuda.synchronize()
        assert step._cparams.flags == 0
        fresh = TrainStep(model, B, P2, with_colour=True, use_graph=False, frozen_mlp=frozen)
        fresh.load_inputs(idx, target, colt)
        fresh.run(); torch.cuda.synchronize()
        assert torch.equal(fresh.verts, step.verts) and torch.equal(fresh.loss3, step.loss3)
        assert not torch.equal(step.verts, first[1])


@pytest.mark.parametrize("use_graph", [False, True])
def test_pipelined_host_feed_matches_serial_steps(use_graph):
    """iterate_from_host (batch k+1 staged on a copy stream while step k runs) gives, batch for batch, the losses and the
    gradients of step_from_host; staging more than two batches ahead is refused."""
    from find_b200 import synthetic as syn
    from find_b200.step import TrainStep
    B, V, P2, n_rows = 3, 700, 600, 9
    dev, model, p, tv, tables, idx, target, colt = _setup(B, V, P2, n_rows, seed=71)
    batches = []
    for k in range(4):
        g = torch.Generator().manual_seed(100 + k)
        batches.append((torch.randperm(n_rows, generator=g)[:B].pin_memory(), syn.make_targets(B, P2, 10 * k).pin_memory(),
                        syn.make_col_target(B, V, 10 * k).pin_memory()))
    step = TrainStep(model, B, P2, with_colour=True, use_graph=use_graph)
    serial, grads = [], []
    for b in batches:
        serial.append(step.step_from_host(*b))
        grads.append(step.flat_grad.clone())
    assert len({round(s["loss"], 6) for s in serial}) == 4                 # the batches really differ
    piped = list(step.iterate_from_host(batches))
    assert len(piped) == 4
    for a, b in zip(serial, piped):
        for k in ("loss", "chamf", "tex"):
            assert abs(a[k] - b[k]) <= 1e-6 * abs(a[k])
    assert rel_l2(step.flat_grad, grads[-1]) < 1e-5
    step.prefetch_from_host(*batches[0]); step.prefetch_from_host(*batches[1])
    with pytest.raises(RuntimeError):
        step.prefetch_from_host(*batches[2])
    assert abs(step.step_prefetched()["loss"] - serial[0]["loss"]) <= 1e-6 * serial[0]["loss"]
    assert abs(step.step_prefetched()["loss"] - serial[1]["loss"]) <= 1e-6 * serial[1]["loss"]
    with pytest.raises(RuntimeError):
        step.step_prefetched()
