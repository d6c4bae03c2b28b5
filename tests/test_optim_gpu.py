"""GPU parity of csrc/optim.cu (FusedAdam / FusedSGD) against torch.optim.Adam / torch.optim.SGD on CPU — the classes the
reference itself steps (/root/reference/src/train/train.py:161-168, trainer.py:122-123), so this oracle IS the reference.
Tolerance: rel-L2 <= 2e-6 on parameters and optimiser state after 25 steps (fp32, operation order differs only by FMA
contraction); state_dict interchange both ways."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _params(shapes, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(*s, generator=g) * 0.1 for s in shapes]


SHAPES = [(256, 515), (256,), (256, 256), (3, 256), (3,), (7, 100), (5, 9), (1,)]


def _run(opt_cpu, opt_gpu, cpu_ps, gpu_ps, steps, seed, sparse_rows=False):
    g = torch.Generator().manual_seed(seed)
    for it in range(steps):
        for pc, pg in zip(cpu_ps, gpu_ps):
            grad = torch.randn(pc.shape, generator=g) * (0.01 + 0.1 * (it % 3))
            if sparse_rows and pc.dim() == 2 and pc.shape[0] > 2:          # table rows outside the batch: zero grad, still move
                grad[it % pc.shape[0]] = 0
                grad[(it + 1) % pc.shape[0]] = 0
            pc.grad = grad.clone()
            pg.grad = grad.cuda()
        opt_cpu.step()
        opt_gpu.step()


@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_fused_adam_matches_torch(wd):
    from find_b200.optim import FusedAdam
    init = _params(SHAPES, 0)
    cpu_ps = [torch.nn.Parameter(p.clone()) for p in init]
    gpu_ps = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    oc = torch.optim.Adam(cpu_ps, lr=5e-4, weight_decay=wd)
    og = FusedAdam(gpu_ps, lr=5e-4, weight_decay=wd)
    _run(oc, og, cpu_ps, gpu_ps, 25, 1, sparse_rows=True)
    for pc, pg in zip(cpu_ps, gpu_ps):
        assert _rel(pg, pc) <= 2e-6
        assert _rel(og.state[pg]["exp_avg"], oc.state[pc]["exp_avg"]) <= 2e-6
        assert _rel(og.state[pg]["exp_avg_sq"], oc.state[pc]["exp_avg_sq"]) <= 2e-6
        assert float(og.state[pg]["step"]) == float(oc.state[pc]["step"]) == 25.0


def test_fused_sgd_momentum_matches_torch():
    from find_b200.optim import FusedSGD
    init = _params([(40, 9), (3,)], 2)
    cpu_ps = [torch.nn.Parameter(p.clone()) for p in init]
    gpu_ps = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    oc = torch.optim.SGD(cpu_ps, lr=1e-3, momentum=0.9)
    og = FusedSGD(gpu_ps, lr=1e-3, momentum=0.9)
    _run(oc, og, cpu_ps, gpu_ps, 25, 3, sparse_rows=True)
    for pc, pg in zip(cpu_ps, gpu_ps):
        assert _rel(pg, pc) <= 2e-6
        assert _rel(og.state[pg]["momentum_buffer"], oc.state[pc]["momentum_buffer"]) <= 2e-6


def test_state_dict_interchange_with_torch():
    """A checkpoint written by the fused class resumes in torch.optim.Adam and vice versa (same trajectory)."""
    from find_b200.optim import FusedAdam
    init = _params([(64, 33), (33,)], 4)
    cpu_ps = [torch.nn.Parameter(p.clone()) for p in init]
    gpu_ps = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    oc, og = torch.optim.Adam(cpu_ps, lr=1e-3), FusedAdam(gpu_ps, lr=1e-3)
    _run(oc, og, cpu_ps, gpu_ps, 5, 5)
    # fused -> torch (GPU torch Adam resumes from the fused state), torch -> fused
    sd_fused, sd_torch = copy.deepcopy(og.state_dict()), copy.deepcopy(oc.state_dict())
    assert set(sd_fused["state"][0].keys()) == set(sd_torch["state"][0].keys()) == {"step", "exp_avg", "exp_avg_sq"}
    gpu2 = [torch.nn.Parameter(p.detach().clone()) for p in gpu_ps]
    ot = torch.optim.Adam(gpu2, lr=1e-3)
    ot.load_state_dict(sd_fused)
    gpu3 = [torch.nn.Parameter(p.detach().clone().cuda()) for p in cpu_ps]
    of = FusedAdam(gpu3, lr=1e-3)
    of.load_state_dict(sd_torch)
    g = torch.Generator().manual_seed(9)
    for _ in range(5):
        for pc, p2, p3 in zip(cpu_ps, gpu2, gpu3):
            grad = torch.randn(pc.shape, generator=g) * 0.05
            pc.grad, p2.grad, p3.grad = grad.clone(), grad.cuda(), grad.cuda()
        oc.step(); ot.step(); of.step()
    for pc, p2, p3 in zip(cpu_ps, gpu2, gpu3):
        assert _rel(p2, pc) <= 3e-6 and _rel(p3, pc) <= 3e-6


def test_many_tensors_chunking_and_graph_capture():
    from find_b200.optim import FusedAdam
    init = _params([(17,)] * 60, 6)                       # more than one 48-tensor table
    cpu_ps = [torch.nn.Parameter(p.clone()) for p in init]
    gpu_ps = [torch.nn.Parameter(p.clone().cuda()) for p in init]
    oc, og = torch.optim.Adam(cpu_ps, lr=1e-2), FusedAdam(gpu_ps, lr=1e-2)
    _run(oc, og, cpu_ps, gpu_ps, 4, 7)
    for pc, pg in zip(cpu_ps, gpu_ps):
        assert _rel(pg, pc) <= 2e-6
    # capture one step of a single-table optimiser in a CUDA graph and replay it: the device step counter advances
    ps = [torch.nn.Parameter(torch.ones(1000, device="cuda"))]
    ps[0].grad = torch.full((1000,), 0.5, device="cuda")
    ref = torch.nn.Parameter(torch.ones(1000)); ref.grad = torch.full((1000,), 0.5)
    o, r = FusedAdam(ps, lr=1e-2), torch.optim.Adam([ref], lr=1e-2)
    o.step(); r.step()                                    # warm-up outside capture (allocates state)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        o.step()                                          # captured, not executed
    for _ in range(3):
        g.replay(); r.step()
    torch.cuda.synchronize()
    assert _rel(ps[0], ref) <= 2e-6


def test_cpu_params_refused():
    from find_b200.optim import FusedAdam
    p = torch.nn.Parameter(torch.zeros(3)); p.grad = torch.ones(3)
    with pytest.raises(RuntimeError):
        FusedAdam([p]).step()
