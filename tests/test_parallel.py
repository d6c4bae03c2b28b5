"""CPU: foot-batch data parallelism bookkeeping (find_b200/parallel.py) — world_size-2 gloo run.

The sharded step on 2 ranks must reproduce the single-rank step on the concatenated batch: allreduced MLP gradients
equal within fp32 summation tolerance, latent/reg gradients equal row-wise and stay rank-local.  The per-rank compute
here is the CPU oracle (this container has no GPU); on the GPU box the same bookkeeping wraps TrainStep."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from find_b200.parallel import allreduce_flat_grad, owner_by_label, shard_by_label, shard_range


def test_shard_range_partitions():
    for n in (0, 1, 7, 8, 64, 255):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_owner_by_label_keeps_shared_rows_on_one_rank():
    labels = ["f3", "f1", "f3", "f2", "f1", "f9", "f3"]
    own = owner_by_label(labels, 2)
    assert set(own) == {"f1", "f2", "f3", "f9"} and set(own.values()) == {0, 1}
    parts = [shard_by_label(labels, r, 2) for r in range(2)]
    assert sorted(parts[0] + parts[1]) == list(range(len(labels)))
    for lab in set(labels):
        ranks = {r for r in range(2) for i in parts[r] if labels[i] == lab}
        assert len(ranks) == 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _problem():
    from find_b200 import synthetic as syn
    from oracle import field_oracle as fo
    B, V, P2 = 4, 96, 80
    p = fo.random_params(3)
    tv, _ = syn.make_template(V)
    lat = syn.make_latents(B, 0)
    tables = {k: lat[k] for k in ("shapevec", "posevec", "texvec", "reg")}
    return B, V, P2, p, fo.fourier_B(), tv, tables, syn.make_targets(B, P2, 0), syn.make_col_target(B, V, 0)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        from oracle import field_oracle as fo
        from oracle import step_oracle as so
        B, V, P2, p, Bm, tv, tables, target, colt = _problem()
        lo, hi = shard_range(B, rank, world)
        idx = torch.arange(lo, hi)
        out = so.train_step(p, Bm, tv, tables, idx, target[lo:hi], colt[lo:hi])
        flat = torch.cat([out["grads"][k].reshape(-1) for k in fo.WEIGHT_KEYS])
        flat.mul_((hi - lo) / B)                       # ShardedTrainStep.weight
        allreduce_flat_grad(flat)
        q.put((rank, flat.numpy(), {k: (v * ((hi - lo) / B)).numpy() for k, v in out["table_grads"].items()}, (lo, hi)))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_matches_single_rank():
    from oracle import field_oracle as fo
    from oracle import step_oracle as so
    torch.set_num_threads(1)
    B, V, P2, p, Bm, tv, tables, target, colt = _problem()
    ref = so.train_step(p, Bm, tv, tables, torch.arange(B), target, colt)
    ref_flat = torch.cat([ref["grads"][k].reshape(-1) for k in fo.WEIGHT_KEYS]).numpy()

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=240) for _ in procs]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    for rank, flat, tgrads, (lo, hi) in res:
        rel = np.linalg.norm(flat - ref_flat) / np.linalg.norm(ref_flat)
        assert rel < 1e-5, rel                          # only the fp32 summation order differs
        for k, g in tgrads.items():
            np.testing.assert_allclose(g[lo:hi], ref["table_grads"][k][lo:hi].numpy(), rtol=2e-4, atol=1e-7)
            other = np.delete(g, np.arange(lo, hi), axis=0)
            assert not other.any()                      # rows of feet owned by the other rank are never touched


@pytest.mark.gpu
@pytest.mark.parametrize("collective", ["p2p", "nvls", "nccl"])
def test_sharded_step_on_two_gpus_matches_single_rank(collective):
    """Two real ranks (torchrun, NCCL rendezvous, the exchange inside the step graph: the library's peer-memory kernel, its NVLS
    variant — which falls back to the peer-memory kernel when the ranks share no multicast object — or NCCL) against the single-rank
    step on the concatenated batch: tools/dp_check.py.  Needs >= 2 visible GPUs; skipped otherwise (the driver's GPU box has one)."""
    import json
    import subprocess
    import sys
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tools", "dp_check.py")]
    res = subprocess.run(cmd, cwd=root, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600,
                         env=dict(os.environ, FINDB200_COLLECTIVE=collective))
    assert res.returncode == 0, res.stderr[-2000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    assert out["ok"] and out["all_ranks_identical"] and out["bucket_rel_l2"] < 1e-4 and out["shapevec_rows_rel_l2"] < 1e-4, out
    assert out["collective"] in ((collective, "p2p") if collective == "nvls" else (collective,)), out
