"""tools/dp_check.py — multi-GPU correctness of the foot-batch sharding on real GPUs (run under torchrun, N >= 2):
every rank runs ShardedTrainStep (graph with the NCCL allreduce captured) on its slice of a global batch; rank 0 also runs
the single-rank TrainStep on the whole batch; the allreduced MLP-gradient bucket must equal the single-rank one (fp32
summation order only) and the rank-local latent gradients must equal the matching rows.  Prints one JSON line on rank 0."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import synthetic as syn                      # noqa: E402
from find_b200.field import NeuralDisplacementField        # noqa: E402
from find_b200.parallel import ShardedTrainStep            # noqa: E402
from find_b200.step import TrainStep                       # noqa: E402


def build(dev, G, V):
    torch.manual_seed(1)
    model = NeuralDisplacementField(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=G, val_size=1,
                                    shapevec_size=100, texvec_size=100, posevec_size=100, device=dev)
    tv, tf = syn.make_template(V)
    model.set_template(tv, tf)
    syn.randomise_last_disp_layer(model.mlp_disp[-1], seed=0)
    lat = syn.make_latents(G, 0)
    with torch.no_grad():
        for k in ("shapevec", "posevec", "texvec", "reg"):
            getattr(model, k).data.copy_(lat[k].to(dev))
    return model


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", lr)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    per, V, P2 = 2, 2562, 3000
    G = per * world
    model = build(dev, G, V)
    target, colt = syn.make_targets(G, P2, 0), syn.make_col_target(G, V, 0)
    lo = rank * per
    step = TrainStep(model, per, P2, with_colour=True, use_graph=True)
    sh = ShardedTrainStep(step, G, collective=os.environ.get("FINDB200_COLLECTIVE", "auto"))
    step.load_inputs(torch.arange(lo, lo + per), target[lo:lo + per], colt[lo:lo + per])
    for _ in range(3):                                       # capture + replays
        sh.run()
    torch.cuda.synchronize()
    flat = step.flat_grad.clone()
    # every rank must hold the same bucket
    ref0 = flat.clone()
    dist.broadcast(ref0, 0)
    same = bool(torch.equal(ref0, flat))
    gathered = [torch.zeros_like(step.grows["shapevec"]) for _ in range(world)]
    dist.all_gather(gathered, step.grows["shapevec"] * sh.weight)
    out = None
    if rank == 0:
        model1 = build(dev, G, V)
        full = TrainStep(model1, G, P2, with_colour=True, use_graph=False)
        full.load_inputs(torch.arange(G), target, colt)
        full.run()
        torch.cuda.synchronize()
        out = dict(world=world, collective=sh.collective, bucket_rel_l2=rel(flat, full.flat_grad), shapevec_rows_rel_l2=rel(torch.cat(gathered), full.grows["shapevec"]))
    oks = [None] * world
    dist.all_gather_object(oks, same)
    if rank == 0:
        out["all_ranks_identical"] = all(oks)
        out["ok"] = bool(out["bucket_rel_l2"] < 1e-4 and out["shapevec_rows_rel_l2"] < 1e-4 and out["all_ranks_identical"])
        print(json.dumps(out))
    sys.stdout.flush()
    # no destroy_process_group(): tearing NCCL down after a graph that captured its collective can block; leave hard
    dist.barrier()
    os._exit(0 if (rank != 0 or out["ok"]) else 1)


if __name__ == "__main__":
    main()
