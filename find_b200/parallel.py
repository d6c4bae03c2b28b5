"""find_b200/parallel.py — foot-batch data parallelism (one process per GPU, torch.distributed).

The reference has no distributed code: its only multi-GPU mechanism is one independent experiment per GPU
(/root/reference/src/train/run_expmt.py:80-99).  The B200 path shards the FEET of a batch across ranks
(SURVEY.md §8(e)): field outputs, registration, Chamfer and every latent / reg gradient of foot b depend on foot b
only, so the single cross-rank quantity is the MLP weight gradient — ONE allreduce(sum) of the flat 868 358-float
bucket ``TrainStep.flat_grad`` per step (NCCL over NVLink on GPUs, gloo in the CPU tests).  Latent and registration
tables stay rank-local: a row is only ever touched by the rank that owns its foot.  Frozen-MLP latent fitting
(train.py:218-224) needs no collective at all.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) slice of ``n_items`` feet for ``rank`` (first ``n % world`` ranks get one more)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def owner_by_label(labels: Sequence[str], world: int) -> Dict[str, int]:
    """Owner rank of every distinct foot ID: all scans that share a shape/tex row (``use_latent_labels``,
    /root/reference/src/data/dataset.py:191-201) land on one rank, so latent tables never need a collective.
    Deterministic: distinct labels in first-appearance order, dealt out in balanced contiguous blocks."""
    uniq: List[str] = []
    seen = set()
    for lab in labels:
        if lab not in seen:
            seen.add(lab)
            uniq.append(lab)
    out = {}
    for r in range(world):
        lo, hi = shard_range(len(uniq), r, world)
        for lab in uniq[lo:hi]:
            out[lab] = r
    return out


def shard_by_label(labels: Sequence[str], rank: int, world: int) -> List[int]:
    """Indices (into ``labels``) of the batch items this rank owns."""
    own = owner_by_label(labels, world)
    return [i for i, lab in enumerate(labels) if own[lab] == rank]


def allreduce_flat_grad(flat_grad: torch.Tensor, group=None, average: bool = False) -> Optional["dist.Work"]:
    """SUM the MLP-gradient bucket over ranks in place (the only data-path collective).  ``average`` divides by the
    world size afterwards (the loss of a sharded batch is a mean over ALL feet; callers that already scaled their
    local loss by 1/world pass False)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
    if average:
        flat_grad.div_(dist.get_world_size(group))
    return None


class ShardedTrainStep:
    """TrainStep over this rank's slice of a global foot batch + the gradient allreduce.

    The global objective is a mean over all ``global_feet`` feet; each rank computes the mean over its
    ``local_feet`` and the bucket is combined with weights ``local_feet / global_feet`` (a plain average when the
    split is even), which reproduces the single-rank gradient of the concatenated batch."""

    def __init__(self, step, global_feet: int, group=None):
        self.step = step
        self.group = group
        self.global_feet = int(global_feet)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.weight = step.B / float(self.global_feet)
        self.graph: Optional[torch.cuda.CUDAGraph] = None

    def _collective(self):
        """weight * flat_grad summed over ranks, in place.  An even split is NCCL's AVG (one kernel, no pre-scale)."""
        fg = self.step.flat_grad
        if fg.is_cuda and abs(self.weight * self.world - 1.0) < 1e-12:
            dist.all_reduce(fg, op=dist.ReduceOp.AVG, group=self.group)
            return
        if self.weight != 1.0:
            fg.mul_(self.weight)
        dist.all_reduce(fg, op=dist.ReduceOp.SUM, group=self.group)

    def run(self):
        """One step of this rank's shard + the gradient allreduce.  With ``step.use_graph`` the kernels of the step AND
        the NCCL allreduce are captured in ONE CUDA graph (a single replay per step: no host launch between the last
        backward kernel and the collective)."""
        step = self.step
        if self.world == 1 or step.frozen:
            return step.run()
        if not step.use_graph:
            step.run()
            self._collective()
            return step.loss3
        with torch.cuda.device(step.dev):
            if self.graph is None:
                step._enqueue()                      # warm-up outside capture: function attributes, NCCL communicator
                self._collective()
                torch.cuda.synchronize(step.dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    step._enqueue()
                    self._collective()
                self.graph = g
            self.graph.replay()
        return step.loss3

    def scale_local_grads(self):
        """Latent / reg row gradients are rank-local; scale them to the global mean as well."""
        if self.weight != 1.0:
            for g in self.step.grows.values():
                g.mul_(self.weight)


__all__ = ["shard_range", "owner_by_label", "shard_by_label", "allreduce_flat_grad", "ShardedTrainStep"]
