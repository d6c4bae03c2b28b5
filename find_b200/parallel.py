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


class P2PAllReduce:
    """The MLP-gradient bucket in symmetric memory + findb200_allreduce_p2p (csrc/allreduce.cu): ONE kernel over NVLink /
    NVSwitch peer memory instead of an NCCL call.  torch's symmetric-memory rendezvous is used only to allocate the bucket and
    to exchange the peer mappings (plumbing); the reduction itself is the library's kernel.  ``bucket`` is the fp32 view of
    length ``n`` the step writes its gradients into."""

    def __init__(self, n: int, device: torch.device, group=None, n_ctas: int = 128):
        import ctypes
        import torch.distributed._symmetric_memory as symm_mem
        from . import _lib
        self.lib = _lib.load()
        self._lib_mod = _lib
        group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.n, self.n_pad = int(n), (int(n) + 3) // 4 * 4
        self.dev, self.n_ctas = device, int(n_ctas)
        flag_words = int(self.lib.findb200_allreduce_p2p_flag_bytes()) // 4
        with torch.cuda.device(device):
            self.buf = symm_mem.empty(self.n_pad, dtype=torch.float32, device=device)
            self.flags = symm_mem.empty(flag_words, dtype=torch.int32, device=device)
            self.buf.zero_()
            self.flags.zero_()
            torch.cuda.synchronize(device)
            self._h_buf = symm_mem.rendezvous(self.buf, group)
            self._h_flags = symm_mem.rendezvous(self.flags, group)
        bp, fp = list(self._h_buf.buffer_ptrs), list(self._h_flags.buffer_ptrs)
        if len(bp) != self.world or len(fp) != self.world or not all(bp) or not all(fp):
            raise RuntimeError("symmetric-memory rendezvous did not map every peer")
        self._bufs = (ctypes.c_void_p * self.world)(*bp)
        self._flags = (ctypes.c_void_p * self.world)(*fp)
        # NVLS: the bucket's multicast mapping, when the ranks share a multicast object (0 otherwise)
        self.mc_ptr = int(getattr(self._h_buf, "multicast_ptr", 0) or 0)
        self.use_mc = False                        # set by ShardedTrainStep once EVERY rank has a multicast mapping
        self.epoch = torch.zeros(4, dtype=torch.int32, device=device)   # [calls done, CTAs done, error code, culprit rank]
        self.bucket = self.buf[: self.n]
        dist.barrier(group)                       # every rank's flag page is zero before anyone signals
        torch.cuda.synchronize(device)

    def enqueue(self, scale: float = 1.0):
        """bucket <- scale * sum over ranks (in place, on the current stream; graph-capturable)."""
        _lib = self._lib_mod
        if self.use_mc:                            # the sum is taken inside the NVSwitch (csrc/allreduce.cu, allreduce_mc_kernel)
            import ctypes
            rc = self.lib.findb200_allreduce_mc(ctypes.c_void_p(self.mc_ptr), self._flags, self.rank, self.world, self.n_pad,
                                                float(scale), _lib.ptr(self.epoch), self.n_ctas, _lib.stream_ptr(self.dev))
            _lib.check(rc, "findb200_allreduce_mc")
            return
        rc = self.lib.findb200_allreduce_p2p(self._bufs, self._flags, self.rank, self.world, self.n_pad, float(scale),
                                             _lib.ptr(self.epoch), self.n_ctas, _lib.stream_ptr(self.dev))
        _lib.check(rc, "findb200_allreduce_p2p")

    def check(self):
        """Synchronising health check: raises if a wait of the kernel gave up (a rank missing for FINDB200_P2P_TIMEOUT_S seconds,
        default 600) or if the ranks were found out of lock-step.  The context stays usable, so a caller may switch the
        ShardedTrainStep to NCCL (``collective="nccl"``) for skew-prone loops."""
        code, who = (int(x) for x in self.epoch[2:4].tolist())
        if code:
            what = {1: f"rank {who} never arrived", 2: f"rank {who} enqueued one call more than rank {self.rank}"}.get(code, "unknown")
            raise RuntimeError(f"findb200_allreduce_p2p failed on rank {self.rank}: {what}")


class ShardedTrainStep:
    """TrainStep over this rank's slice of a global foot batch + the gradient allreduce.

    The global objective is a mean over all ``global_feet`` feet; each rank computes the mean over its
    ``local_feet`` and the bucket is combined with weights ``local_feet / global_feet`` (a plain average when the
    split is even), which reproduces the single-rank gradient of the concatenated batch."""

    def __init__(self, step, global_feet: int, group=None, collective: str = "auto"):
        """``collective``: "p2p" = the library's own peer-memory allreduce kernel (csrc/allreduce.cu), "nvls" = the same
        exchange with the sum taken inside the NVSwitch (multimem.ld_reduce / multimem.st on the bucket's multicast mapping;
        falls back to "p2p" when a rank has no multicast mapping), "nccl" = torch.distributed.all_reduce, "auto" = p2p when
        every rank can map its peers (symmetric memory; nvls from 8 ranks on when every rank has the multicast mapping), else
        nccl."""
        self.step = step
        self.group = group
        self.global_feet = int(global_feet)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.weight = step.B / float(self.global_feet)
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self._graphs = {}
        self.p2p: Optional[P2PAllReduce] = None
        self.collective = "none"
        if self.world > 1 and getattr(step, "mesh_terms", None) is not None:
            # every rank must draw its OWN surface samples: fold the rank into the key of the Philox stream
            rank = dist.get_rank(group)
            step.mesh_terms.seed = (int(step.mesh_terms.seed) + 0x9E3779B97F4A7C15 * (rank + 1)) & (2 ** 64 - 1)
        if self.world > 1 and not step.frozen:
            self.collective = "nccl"
            if collective in ("auto", "p2p", "nvls") and step.flat_grad.is_cuda:
                err = None
                try:
                    self.p2p = P2PAllReduce(step.flat_grad.numel(), step.dev, group)
                except Exception as e:              # noqa: BLE001 — no peer mapping on this box: NCCL carries the exchange
                    err = e
                ok = torch.tensor([1 if self.p2p is not None else 0], device=step.dev)
                dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
                if int(ok.item()) == 1:
                    self._adopt_bucket()
                    self.collective = "p2p"
                    # "auto": the in-switch sum pays from 8 ranks on (isolated 22.4 us against 25.2 us at N = 8, equal at N = 4,
                    # 23.1 against 19.5 us at N = 2: profiles/r2_allreduce_nvls.md)
                    if collective == "nvls" or (collective == "auto" and self.world >= 8):
                        mc = torch.tensor([1 if self.p2p.mc_ptr else 0], device=step.dev)
                        dist.all_reduce(mc, op=dist.ReduceOp.MIN, group=group)
                        if int(mc.item()) == 1:
                            self.p2p.use_mc = True
                            self.collective = "nvls"
                else:
                    self.p2p = None
                    if collective in ("p2p", "nvls"):
                        raise RuntimeError(f"peer-memory allreduce unavailable: {err}")

    def _adopt_bucket(self):
        """Make the step write its MLP gradients straight into the symmetric bucket (must happen before its first enqueue)."""
        step = self.step
        if step._cparams is not None:
            raise RuntimeError("ShardedTrainStep must wrap a TrainStep before its first run")
        step.flat_grad = self.p2p.bucket
        step.grad_views, off = [], 0
        for prm in step.mlp_params:
            step.grad_views.append(step.flat_grad[off:off + prm.numel()].view_as(prm))
            off += prm.numel()

    def _collective(self):
        """weight * flat_grad summed over ranks, in place.  An even split is NCCL's AVG (one kernel, no pre-scale)."""
        if self.p2p is not None:
            self.p2p.enqueue(self.weight)
            return
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
            if step.graphs_stale():                  # template changed: recompute the cached tiles, re-capture
                self._graphs = {}
            self.graph = self._graphs.get(step._cur)         # one graph per input set of the step (pipelined host feed)
            if self.graph is None:
                step._enqueue()                      # warm-up outside capture: function attributes, NCCL communicator
                self._collective()
                torch.cuda.synchronize(step.dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    step._enqueue()
                    self._collective()
                self.graph = self._graphs[step._cur] = g
            self.graph.replay()
        return step.loss3

    def scale_local_grads(self):
        """Latent / reg row gradients are rank-local; scale them to the global mean as well."""
        if self.weight != 1.0:
            for g in self.step.grows.values():
                g.mul_(self.weight)
            for g in self.step.table_grads.values():
                g.mul_(self.weight)


__all__ = ["shard_range", "owner_by_label", "shard_by_label", "allreduce_flat_grad", "ShardedTrainStep", "P2PAllReduce"]
