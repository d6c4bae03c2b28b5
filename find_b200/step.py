"""find_b200/step.py — FIND's 3-D training step as ONE enqueue of sm_100a kernels (findb200_train_step).

What it replaces, per step of ``Trainer.train_epoch`` (/root/reference/src/train/trainer.py:97-121) with the
``chamf`` (+ colour) losses on:
    batch.update(sample_latent_vectors(...))      trainer.py:97      -> findb200_gather_rows (4 tables)
    loss, loss_dict = self.model(batch, ...)      trainer.py:106     -> field fwd + reg transform + Chamfer (+ colour MSE)
    loss.backward()                               trainer.py:121     -> Chamfer bwd + field bwd (+ latent scatter-add)
The optimiser step stays with torch (``.grad`` of every parameter is populated exactly like autograd would).

``TrainStep`` owns static device buffers, so the enqueue can be captured once in a CUDA graph and replayed; the
per-step inputs are the foot indices, the target clouds and (optionally) colour targets.  No CPU path exists.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import torch

from . import _lib
from .field import NeuralDisplacementField
from .field_ops import FIELD_ENC_CACHED, FIELD_TRUNK_CACHED, FieldGradsC, N_PARAMS, _make_params
from .losses import CHAMF_WEIGHT, TEX_WEIGHT

SMOOTH_WEIGHT = 1000.0   # opts.py:97 (--weight_smooth default)


class GatherJobC(ctypes.Structure):             # findb200_gather_job (include/find_b200.h)
    _fields_ = [("table", ctypes.c_void_p), ("n_rows", ctypes.c_int), ("width", ctypes.c_int), ("out", ctypes.c_void_p),
                ("idx", ctypes.c_void_p)]


class ScatterJobC(ctypes.Structure):            # findb200_scatter_job (include/find_b200.h)
    _fields_ = [("g", ctypes.c_void_p), ("n_rows", ctypes.c_int), ("width", ctypes.c_int), ("grad_table", ctypes.c_void_p),
                ("idx", ctypes.c_void_p)]


TABLE_KEYS = ("shapevec", "posevec", "texvec", "reg")


class MeshTermsC(ctypes.Structure):             # findb200_mesh_terms (include/find_b200.h)
    _fields_ = [("faces", ctypes.c_void_p), ("F", ctypes.c_int), ("edges", ctypes.c_void_p), ("E", ctypes.c_int),
                ("vf_ptr", ctypes.c_void_p), ("vf_face", ctypes.c_void_p), ("vf_corner", ctypes.c_void_p), ("nnz", ctypes.c_int),
                ("num_samples", ctypes.c_int), ("seed", ctypes.c_ulonglong), ("draw_counter", ctypes.c_void_p),
                ("w_smooth", ctypes.c_float), ("loss_extra", ctypes.c_void_p)]


class TrainStep:
    def __init__(self, model: NeuralDisplacementField, nfeet: int, P2: int, with_colour: bool = True,
                 frozen_mlp: bool = False, is_train: bool = True, weight_chamf: float = CHAMF_WEIGHT,
                 weight_tex: float = TEX_WEIGHT, use_graph: bool = True, num_samples: Optional[int] = None,
                 weight_smooth: float = 0.0, seed: int = 0):
        """``num_samples``: Chamfer on that many surface samples of the deformed template (losses.py:67; None = on the
        vertices, BASELINE.json's configuration).  ``weight_smooth``: weight of MeshSmoothnessLoss (opts.py:97 default
        1000 when --smooth_loss is on, cfgs/train_3d.yaml:21; 0 = off).  ``seed``: key of the sample draw; a device
        counter advances the draw every step, also under graph replay."""
        self.lib = _lib.load()
        self.model = model
        dev = model.template_verts.device
        _lib.require_cuda(model.template_verts)
        self.dev = dev
        self.B, self.P2 = int(nfeet), int(P2)
        self.V = int(model.template_verts.shape[1])
        self.with_colour, self.frozen = bool(with_colour), bool(frozen_mlp)
        self.w_chamf, self.w_tex = float(weight_chamf), float(weight_tex)
        sfx = "" if is_train else "_val"
        self.tables = {k: getattr(model, k + sfx) for k in TABLE_KEYS}
        f32 = dict(dtype=torch.float32, device=dev)
        B, V = self.B, self.V
        # per-step inputs (static addresses).  One index row per table: tables with labels are indexed by batch[vec.key],
        # the others by batch['idx'] (sample_latent_vectors, trainer.py:29-46); load_inputs() fills all four from `idx`
        # unless told otherwise.
        # Two input sets can exist (the second is created by prefetch_from_host): the pipelined host feed copies batch k+1
        # straight into the set that step k is NOT reading, and every set has its own captured graph, so no device-to-device
        # copy sits between the H2D and the step.  ``idx4`` / ``idx`` / ``target`` / ``col_target`` are the CURRENT set's.
        self._sets = [self._new_input_set()]
        self._cur = 0
        # gathered rows, outputs, gradients
        self.rows = {k: torch.zeros(B, t.data.shape[1], **f32) for k, t in self.tables.items() if t is not None}
        self.grows = {k: torch.zeros_like(v) for k, v in self.rows.items()}
        # dense table gradients, written (not accumulated) by the scatter at the end of every enqueue: what
        # loss.backward() leaves in LatentVector.data.grad after zero_grad() (trainer.py:99-100,121)
        self.table_grads = {k: torch.zeros_like(t.data) for k, t in self.tables.items() if t is not None}
        self.disp, self.col, self.verts = (torch.empty(B, V, 3, **f32) for _ in range(3))
        self.mlp_params = model.mlp_parameters()
        n_flat = sum(p.numel() for p in self.mlp_params)
        self.flat_grad = torch.zeros(n_flat, **f32)          # the data-parallel allreduce bucket
        self.grad_views, off = [], 0
        for p in self.mlp_params:
            self.grad_views.append(self.flat_grad[off:off + p.numel()].view_as(p))
            off += p.numel()
        self.S = int(num_samples) if num_samples else 0
        self.w_smooth = float(weight_smooth)
        self.mesh_terms = None
        if self.S > 0 or self.w_smooth != 0.0:
            from .mesh import MeshTopology
            faces = model.template_faces.reshape(-1, 3)
            if faces.shape[0] < 2:
                raise RuntimeError("surface sampling / smoothness need the template faces (model.set_template(verts, faces))")
            self.topo = MeshTopology(faces.to(dev), V)
            self.draw_counter = torch.zeros(1, dtype=torch.int64, device=dev)
            self.loss_extra = torch.zeros(2, **f32)
            t, mt = self.topo, MeshTermsC()
            mt.faces, mt.F, mt.edges, mt.E = t.faces.data_ptr(), t.F, t.edges.data_ptr(), t.E
            mt.vf_ptr, mt.vf_face, mt.vf_corner, mt.nnz = t.vf_ptr.data_ptr(), t.vf_face.data_ptr(), t.vf_corner.data_ptr(), t.nnz
            mt.num_samples, mt.seed, mt.draw_counter = self.S, int(seed) & (2 ** 64 - 1), self.draw_counter.data_ptr()
            mt.w_smooth, mt.loss_extra = self.w_smooth, self.loss_extra.data_ptr()
            self.mesh_terms = mt
            nbytes = int(self.lib.findb200_train_step_mesh_workspace_bytes(B, V, self.P2, self.S,
                                                                           t.nnz if self.w_smooth != 0.0 else 0, t.F))
        else:
            nbytes = int(self.lib.findb200_train_step_workspace_bytes(B, V, self.P2))
        self.ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        self.kernel_launches = 0         # filled by the first enqueue (counted from the C ABI's launch list)
        self._graphs = [None, None]      # one captured step per input set
        self.use_graph = use_graph
        self._cparams = None
        self._cgrads = None
        self._jobs = [None, None]        # gather / scatter job tables per input set (they hold the index rows' addresses)
        self._cache_key = None           # what the cached tiles in the workspace were computed from

    # ---- the enqueue ---------------------------------------------------------------------------------------
    def _structs(self):
        m = self.model
        mods = m.mlp_modules()
        for p in self.mlp_params:
            if not p.is_contiguous() or p.dtype != torch.float32:
                raise RuntimeError("MLP parameters must be contiguous fp32")
        S = self.rows["shapevec"].shape[1] if "shapevec" in self.rows else 0
        P = self.rows["posevec"].shape[1] if "posevec" in self.rows else 0
        T = self.rows["texvec"].shape[1] if "texvec" in self.rows else 0
        self._B = m.encoder[0].B_on(self.dev)
        self._avg = m.avg_col.detach() if m.use_avg_colour else None
        cp = _make_params([x.weight.detach() for x in mods], [x.bias.detach() for x in mods], self._B, self._avg, S, P, T)
        cg = None
        if not self.frozen:
            cg = FieldGradsC()
            for i in range(N_PARAMS):
                cg.weight[i] = self.grad_views[i].data_ptr()
                cg.bias[i] = self.grad_views[N_PARAMS + i].data_ptr()
        return cp, cg

    def _new_input_set(self):
        f32 = dict(dtype=torch.float32, device=self.dev)
        return dict(idx4=torch.zeros(len(TABLE_KEYS), self.B, dtype=torch.int64, device=self.dev),
                    target=torch.zeros(self.B, self.P2, 3, **f32),
                    col_target=torch.zeros(self.B, self.V, 3, **f32) if self.with_colour else None,
                    loss3=torch.zeros(3, **f32))       # [total, chamfer, colour]: written by the step that reads this set

    @property
    def idx4(self):
        return self._sets[self._cur]["idx4"]

    @property
    def idx(self):                                 # the batch['idx'] row (kept under its old name)
        return self._sets[self._cur]["idx4"][0]

    @property
    def target(self):
        return self._sets[self._cur]["target"]

    @property
    def col_target(self):
        return self._sets[self._cur]["col_target"]

    @property
    def loss3(self):
        return self._sets[self._cur]["loss3"]

    @property
    def graph(self):
        return self._graphs[self._cur]

    def _enqueue(self):
        lib, st = self.lib, _lib.stream_ptr(self.dev)
        if self._cparams is None:
            self._cparams, self._cgrads = self._structs()
        if self._jobs[self._cur] is None:
            live = [(k, t) for k, t in self.tables.items() if t is not None]
            jobs = (GatherJobC * len(live))()
            sjobs = (ScatterJobC * len(live))()
            for j, (k, t) in enumerate(live):
                irow = self.idx4[TABLE_KEYS.index(k)].data_ptr()
                jobs[j].table, jobs[j].n_rows, jobs[j].width = t.data.data_ptr(), t.data.shape[0], t.data.shape[1]
                jobs[j].out, jobs[j].idx = self.rows[k].data_ptr(), irow
                sjobs[j].g, sjobs[j].n_rows, sjobs[j].width = self.grows[k].data_ptr(), t.data.shape[0], t.data.shape[1]
                sjobs[j].grad_table, sjobs[j].idx = self.table_grads[k].data_ptr(), irow
            self._jobs[self._cur] = ((jobs, len(live)), (sjobs, len(live)))
        gather, scatter = self._jobs[self._cur]
        if gather[1]:
            rc = lib.findb200_gather_rows_multi(gather[0], gather[1], _lib.ptr(self.idx), self.B, st)
            _lib.check(rc, "findb200_gather_rows_multi")
        # Workspace caches (include/find_b200.h FINDB200_FIELD_*): the template never changes during training, so after the
        # first enqueue the Fourier-encoded tiles are reused; with a frozen MLP the whole trunk output is.  Any in-place
        # change of the template (or of the frozen weights) is seen through the tensors' version counters.
        key = self._current_cache_key()
        flags = 0
        if self._cache_key == key:
            flags = FIELD_ENC_CACHED | (FIELD_TRUNK_CACHED if self.frozen else 0)
        self._cparams.flags = flags
        self._cache_key = None                   # set again below, once the enqueue has succeeded
        r, g = self.rows, self.grows
        args = (ctypes.byref(self._cparams), _lib.ptr(self.model.template_verts), _lib.ptr(r.get("shapevec")),
                _lib.ptr(r.get("posevec")), _lib.ptr(r.get("texvec")), _lib.ptr(r.get("reg")), _lib.ptr(self.target), None,
                _lib.ptr(self.col_target), self.B, self.V, self.P2, self.w_chamf, self.w_tex if self.with_colour else 0.0,
                _lib.ptr(self.loss3), _lib.ptr(self.disp), _lib.ptr(self.col), _lib.ptr(self.verts),
                ctypes.byref(self._cgrads) if self._cgrads is not None else None, _lib.ptr(g.get("shapevec")),
                _lib.ptr(g.get("posevec")), _lib.ptr(g.get("texvec")), _lib.ptr(g.get("reg")))
        if self.mesh_terms is None:
            rc = lib.findb200_train_step(*args, _lib.ptr(self.ws), self.ws.numel(), st)
            _lib.check(rc, "findb200_train_step")
        else:
            rc = lib.findb200_train_step_mesh(*args, ctypes.byref(self.mesh_terms), _lib.ptr(self.ws), self.ws.numel(), st)
            _lib.check(rc, "findb200_train_step_mesh")
        if scatter[1]:
            # autograd of the gather (trainer.py:121): dense table gradients, part of the step (and of its CUDA graph)
            rc = lib.findb200_scatter_rows_multi(scatter[0], scatter[1], _lib.ptr(self.idx), self.B, 0, st)
            _lib.check(rc, "findb200_scatter_rows_multi")
        self._cache_key = key

    def _current_cache_key(self):
        tv = self.model.template_verts
        key = (tv.data_ptr(), tv._version)
        if self.frozen:
            key = key + tuple((p.data_ptr(), p._version) for p in self.mlp_params)
        return key

    def run(self) -> torch.Tensor:
        """Enqueue one step on the current stream (graph replay after the first call) -> loss3 (device)."""
        with torch.cuda.device(self.dev):
            if not self.use_graph:
                self._enqueue()
            elif self.graphs_stale() or self.graph is None:      # (stale first: it drops the graphs of BOTH input sets)
                self._enqueue()                      # warm-up outside capture (sets function attributes)
                torch.cuda.synchronize(self.dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._enqueue()
                self._graphs[self._cur] = g
                g.replay()
            else:
                self.graph.replay()
        return self.loss3

    def graphs_stale(self) -> bool:
        """True (once) when the template — or the frozen weights — changed since the cached tiles were computed: every captured
        graph is dropped and the next enqueue recomputes the caches."""
        if self._cache_key is not None and self._cache_key != self._current_cache_key():
            self._cache_key = None
            self._graphs = [None, None]
            return True
        return False

    # ---- host-facing conveniences --------------------------------------------------------------------------
    def _check_rows(self, idx: torch.Tensor, key: str = "idx"):
        """Host-resident indices are validated here (torch's `self.data[idx]` would raise); device-resident ones by the kernels."""
        if idx.is_cuda:
            return
        if idx.numel() != self.B:
            raise RuntimeError(f"TrainStep: {key} has {idx.numel()} entries, the step was built for {self.B} feet")
        lo, hi = int(idx.min()), int(idx.max())
        tabs = [(k, t) for k, t in self.tables.items() if t is not None and (key == "idx" or k == key)]
        for k, t in tabs:
            if lo < 0 or hi >= t.data.shape[0]:
                raise IndexError(f"TrainStep: {key} range [{lo}, {hi}] is outside table '{k}' with {t.data.shape[0]} rows")

    def load_inputs(self, idx: torch.Tensor, target: torch.Tensor, col_target: Optional[torch.Tensor] = None,
                    table_idx: Optional[Dict[str, torch.Tensor]] = None):
        """Copy one batch (host pinned or device tensors) into the static input buffers, asynchronously.  ``table_idx``:
        per-table row indices for tables that are indexed by label rather than by batch['idx'] (see ``rows_of_batch``)."""
        self._check_rows(idx)
        for r in range(len(TABLE_KEYS)):
            self.idx4[r].copy_(idx, non_blocking=True)
        for k, v in (table_idx or {}).items():
            self._check_rows(v, k)
            self.idx4[TABLE_KEYS.index(k)].copy_(v, non_blocking=True)
        self.target.copy_(target, non_blocking=True)
        if self.with_colour and col_target is not None:
            self.col_target.copy_(col_target, non_blocking=True)

    def rows_of_batch(self, batch) -> Dict[str, torch.Tensor]:
        """sample_latent_vectors (trainer.py:29-46) as row indices: a labelled table is indexed by batch[vec.key] through
        its label list, every other table by batch['idx'].  -> {table: int64 [B]} for ``load_inputs(table_idx=...)``."""
        out = {}
        for k, t in self.tables.items():
            if t is not None and t.labels is not None:
                assert t.key in batch, f"Trying to sample from latent vector {t.key} using keys, but not found in dataset"
                out[k] = torch.tensor([t._lookup(lab) for lab in batch[t.key]], dtype=torch.int64)
        return out

    def populate_grads(self):
        """Make ``.grad`` of every parameter what autograd would have produced (dense table grads, trainer.py:121).  The
        gradients were already written by the enqueue (MLP bucket by the backward kernels, tables by the in-step scatter);
        this only attaches the buffers — no launch."""
        if not self.frozen:
            for p, gv in zip(self.mlp_params, self.grad_views):
                p.grad = gv
        for k, t in self.tables.items():
            if t is not None:
                t.data.grad = self.table_grads[k]

    # ---- pipelined host feed: the NEXT batch crosses PCIe on a copy stream while the current step computes ---------------
    def prefetch_from_host(self, idx: torch.Tensor, target: torch.Tensor, col_target: Optional[torch.Tensor] = None):
        """Stage one batch (pinned host tensors) into one of two device staging sets on a side stream.  Pair every call with
        one ``step_prefetched()``; at most two batches may be staged ahead."""
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
            if len(self._sets) < 2:
                self._sets.append(self._new_input_set())
            self._ready = [torch.cuda.Event() for _ in range(2)]
            self._consumed = [torch.cuda.Event() for _ in range(2)]
            self._staged, self._stage_next = [], 0
        if len(self._staged) >= 2:
            raise RuntimeError("prefetch_from_host: two batches are already staged; call step_prefetched() first")
        self._check_rows(idx)
        k = self._stage_next
        st = self._sets[k]
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._consumed[k])          # the step that last read this set has finished
            for r in range(len(TABLE_KEYS)):
                st["idx4"][r].copy_(idx, non_blocking=True)
            st["target"].copy_(target, non_blocking=True)
            if st["col_target"] is not None and col_target is not None:
                st["col_target"].copy_(col_target, non_blocking=True)
            self._ready[k].record(self._copy_stream)
        self._staged.append(k)
        self._stage_next = k ^ 1

    def _launch_prefetched(self, runner=None) -> int:
        """Enqueue one step on the oldest staged batch plus an asynchronous D2H copy of its losses; -> the loss slot to collect."""
        if not getattr(self, "_staged", None):
            raise RuntimeError("step_prefetched: nothing staged; call prefetch_from_host() first")
        if getattr(self, "_loss_host", None) is None:
            self._loss_host = [torch.empty(3, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._loss_done = [torch.cuda.Event() for _ in range(2)]
            self._d2h_stream = torch.cuda.Stream(device=self.dev)
        k = self._staged.pop(0)
        cur = torch.cuda.current_stream(self.dev)
        cur.wait_event(self._ready[k])
        cur.wait_event(self._loss_done[k])                           # set k's losses of two steps ago have left the device
        self._cur = k                                                # the step reads set k in place (its own captured graph)
        (runner or self).run()
        self._consumed[k].record(cur)
        # the losses go home on their own stream (every set has its own loss3), so the next step's graph follows this one directly
        self._d2h_stream.wait_event(self._consumed[k])
        with torch.cuda.stream(self._d2h_stream):
            self._loss_host[k].copy_(self._sets[k]["loss3"], non_blocking=True)
            self._loss_done[k].record(self._d2h_stream)
        return k

    def _collect(self, j: int) -> Dict[str, float]:
        self._loss_done[j].synchronize()
        total, chamf, tex = self._loss_host[j].tolist()
        return {"loss": total, "chamf": chamf, "tex": tex}

    def step_prefetched(self, runner=None) -> Dict[str, float]:
        """Run one step on the oldest staged batch (``runner``: a ShardedTrainStep wrapping this step, default this step)
        and read the losses back (one D2H read, like trainer.py:128-130)."""
        return self._collect(self._launch_prefetched(runner))

    def iterate_from_host(self, batches, runner=None, deferred: bool = True):
        """for losses in step.iterate_from_host(loader): ... — ``batches`` yields (idx, target, col_target) pinned host tensors.
        Batch k+1 is staged on the copy stream before step k is enqueued, and (``deferred``) the losses of step k are read
        from pinned memory only after step k+1 has been enqueued, so neither PCIe nor the host's launch work is exposed.
        Yields one loss dict per batch, in order; gradients / outputs on the device are those of the LAST enqueued step."""
        it = iter(batches)
        try:
            self.prefetch_from_host(*next(it))
        except StopIteration:
            return
        pending = None
        for nxt in it:
            self.prefetch_from_host(*nxt)
            tok = self._launch_prefetched(runner)
            if pending is not None:
                yield self._collect(pending)
                pending = None
            if deferred:
                pending = tok
            else:
                yield self._collect(tok)
        tok = self._launch_prefetched(runner)
        if pending is not None:
            yield self._collect(pending)
        yield self._collect(tok)

    def step_from_host(self, idx: torch.Tensor, target: torch.Tensor, col_target: Optional[torch.Tensor] = None
                       ) -> Dict[str, float]:
        """The public end-to-end call: host batch in, losses out (one D2H read, like trainer.py:128-130)."""
        self.load_inputs(idx, target, col_target)
        self.run()
        total, chamf, tex = self.loss3.tolist()
        return {"loss": total, "chamf": chamf, "tex": tex}


__all__ = ["TrainStep"]
