"""find_b200/optim.py — torch.optim.Adam / torch.optim.SGD(momentum) as FIND steps them, one sm_100a launch per optimiser.

Replaces the four optimisers of /root/reference/src/train/train.py:161-168 (``Adam(main_params, lr_net)``,
``SGD(reg_params, lr_reg, momentum=0.9)``, ``Adam``/``SGD(latent_params)``) as stepped at
/root/reference/src/train/trainer.py:99-100,122-123 (``zero_grad`` / ``step``).  Same constructor arguments, same
``state_dict`` layout (``step``, ``exp_avg``, ``exp_avg_sq`` / ``momentum_buffer``) so checkpoints interchange with the
torch classes; updates are dense like torch's on dense ``.grad`` (table rows outside the batch keep moving by momentum).
All arithmetic happens in libfindb200.so (csrc/optim.cu); CPU parameters raise.
"""
from __future__ import annotations

import ctypes
from typing import List

import torch

from . import _lib


class OptimTensorC(ctypes.Structure):           # findb200_optim_tensor (include/find_b200.h)
    _fields_ = [("param", ctypes.c_void_p), ("grad", ctypes.c_void_p), ("m", ctypes.c_void_p), ("v", ctypes.c_void_p),
                ("numel", ctypes.c_longlong)]


MAX_TENSORS = 48


class _FusedBase(torch.optim.Optimizer):
    _m_key, _v_key, _has_step = "exp_avg", "exp_avg_sq", True

    def _prepare(self, gi, group):
        """-> (params with grads, device, device step counter) of group gi; private counters stay out of param_groups so
        state_dict() keeps torch's layout."""
        ps: List[torch.Tensor] = [p for p in group["params"] if p.grad is not None]
        if not ps:
            return None
        dev = _lib.require_cuda(*ps)
        for p in ps:
            if p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("fused optimisers need contiguous fp32 parameters")
            if p.grad.is_sparse:
                raise RuntimeError("sparse gradients are not supported (the reference uses dense .grad)")
        priv = self.__dict__.setdefault("_fb_priv", {})
        if gi not in priv or priv[gi]["step_dev"].device != dev:
            # resume from a loaded state_dict: torch's Adam keeps `step` per parameter, SGD only the buffer
            steps = [int(self.state[p]["step"]) for p in ps if "step" in self.state.get(p, {})]
            if not steps and any(self._m_key in self.state.get(p, {}) for p in ps):
                steps = [1]
            n0 = max(steps) if steps else 0
            priv[gi] = dict(step_dev=torch.tensor([n0], dtype=torch.int64, device=dev), host_step=n0)
        return ps, dev, priv[gi]

    def _table(self, ps):
        arr = (OptimTensorC * len(ps))()
        keep = []
        for i, p in enumerate(ps):
            st = self.state[p]
            if self._m_key not in st:
                st[self._m_key] = torch.zeros_like(p, memory_format=torch.preserve_format)
            if self._v_key and self._v_key not in st:
                st[self._v_key] = torch.zeros_like(p, memory_format=torch.preserve_format)
            g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
            keep.append(g)
            arr[i].param, arr[i].grad, arr[i].m = p.data_ptr(), g.data_ptr(), st[self._m_key].data_ptr()
            arr[i].v = st[self._v_key].data_ptr() if self._v_key else None
            arr[i].numel = p.numel()
        return arr, keep

    def _launch(self, lib, arr, n, group, step_dev, dev):
        raise NotImplementedError

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        lib = _lib.load()
        for gi, group in enumerate(self.param_groups):
            got = self._prepare(gi, group)
            if got is None:
                continue
            ps, dev, priv = got
            for c0 in range(0, len(ps), MAX_TENSORS):
                chunk = ps[c0:c0 + MAX_TENSORS]
                arr, keep = self._table(chunk)
                # every chunk of one step must see the same count: later chunks read a rewound copy
                step_dev = priv["step_dev"] if c0 == 0 else priv["step_dev"] - 1
                with torch.cuda.device(dev):
                    self._launch(lib, arr, len(chunk), group, step_dev, dev)
            priv["host_step"] += 1
            if self._has_step:                      # torch's Adam keeps `step` per parameter (host-side count, no sync)
                for p in ps:
                    self.state[p]["step"] = torch.tensor(float(priv["host_step"]))
        return loss


class FusedAdam(_FusedBase):
    """torch.optim.Adam(params, lr, betas, eps, weight_decay) — amsgrad / maximize / capturable are not on FIND's path."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        if lr < 0 or eps < 0 or not (0 <= betas[0] < 1) or not (0 <= betas[1] < 1) or weight_decay < 0:
            raise ValueError("Invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))

    def _launch(self, lib, arr, n, group, step_dev, dev):
        rc = lib.findb200_adam_step(arr, n, group["lr"], group["betas"][0], group["betas"][1], group["eps"],
                                    group["weight_decay"], _lib.ptr(step_dev), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_adam_step")


class FusedSGD(_FusedBase):
    """torch.optim.SGD(params, lr, momentum, weight_decay) with momentum > 0 (dampening 0, no nesterov)."""
    _m_key, _v_key, _has_step = "momentum_buffer", None, False

    def __init__(self, params, lr=1e-3, momentum=0.9, weight_decay=0.0):
        if lr < 0 or momentum <= 0 or weight_decay < 0:
            raise ValueError("FusedSGD needs lr >= 0, momentum > 0, weight_decay >= 0")
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay))

    def _launch(self, lib, arr, n, group, step_dev, dev):
        rc = lib.findb200_sgd_momentum_step(arr, n, group["lr"], group["momentum"], group["weight_decay"],
                                            _lib.ptr(step_dev), _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_sgd_momentum_step")


__all__ = ["FusedAdam", "FusedSGD"]
