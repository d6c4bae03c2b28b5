"""find_b200/field_ops.py — torch.autograd bridge to the displacement-field kernels (C ABI in include/find_b200.h).

``field_apply`` is the differentiable op the ``NeuralDisplacementField`` drop-in (find_b200/field.py) is built on:
it replaces NeuralDisplacementField.forward + the registration transform of get_meshes
(/root/reference/src/model/model.py:393-453, 481-491) and their autograd.  All arithmetic happens in
libfindb200.so; torch only owns the memory.  There is no CPU path: CPU tensors raise.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib

PARAM_NAMES = ["base.0", "base.2", "base.4", "base.6", "base.8",
               "mlp_disp.0", "mlp_disp.2", "mlp_disp.4", "mlp_disp.6",
               "mlp_col.0", "mlp_col.2", "mlp_col.4", "mlp_col.6"]
N_PARAMS = len(PARAM_NAMES)
WIDTH = 256
ENC = 515


class FieldParamsC(ctypes.Structure):
    _fields_ = [("weight", ctypes.c_void_p * N_PARAMS), ("bias", ctypes.c_void_p * N_PARAMS),
                ("fourier_B", ctypes.c_void_p), ("avg_col", ctypes.c_void_p),
                ("shapevec_size", ctypes.c_int), ("posevec_size", ctypes.c_int), ("texvec_size", ctypes.c_int),
                ("flags", ctypes.c_int)]


FIELD_ENC_CACHED, FIELD_TRUNK_CACHED = 1, 2


class FieldGradsC(ctypes.Structure):
    _fields_ = [("weight", ctypes.c_void_p * N_PARAMS), ("bias", ctypes.c_void_p * N_PARAMS)]


def _f32c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if t is None:
        return None
    t = t.detach()
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _expected_shapes(S: int, P: int, T: int):
    ins = [ENC, WIDTH, WIDTH, WIDTH, WIDTH, WIDTH + S + P, WIDTH, WIDTH, WIDTH, WIDTH + T, WIDTH, WIDTH, WIDTH]
    outs = [WIDTH] * 8 + [3] + [WIDTH] * 3 + [3]
    return ins, outs


def _make_params(weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor], fourier_B: torch.Tensor,
                 avg_col: Optional[torch.Tensor], S: int, P: int, T: int) -> FieldParamsC:
    ins, outs = _expected_shapes(S, P, T)
    c = FieldParamsC()
    for i in range(N_PARAMS):
        if tuple(weights[i].shape) != (outs[i], ins[i]) or tuple(biases[i].shape) != (outs[i],):
            raise RuntimeError(f"{PARAM_NAMES[i]}: expected weight {(outs[i], ins[i])} / bias {(outs[i],)}, got "
                               f"{tuple(weights[i].shape)} / {tuple(biases[i].shape)}")
        c.weight[i] = weights[i].data_ptr()
        c.bias[i] = biases[i].data_ptr()
    if tuple(fourier_B.shape) != (3, WIDTH):
        raise RuntimeError(f"fourier_B must be [3, {WIDTH}], got {tuple(fourier_B.shape)}")
    c.fourier_B = fourier_B.data_ptr()
    c.avg_col = None if avg_col is None else avg_col.data_ptr()
    c.shapevec_size, c.posevec_size, c.texvec_size = S, P, T
    c.flags = 0
    return c


def workspace_bytes(nfeet: int, V: int, pos_shared: bool, for_backward: bool) -> int:
    return int(_lib.load().findb200_field_workspace_bytes(nfeet, V, int(pos_shared), int(for_backward)))


class _FieldFunction(torch.autograd.Function):
    """(pos, z_shape, z_pose, z_tex, reg, fourier_B, avg_col, flags, *26 params) -> (disp, col, verts)."""

    @staticmethod
    def forward(ctx, pos, z_shape, z_pose, z_tex, reg, fourier_B, avg_col, flags, *params):
        want_col, want_verts, frozen = flags
        lib = _lib.load()
        dev = _lib.require_cuda(pos, z_shape, z_pose, z_tex, reg, fourier_B, avg_col, *params)
        if pos.dim() != 3 or pos.shape[-1] != 3:
            raise RuntimeError(f"pos must be [B or 1, V, 3], got {tuple(pos.shape)}")
        lat = [z for z in (z_shape, z_pose, z_tex) if z is not None]
        nfeet = lat[0].shape[0] if lat else pos.shape[0]
        for z in lat:
            if z.dim() != 2 or z.shape[0] != nfeet:
                raise RuntimeError("latent codes must be [B, size] with the same B")
        if pos.shape[0] not in (1, nfeet):
            raise RuntimeError(f"pos batch {pos.shape[0]} does not match latent batch {nfeet}")
        shared = pos.shape[0] == 1
        V = pos.shape[1]
        if reg is not None and tuple(reg.shape) != (nfeet, 9):
            raise RuntimeError(f"reg must be [{nfeet}, 9], got {tuple(reg.shape)}")
        S = 0 if z_shape is None else z_shape.shape[1]
        P = 0 if z_pose is None else z_pose.shape[1]
        T = 0 if z_tex is None else z_tex.shape[1]

        posc, zs, zp, zt, regc = (_f32c(t) for t in (pos, z_shape, z_pose, z_tex, reg))
        Bc, avg = _f32c(fourier_B), _f32c(avg_col)
        pc = [_f32c(t) for t in params]
        weights, biases = pc[:N_PARAMS], pc[N_PARAMS:]
        cparams = _make_params(weights, biases, Bc, avg, S, P, T)

        needs_grad = any(ctx.needs_input_grad)
        disp = torch.empty((nfeet, V, 3), dtype=torch.float32, device=dev)
        col = torch.empty((nfeet, V, 3), dtype=torch.float32, device=dev) if want_col else None
        verts = torch.empty((nfeet, V, 3), dtype=torch.float32, device=dev) if want_verts else None
        ws = torch.empty(workspace_bytes(nfeet, V, shared, needs_grad), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.findb200_field_forward(ctypes.byref(cparams), _lib.ptr(posc), int(shared), _lib.ptr(zs), _lib.ptr(zp),
                                            _lib.ptr(zt), _lib.ptr(regc), nfeet, V, int(needs_grad), _lib.ptr(disp),
                                            _lib.ptr(col), _lib.ptr(verts), _lib.ptr(ws), ws.numel(),
                                            _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_field_forward")
        if needs_grad:
            ctx.save_for_backward(posc, zs, zp, zt, regc, Bc, avg, disp, col, ws, *pc)
            ctx.meta = (nfeet, V, shared, S, P, T, frozen)
        outs = (disp, col if want_col else disp.new_empty(0), verts if want_verts else disp.new_empty(0))
        if not want_col:
            ctx.mark_non_differentiable(outs[1])
        if not want_verts:
            ctx.mark_non_differentiable(outs[2])
        return outs

    @staticmethod
    def backward(ctx, g_disp, g_col, g_verts):
        lib = _lib.load()
        posc, zs, zp, zt, regc, Bc, avg, disp, col, ws, *pc = ctx.saved_tensors
        nfeet, V, shared, S, P, T, frozen = ctx.meta
        dev = disp.device
        weights, biases = pc[:N_PARAMS], pc[N_PARAMS:]
        cparams = _make_params(weights, biases, Bc, avg, S, P, T)
        need = ctx.needs_input_grad
        need_params = any(need[8:]) and not frozen

        def g(t, ref):
            if t is None or ref is None or t.numel() == 0:
                return None
            return _f32c(t)

        gd, gc, gv = g(g_disp, disp), g(g_col, col), g(g_verts, disp)
        grads_c, gws, gbs = None, [None] * N_PARAMS, [None] * N_PARAMS
        if need_params:
            grads_c = FieldGradsC()
            for i in range(N_PARAMS):
                gws[i] = torch.empty_like(weights[i])
                gbs[i] = torch.empty_like(biases[i])
                grads_c.weight[i] = gws[i].data_ptr()
                grads_c.bias[i] = gbs[i].data_ptr()
        gzs = torch.empty_like(zs) if zs is not None else None
        gzp = torch.empty_like(zp) if zp is not None else None
        gzt = torch.empty_like(zt) if zt is not None else None
        greg = torch.empty_like(regc) if regc is not None else None
        with torch.cuda.device(dev):
            rc = lib.findb200_field_backward(ctypes.byref(cparams), _lib.ptr(posc), int(shared), _lib.ptr(zs), _lib.ptr(zp),
                                             _lib.ptr(zt), _lib.ptr(regc), nfeet, V, _lib.ptr(disp), _lib.ptr(col),
                                             _lib.ptr(gd), _lib.ptr(gc), _lib.ptr(gv),
                                             ctypes.byref(grads_c) if grads_c is not None else None, _lib.ptr(gzs),
                                             _lib.ptr(gzp), _lib.ptr(gzt), _lib.ptr(greg), _lib.ptr(ws), ws.numel(),
                                             _lib.stream_ptr(dev))
        _lib.check(rc, "findb200_field_backward")
        out = [None, gzs if need[1] else None, gzp if need[2] else None, gzt if need[3] else None,
               greg if need[4] else None, None, None, None]
        for i in range(N_PARAMS):
            out.append(gws[i] if need[8 + i] else None)
        for i in range(N_PARAMS):
            out.append(gbs[i] if need[8 + N_PARAMS + i] else None)
        return tuple(out)


def field_apply(pos: torch.Tensor, z_shape: Optional[torch.Tensor], z_pose: Optional[torch.Tensor],
                z_tex: Optional[torch.Tensor], reg: Optional[torch.Tensor], fourier_B: torch.Tensor,
                avg_col: Optional[torch.Tensor], weights: Sequence[torch.Tensor], biases: Sequence[torch.Tensor],
                want_col: bool = True, want_verts: bool = True, frozen_mlp: bool = False
                ) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor]]:
    """disp, col, verts = field(pos; latents, reg).  pos [1 or B, V, 3]; latents [B, size] or None; reg [B, 9] or None.

    ``frozen_mlp`` skips every weight gradient (latent-code fitting, /root/reference/src/train/train.py:218-224)."""
    if len(weights) != N_PARAMS or len(biases) != N_PARAMS:
        raise RuntimeError(f"need {N_PARAMS} weights and biases in the order {PARAM_NAMES}")
    disp, col, verts = _FieldFunction.apply(pos, z_shape, z_pose, z_tex, reg, fourier_B, avg_col,
                                            (bool(want_col), bool(want_verts), bool(frozen_mlp)), *weights, *biases)
    return disp, (col if want_col else None), (verts if want_verts else None)
