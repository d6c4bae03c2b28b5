"""find_b200/_lib.py — loads libfindb200.so (the C ABI in include/find_b200.h) through ctypes.

There is NO fallback: if the CUDA library is missing or a call fails, a RuntimeError is raised.  torch is
used only as the owner of device memory and streams; every pointer handed to the library is a raw
``data_ptr()`` and every call is enqueued on torch's current CUDA stream.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfindb200.so")

_lib = None

c_void_p, c_int, c_size_t, c_float, c_char_p = (ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_float,
                                                ctypes.c_char_p)

# name -> (restype, argtypes): mirrors include/find_b200.h exactly (checked by tests/test_abi.py)
SIGNATURES = {
    "findb200_last_error": (c_char_p, []),
    "findb200_version": (c_char_p, []),
    "findb200_knn_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "findb200_knn1": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 3 + [c_size_t, c_void_p]),
    "findb200_knn1_static": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 3),
    "findb200_knn1_backward": (c_int, [c_void_p] * 6 + [c_int] * 3 + [c_void_p] * 3),
    "findb200_chamfer_forward": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 6 + [c_size_t, c_void_p]),
    "findb200_field_workspace_bytes": (c_size_t, [c_int] * 4),
    "findb200_field_forward": (c_int, [c_void_p] * 2 + [c_int] + [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 4
                               + [c_size_t, c_void_p]),
    "findb200_field_backward": (c_int, [c_void_p] * 2 + [c_int] + [c_void_p] * 4 + [c_int] * 2 + [c_void_p] * 11
                                + [c_size_t, c_void_p]),
    "findb200_train_step_workspace_bytes": (c_size_t, [c_int] * 3),
    "findb200_train_step": (c_int, [c_void_p] * 9 + [c_int] * 3 + [c_float] * 2 + [c_void_p] * 10 + [c_size_t, c_void_p]),
    "findb200_train_step_mesh_workspace_bytes": (c_size_t, [c_int] * 6),
    "findb200_train_step_mesh": (c_int, [c_void_p] * 9 + [c_int] * 3 + [c_float] * 2 + [c_void_p] * 11 + [c_size_t, c_void_p]),
    "findb200_gather_rows": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "findb200_gather_rows_multi": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "findb200_scatter_add_rows": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "findb200_scatter_rows_multi": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p]),
    "findb200_chain_trace": (c_int, [c_void_p]),
    "findb200_profile_enable": (c_int, [c_int]),
    "findb200_profile_read": (c_int, [c_char_p, c_size_t]),
    "findb200_umma_selftest": (c_int, [c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "findb200_umma_rate": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "findb200_sync_probe": (c_int, [c_int, c_int, c_void_p, c_void_p]),
    "findb200_fp32_peak": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p]),
    "findb200_l2_stream": (c_int, [c_void_p, c_size_t] + [c_int] * 5 + [c_void_p, c_void_p]),
    "findb200_set_knn_mode": (c_int, [c_int, c_int]),
    "findb200_knn_stats": (c_int, [c_void_p, c_void_p, c_void_p]),
    "findb200_sample_workspace_bytes": (c_size_t, [c_int, c_int]),
    "findb200_sample_points_forward": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p] + [c_int] * 4 + [ctypes.c_ulonglong]
                                       + [c_void_p] * 7 + [c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "findb200_sample_points_backward": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p] + [c_int] * 4
                                        + [c_float, c_int, c_void_p, c_void_p]),
    "findb200_mesh_smooth_workspace_bytes": (c_size_t, [c_int] * 3),
    "findb200_mesh_smooth": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                     c_int, c_float, c_float, c_float, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                     c_size_t, c_void_p]),
    "findb200_silhouette_workspace_bytes": (c_size_t, [c_int] * 2),
    "findb200_silhouette_saved_bytes": (c_size_t, [c_int] * 2),
    "findb200_silhouette_forward": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_float, c_float, c_int, c_int]
                                    + [c_void_p] * 6 + [c_size_t, c_void_p]),
    "findb200_silhouette_backward": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_float, c_float, c_int, c_int]
                                     + [c_void_p] * 4 + [c_size_t, c_void_p]),
    "findb200_silhouette_status": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "findb200_rasterize_nearest": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_float, c_int, c_int] + [c_void_p] * 5
                                   + [c_size_t, c_void_p]),
    "findb200_vertex_normals": (c_int, [c_void_p, c_int, c_int] + [c_void_p] * 6),
    "findb200_phong_shade": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_void_p] * 4 + [c_int] + [c_void_p] * 4),
    "findb200_phong_shade_backward": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p] * 5 + [c_int] + [c_void_p] * 7),
    "findb200_vertex_normals_backward": (c_int, [c_void_p, c_int, c_int] + [c_void_p] * 8),
    "findb200_points_render": (c_int, [c_void_p, c_void_p] + [c_int] * 5 + [c_float, c_int, c_void_p, c_void_p]),
    "findb200_feature_blend_forward": (c_int, [c_void_p, c_int, c_int, c_void_p] + [c_int] * 5 + [c_float, c_float, c_int, c_int] + [c_float] * 3 + [c_void_p] * 6 + [c_size_t, c_void_p]),
    "findb200_feature_blend_backward": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p] + [c_int] * 5 + [c_float, c_float, c_int, c_int] + [c_float] * 3 + [c_void_p] * 9 + [c_size_t, c_void_p]),
    "findb200_adam_step": (c_int, [c_void_p, c_int] + [ctypes.c_double] * 5 + [c_void_p, c_void_p]),
    "findb200_sgd_momentum_step": (c_int, [c_void_p, c_int] + [ctypes.c_double] * 3 + [c_void_p, c_void_p]),
    "findb200_allreduce_p2p_flag_bytes": (c_size_t, []),
    "findb200_allreduce_p2p": (c_int, [c_void_p, c_void_p, c_int, c_int, ctypes.c_longlong, c_float, c_void_p, c_int, c_void_p]),
    "findb200_allreduce_mc": (c_int, [c_void_p, c_void_p, c_int, c_int, ctypes.c_longlong, c_float, c_void_p, c_int, c_void_p]),
    "findb200_train_step_knn_stats": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "findb200_set_sort_cta_max": (c_int, [c_int]),
    "findb200_set_step_overlap": (c_int, [c_int, c_float]),
    "findb200_chamfer_train": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_float] + [c_void_p] * 2 + [c_size_t, c_float] + [c_void_p] * 7 + [c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p]),
    "findb200_chamfer_backward": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 3 + [c_float] + [c_void_p] * 3),
}


def load(build_if_missing: bool = True):
    """Return the ctypes handle of libfindb200.so; raise RuntimeError if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -m find_b200.build`")
        from . import build as _build
        _build.build()
    try:
        lib = ctypes.CDLL(LIB_PATH)
    except OSError as e:  # pragma: no cover
        raise RuntimeError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise RuntimeError(f"{LIB_PATH} does not export {name}; rebuild with `python -m find_b200.build`") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().findb200_last_error().decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {last_error()}")


def ptr(t: Optional[torch.Tensor]):
    """Raw device pointer (None -> NULL)."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_ptr(device: torch.device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(*tensors: Optional[torch.Tensor]) -> torch.device:
    """The product path has no CPU implementation: refuse loudly instead of silently falling back."""
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("find_b200 runs on CUDA (sm_100a) only: got a CPU tensor; there is no CPU fallback")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} vs {t.device}")
    if dev is None:
        raise RuntimeError("no tensor given")
    return dev
