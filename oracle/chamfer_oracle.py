"""oracle/chamfer_oracle.py — TEST INFRASTRUCTURE ONLY.

ctypes front-end of ``oracle/chamfer_oracle.c`` (CPU restatement of PyTorch3D ``knn_points(K=1)`` /
``chamfer_distance`` as reached from the reference at ``src/model/losses.py:77,85,88`` and
``src/eval/eval_3d.py:151,159,176-177``).  PARITY UNPINNED (PyTorch3D is absent; see the C header).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` legs
may import this module.  Nothing under ``find_b200/`` does.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libchamfer_oracle.so")
_lib = None

MODE_FMA_CHAIN = 0   # pytorch3d CUDA build (knn.cu under nvcc -fmad=true): the parity contract
MODE_SEPARATE = 1    # pytorch3d CPU build (knn_cpu.cpp on x86, no contraction)


def build(force: bool = False) -> str:
    """Compile the C checker in place (gcc only; seconds)."""
    src = os.path.join(_HERE, "chamfer_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_SO)
        f32p, i64p, f64p = (ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int64),
                            ctypes.POINTER(ctypes.c_double))
        lib.oracle_knn1.argtypes = [f32p, f32p, i64p, i64p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64,
                                    ctypes.c_int, ctypes.c_int, f32p, i64p]
        lib.oracle_knn1_backward.argtypes = [f32p, f32p, i64p, i64p, i64p, f32p, ctypes.c_int64,
                                             ctypes.c_int64, ctypes.c_int64, f32p, f32p]
        lib.oracle_chamfer.argtypes = [f32p, f32p, i64p, i64p, ctypes.c_int64, ctypes.c_int64,
                                       ctypes.c_int64, ctypes.c_int, ctypes.c_int, f64p, f32p, i64p, f32p,
                                       i64p, f32p, f32p]
        lib.oracle_max_threads.restype = ctypes.c_int
        _lib = lib
    return _lib


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _ptr(a, ty):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(ty))


def _lens(l, N):
    if l is None:
        return None
    l = np.ascontiguousarray(np.asarray(l, dtype=np.int64))
    assert l.shape == (N,)
    return l


def max_threads() -> int:
    return int(_load().oracle_max_threads())


def knn1(p1, p2, lengths1=None, lengths2=None, mode: int = MODE_FMA_CHAIN, threads: int = 1):
    """-> (dists [N,P1] float32 squared-L2, idx [N,P1] int64)."""
    lib = _load()
    p1, p2 = _f32(p1), _f32(p2)
    N, P1, _ = p1.shape
    P2 = p2.shape[1]
    assert p1.shape[2] == 3 and p2.shape[2] == 3 and p2.shape[0] == N
    l1, l2 = _lens(lengths1, N), _lens(lengths2, N)
    d = np.empty((N, P1), np.float32)
    i = np.empty((N, P1), np.int64)
    rc = lib.oracle_knn1(_ptr(p1, ctypes.c_float), _ptr(p2, ctypes.c_float), _ptr(l1, ctypes.c_int64),
                         _ptr(l2, ctypes.c_int64), N, P1, P2, mode, threads,
                         _ptr(d, ctypes.c_float), _ptr(i, ctypes.c_int64))
    if rc:
        raise ValueError(f"oracle_knn1 failed rc={rc}")
    return d, i


def knn1_backward(p1, p2, idx, grad_dists, lengths1=None, lengths2=None):
    lib = _load()
    p1, p2, g = _f32(p1), _f32(p2), _f32(grad_dists)
    idx = np.ascontiguousarray(idx, dtype=np.int64)
    N, P1, _ = p1.shape
    P2 = p2.shape[1]
    l1, l2 = _lens(lengths1, N), _lens(lengths2, N)
    g1 = np.empty_like(p1)
    g2 = np.empty_like(p2)
    lib.oracle_knn1_backward(_ptr(p1, ctypes.c_float), _ptr(p2, ctypes.c_float), _ptr(l1, ctypes.c_int64),
                             _ptr(l2, ctypes.c_int64), _ptr(idx, ctypes.c_int64), _ptr(g, ctypes.c_float),
                             N, P1, P2, _ptr(g1, ctypes.c_float), _ptr(g2, ctypes.c_float))
    return g1, g2


def chamfer(x, y, x_lengths=None, y_lengths=None, mode: int = MODE_FMA_CHAIN, threads: int = 1,
            with_grad: bool = True):
    """chamfer_distance(x, y) with mean/mean reductions.

    -> dict(loss float64, dist_x, idx_x, dist_y, idx_y, grad_x, grad_y)  (grads for upstream grad 1)."""
    lib = _load()
    x, y = _f32(x), _f32(y)
    N, P1, _ = x.shape
    P2 = y.shape[1]
    lx, ly = _lens(x_lengths, N), _lens(y_lengths, N)
    loss = ctypes.c_double(0.0)
    dx, ix = np.empty((N, P1), np.float32), np.empty((N, P1), np.int64)
    dy, iy = np.empty((N, P2), np.float32), np.empty((N, P2), np.int64)
    gx = np.empty_like(x) if with_grad else None
    gy = np.empty_like(y) if with_grad else None
    rc = lib.oracle_chamfer(_ptr(x, ctypes.c_float), _ptr(y, ctypes.c_float), _ptr(lx, ctypes.c_int64),
                            _ptr(ly, ctypes.c_int64), N, P1, P2, mode, threads, ctypes.byref(loss),
                            _ptr(dx, ctypes.c_float), _ptr(ix, ctypes.c_int64), _ptr(dy, ctypes.c_float),
                            _ptr(iy, ctypes.c_int64), _ptr(gx, ctypes.c_float), _ptr(gy, ctypes.c_float))
    if rc:
        raise ValueError(f"oracle_chamfer failed rc={rc}")
    return dict(loss=loss.value, dist_x=dx, idx_x=ix, dist_y=dy, idx_y=iy, grad_x=gx, grad_y=gy)
