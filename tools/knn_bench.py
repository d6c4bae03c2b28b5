"""tools/knn_bench.py — time findb200_chamfer_forward (both directions) on synthetic foot clouds in every kNN mode and
report executed vs algorithmic pair evaluations.  GPU only.  Usage: python tools/knn_bench.py [B V P2]..."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib, synthetic as syn  # noqa: E402


def run(B, V, P2, mode, kq, iters=int(os.environ.get('KB_ITERS', '20'))):
    lib = _lib.load()
    dev = torch.device("cuda:0")
    lib.findb200_set_knn_mode(mode, kq)
    torch.manual_seed(0)
    tv, _ = syn.make_template(V)
    V = tv.shape[0]
    x = (tv[None] * (1 + 0.02 * torch.randn(B, 1, 3))).contiguous().to(dev)
    y = syn.make_targets(B, P2, 0).to(dev)
    nb = lib.findb200_knn_workspace_bytes(B, V, P2)
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    loss = torch.empty((), device=dev)
    ix = torch.empty((B, V), dtype=torch.int64, device=dev)
    iy = torch.empty((B, P2), dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = _lib.stream_ptr(dev)

    def call():
        rc = lib.findb200_chamfer_forward(_lib.ptr(x), _lib.ptr(y), None, None, B, V, P2, _lib.ptr(loss), None, _lib.ptr(ix),
                                          None, _lib.ptr(iy), _lib.ptr(ws), nb, st)
        _lib.check(rc, "chamfer_forward")
    for _ in range(3):
        call()
    ts = []
    for _ in range(iters):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); call(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    stats = (ctypes.c_ulonglong * 4)()
    lib.findb200_knn_stats(_lib.ptr(ws), stats, st)
    lib.findb200_profile_enable(1)
    call()
    buf = ctypes.create_string_buffer(1 << 14)
    lib.findb200_profile_read(buf, len(buf))
    lib.findb200_profile_enable(0)
    prof = {l.split(",")[0]: float(l.split(",")[2]) for l in buf.value.decode().strip().splitlines() if l}
    alg = 2 * B * V * P2
    return {"B": B, "V": V, "P2": P2, "mode": mode, "kq": kq, "ms_median": ts[len(ts) // 2], "ms_min": ts[0],
            "pairs_alg": alg, "pairs_exec": int(stats[0]) if mode != 1 else alg, "slow_queries": int(stats[1]), "max_evals": int(stats[2]), "groups": int(stats[3]),
            "kernels_ms": prof, "loss": loss.item(), "idx_sum": int(ix.sum().item() + iy.sum().item())}


if __name__ == "__main__":
    shapes = [(8, 10242, 10000), (32, 10242, 20000), (4, 50000, 100000)]
    if len(sys.argv) > 3:
        a = [int(v) for v in sys.argv[1:]]
        shapes = [tuple(a[i:i + 3]) for i in range(0, len(a), 3)]
    for s in shapes:
        modes = [tuple(int(v) for v in m.split(':')) for m in os.environ.get('KB_MODES', '1:0,0:1,0:2,0:4').split(',')]
        for mode, kq in modes:
            print(json.dumps(run(*s, mode, kq)), flush=True)
