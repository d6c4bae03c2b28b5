"""tools/cta2_probe.py — does a plain cp.async.bulk accept an mbarrier of the PEER CTA (self-test mode 3)?  Run in its own
process: a wrong guess traps the context."""
import json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_umma_gpu import _run_pair  # noqa: E402

rng = np.random.RandomState(3)
A = (rng.randn(256, 256) * 0.5).astype(np.float16)
W = (rng.randn(256, 256) * 0.1).astype(np.float16)
ref = A.astype(np.float32) @ W.astype(np.float32).T
out = {}
for mode in (2, 3):
    try:
        D = _run_pair(mode, A, W)
        out[f"mode{mode}"] = {"max_abs_err": float(np.abs(D - ref).max()), "ok": bool(np.allclose(D, ref, rtol=1e-3, atol=1e-3))}
    except Exception as e:  # noqa: BLE001
        out[f"mode{mode}"] = {"error": str(e)[:300]}
        break
print(json.dumps(out))
