"""tools/l2_probe.py — L2 -> SMEM bulk-copy bandwidth when every CTA streams the same region (weight streaming)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib  # noqa: E402

lib = _lib.load()
dev = torch.device("cuda:0")
sms = torch.cuda.get_device_properties(0).multi_processor_count
scratch = torch.zeros(16, device=dev)
out = {}
src = torch.zeros(1792 * 1024, dtype=torch.uint8, device=dev)
for ctas_per_sm, chunk_kb, stages, issuers in ((1, 16, 6, 1), (1, 32, 3, 1), (1, 8, 8, 1), (1, 4, 8, 1), (1, 16, 3, 2),
                                               (1, 16, 3, 4), (1, 8, 4, 4), (2, 16, 3, 2), (1, 32, 3, 2), (1, 2, 8, 4)):
    blocks = sms * ctas_per_sm
    iters = 65536 // chunk_kb
    def run():
        _lib.check(lib.findb200_l2_stream(_lib.ptr(src), src.numel(), chunk_kb * 1024, blocks, iters, stages, issuers,
                                          _lib.ptr(scratch), _lib.stream_ptr(dev)), "l2_stream")
    run(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); run(); e1.record(); torch.cuda.synchronize()
    t = e0.elapsed_time(e1) * 1e-3
    tot = blocks * issuers * iters * chunk_kb * 1024
    out[f"cta{ctas_per_sm}_chunk{chunk_kb}KB_st{stages}_iss{issuers}"] = {
        "TBps": round(tot / t / 1e12, 2), "B_per_ns_per_SM": round(tot / t / 1e9 / sms, 1)}
print(json.dumps(out, indent=1))
