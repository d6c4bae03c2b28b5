"""tools/sync_probe.py — cycles per synchronisation primitive of the chain kernels' issue path (findb200_sync_probe)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib
lib = _lib.load()
dev = torch.device("cuda:0")
names = ["commit cg1 (idle pipe)", "commit cg2 multicast (idle pipe)", "mbarrier.try_wait (complete)", "ld.volatile.shared", "mbarrier.arrive local",
         "mbarrier.arrive peer CTA", "fence.proxy.async", "tcgen05.fence::after_thread_sync", "4 MMA cg2 + commit_mc + wait (round trip)",
         "4 MMA cg2 + commit_mc (pipelined)", "read %cluster_ctaid", "clock64"]
buf = torch.zeros(2, dtype=torch.int64, device=dev)
out = {}
for op, name in enumerate(names):
    _lib.check(lib.findb200_sync_probe(op, 256, _lib.ptr(buf), _lib.stream_ptr(dev)), "sync_probe")
    torch.cuda.synchronize()
    cyc, n = buf.tolist()
    out[name] = round(cyc / n, 1)
print(json.dumps(out, indent=1))
