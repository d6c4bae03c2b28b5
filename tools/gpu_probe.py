"""tools/gpu_probe.py — quick on-box numbers: FP32 peak (FFMA / FFMA2) and Chamfer kernel throughput."""
import json
import sys
import os
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from find_b200 import _lib, chamfer_distance  # noqa: E402


def timeit(fn, warm=3, iters=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    lib = _lib.load()
    dev = torch.device("cuda:0")
    props = torch.cuda.get_device_properties(0)
    out = {"gpu": props.name, "sms": props.multi_processor_count}
    scratch = torch.zeros(16, device=dev)
    blocks, iters = props.multi_processor_count * 8, 4096
    for mode, name in ((0, "ffma"), (1, "ffma2")):
        t = timeit(lambda: lib.findb200_fp32_peak(mode, blocks, iters, _lib.ptr(scratch), _lib.stream_ptr(dev)))
        lanes = 2 if mode else 1
        out[f"{name}_tflops"] = blocks * 256 * iters * 64 * lanes * 2 / t / 1e12
    for (n, p1, p2) in ((8, 10242, 10000), (8, 10242, 20000), (4, 50000, 100000)):
        g = torch.Generator(device="cpu").manual_seed(0)
        x = (torch.randn(n, p1, 3, generator=g) * 0.05).to(dev)
        y = (torch.randn(n, p2, 3, generator=g) * 0.05).to(dev)
        t = timeit(lambda: chamfer_distance(x, y), iters=5)
        out[f"chamfer_fwd_{n}x{p1}x{p2}"] = {"ms": t * 1e3, "gpairs_per_s": 2 * n * p1 * p2 / t / 1e9}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
