"""tools/rank_spread.py — how much of the N = 8 "scaling loss" is workload variance between ranks?  On ONE GPU, build the
batch every rank of an 8-rank job would own (bench.build_job(rank, world=8)) and time its step alone (graph replay, CUDA
events, L2 flushed between steps).  A synchronous data-parallel step runs at the pace of its slowest rank."""
import json
import os
import statistics
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench                                                 # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    wl = bench.WORKLOADS[os.environ.get("WORKLOAD", "cfg2")]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = []
    for rank in range(8):
        model, step, host = bench.build_job(wl, rank, 8, dev)
        step.load_inputs(host["idx"], host["target"], host["col_target"])
        for _ in range(5):
            step.run()
        ts = []
        for _ in range(40):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); step.run(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        out.append(round(statistics.median(ts), 4))
        del step, model
        torch.cuda.empty_cache()
    print(json.dumps(dict(ms_per_step_by_rank_batch=out, slowest=max(out), fastest=min(out), mean=round(sum(out) / 8, 4))))


if __name__ == "__main__":
    main()
