"""tools/launches_md.py — turn an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py into the per-step
share table committed under profiles/.  usage: launches_md.py launches.csv bench.json steps_captured title > out.md"""
import csv
import json
import re
import sys
from collections import OrderedDict


def short(name: str) -> str:
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    return name


def main():
    path, bench, nsteps, title = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
    rows = []
    with open(path) as fh:
        lines = [l for l in fh if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((short(r["Kernel Name"]), float(r["Metric Value"]) / 1e3))
    ours = OrderedDict()
    other = OrderedDict()
    for n, us in rows:
        d = other if (n.startswith(("at::", "native::", "at_cuda_detail::")) or "probe" in n or "peak" in n) else ours
        d.setdefault(n, []).append(us)
    b = json.loads(open(bench).read().strip().splitlines()[-1])
    step_ms = b["ms_per_step"]
    total = sum(sum(v) for v in ours.values()) / nsteps
    print(f"# {title}\n")
    print(f"{len(rows)} launches captured over {nsteps} steps; per-launch times are cold-cache and serialised, so compare SHARES.\n")
    print(f"Sum of find_b200 kernels per step: {total:.1f} us (bench.py CUDA-event step time, graph replay: {1e3 * step_ms:.1f} us).\n")
    print("| kernel | launches/step | us/launch | us/step | share of step |")
    print("|---|---|---|---|---|")
    for n, v in sorted(ours.items(), key=lambda kv: -sum(kv[1])):
        per_step = sum(v) / nsteps
        print(f"| `{n}` | {len(v) / nsteps:.2f} | {sum(v) / len(v):.1f} | {per_step:.1f} | {100 * per_step / total:.1f}% |")
    print("\nbench.py live shares (launch profiler, CUDA events on the launching stream, same command without ncu):\n")
    print("| bench name | ms/step | share |")
    print("|---|---|---|")
    for k in sorted(b["kernels"], key=lambda k: -k["ms_per_step"]):
        print(f"| {k['name']} | {k['ms_per_step']:.4f} | {100 * k['share']:.1f}% |")
    if other:
        print("\nNot ours (outside the timed step): " + ", ".join(f"`{n}` x{len(v)}" for n, v in other.items()))


if __name__ == "__main__":
    main()
