#!/usr/bin/env python
"""bench.py — FIND field + Chamfer train-step throughput on B200 (BASELINE.json metric), one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W

A "step" = latent-row gather -> field forward on the shared template -> registration transform -> bidirectional Chamfer
against the target clouds (+ colour MSE) -> backward of all of it incl. dW / latent / reg gradients -> (N > 1) one
allreduce of the flat MLP-gradient bucket.  Optimiser excluded.  Weak scaling: every rank owns
`feet_per_gpu` feet of the global batch.  Prints ONE JSON line on rank 0 (contract in the task statement):
  value      foot-vertices/s, inputs resident in HBM, CUDA events per step, L2 flushed between steps, max over ranks
  e2e        same metric through the public training-loop call (TrainStep.iterate_from_host): every step copies its own
             pinned-host inputs H2D and its losses D2H inside the timed region; the copy of batch k+1 overlaps step k and
             the losses of step k are read after step k+1 is enqueued (e2e.isolated_ms_per_step = the serial form)
  roofline   the dominant kernel of the step, timed live with CUDA events by the library's launch profiler
  cpu_baseline  the CPU oracle port (oracle/step_oracle.py) timed on this box's host cores on a bounded sample
`--impl reference` times that CPU port as its own arm (the reference is pure Python + PyTorch3D and cannot travel to
the GPU box; see DESIGN.md "Reference arm").
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "field+Chamfer train-step foot-vertices/s"
UNIT = "vertices/s"

WORKLOADS = {
    # name: (feet per GPU, V, P2, frozen_mlp)   — BASELINE.json configs
    "cfg2": dict(feet=8, V=10242, P2=10000, frozen=False, desc="FIND field fwd+bwd+Chamfer, batch 8 feet, 10k verts, 1xB200"),
    "cfg3": dict(feet=8, V=10242, P2=20000, frozen=False, desc="batch 64 feet, 10k verts, 20k target points, foot-batch sharded (8 feet/GPU)"),
    "cfg4": dict(feet=4, V=50000, P2=100000, frozen=False, desc="high-res template 50k verts, 100k target points, batch 32 (4 feet/GPU)"),
    "cfg5": dict(feet=32, V=10242, P2=10000, frozen=True, desc="latent-code fitting (frozen MLP), batch 256 feet (32 feet/GPU)"),
}

# algorithmic MACs of the tensor-core kernels (the latent columns of mlp_*.0 are folded into
# a per-foot bias and x.B is done with FMAs, so neither is credited)
MAC_TRUNK_FWD = 515 * 256 + 4 * 65536            # per template vertex
MAC_HEAD_FWD = 3 * 65536 + 768                   # per foot-vertex per head
MAC_HEAD_BWD = 768 + 2 * 65536                   # dX chain of one head
MAC_TRUNK_BWD = 2 * 65536 + 4 * 65536            # both heads' dX into the trunk + 4 trunk layers
MAC_DW_HEAD = 3 * 65536 + 768                    # per foot-vertex per head
MAC_DW_TRUNK = 4 * 65536 + 515 * 256             # per template vertex
FLOP_PER_PAIR = 8                                # 3 sub + 1 mul + 2 fma(=4) ... counted as 8 flop / pair-evaluation (§8(d))


class ClockSampler:
    """SM clock / throttle reasons / power sampled DURING the timed region (B200_PROFILING.md's clocks line).  The timed
    region is short (tens of ms), so the sampler reads NVML in-process every ~2 ms (nvidia_ml_py); `nvidia-smi -lms 100`
    is the fallback when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []
        self.samples, self.stop_flag, self.h, self.nv = [], False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nv = pynvml
        except Exception:
            self.nv = None

    def start(self):
        if self.nv is not None:
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _poll(self):
        nv, h = self.nv, self.h
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8,
                "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}
        smax = None
        try:
            smax = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
        except Exception:
            pass
        while not self.stop_flag:
            try:
                sm = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1e3
                except Exception:
                    pw = None
                self.samples.append((sm, smax, pw, [n for n, b in bits.items() if mask & b]))
            except Exception:
                break
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        sm, smax, reasons, power = [], [], set(), []
        if self.nv is not None:
            self.stop_flag = True
            self.th.join(timeout=2)
            for c, m, pw, rs in self.samples:
                sm.append(c)
                if m: smax.append(m)
                if pw is not None: power.append(pw)
                reasons.update(rs)
            how = "nvml in-process, 2 ms cadence"
        else:
            if self.proc is None:
                return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            for ln in self.lines:
                f = [x.strip() for x in ln.split(",")]
                if len(f) < 7:
                    continue
                try:
                    sm.append(float(f[0])); smax.append(float(f[1])); power.append(float(f[2]))
                except ValueError:
                    continue
                for n, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            how = "nvidia-smi -lms 100"
        # "under load" = samples in the upper half of the observed clock range (idle samples sit at 120-300 MHz)
        load = [c for c in sm if c >= 0.5 * max(sm)] if sm else []
        return {"sm_mhz": statistics.median(load) if load else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons), "how": how}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            d = json.load(fh)
        return dict(hbm_gbs=d["hbm_gbs"], bf16_burst=d["bf16_tflops"], bf16_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, bf16_burst=1590.0, bf16_sustained=1400.0, source="fallback (B200_PROFILING.md)")


# ---------------------------------------------------------------------------------------------------------------
def build_job(wl, rank, world, device):
    """Model + TrainStep + this rank's host-side inputs (pinned)."""
    from find_b200 import synthetic as syn
    from find_b200.field import NeuralDisplacementField
    from find_b200.step import TrainStep
    feet, V, P2 = wl["feet"], wl["V"], wl["P2"]
    global_feet = feet * world
    foot0 = rank * feet
    model = NeuralDisplacementField(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=global_feet,
                                    val_size=1, shapevec_size=100, texvec_size=100, posevec_size=100, device=device)
    tv, tf = syn.make_template(V)
    model.set_template(tv, tf)
    syn.randomise_last_disp_layer(model.mlp_disp[-1], seed=0)
    lat = syn.make_latents(feet, foot0)
    with torch.no_grad():
        for k in ("shapevec", "posevec", "texvec", "reg"):
            getattr(model, k).data[foot0:foot0 + feet] = lat[k].to(device)
    step = TrainStep(model, feet, P2, with_colour=True, frozen_mlp=wl["frozen"])
    host = dict(idx=torch.arange(foot0, foot0 + feet, dtype=torch.int64).pin_memory(),
                target=syn.make_targets(feet, P2, foot0).pin_memory(),
                col_target=syn.make_col_target(feet, V, foot0).pin_memory())
    return model, step, host


def kernel_flops(name, wl, tpf_pad_rows=None):
    """Algorithmic FLOPs of ONE launch of a named kernel for this rank's workload (None: not a math kernel)."""
    B, V, P2 = wl["feet"], wl["V"], wl["P2"]
    fv = B * V
    if name == "trunk_fwd": return 2.0 * MAC_TRUNK_FWD * V
    if name == "head_fwd": return 2.0 * MAC_HEAD_FWD * fv * 2
    if name == "head_bwd": return 2.0 * MAC_HEAD_BWD * fv * 2
    if name == "trunk_bwd": return 2.0 * MAC_TRUNK_BWD * V
    if name == "dw_gemms":
        return 2.0 * (MAC_DW_HEAD * fv * 2 + MAC_DW_TRUNK * V) if not wl["frozen"] else None
    if name == "knn_main": return float(FLOP_PER_PAIR) * 2.0 * B * V * P2
    return None


def run_ours(args, wl, rank, world, local_rank):
    import torch.distributed as dist
    from find_b200 import _lib
    from find_b200.parallel import ShardedTrainStep
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    lib = _lib.load()
    peaks = load_peaks()
    model, step, host = build_job(wl, rank, world, device)
    sharded = ShardedTrainStep(step, wl["feet"] * world, collective=os.environ.get("FINDB200_COLLECTIVE", "auto"))
    feet, V, P2 = wl["feet"], wl["V"], wl["P2"]
    fv_global = feet * world * V
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)     # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    step.load_inputs(host["idx"], host["target"], host["col_target"])
    for _ in range(max(args.warmup, 3)):
        sharded.run()
    barrier()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()

    # ---- (1) device-resident: K steps, CUDA events around each, L2 flushed between ------------------------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.perf_counter()
    for a, b in ev:
        flush.zero_()
        a.record()
        sharded.run()
        b.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    per_step = sorted(a.elapsed_time(b) for a, b in ev)                     # this rank's steps (median + p10 / p90)
    step_stats = {"median": per_step[len(per_step) // 2], "p10": per_step[len(per_step) // 10],
                  "p90": per_step[min(len(per_step) - 1, (9 * len(per_step)) // 10)], "min": per_step[0], "max": per_step[-1]}
    ms_dev = sum(per_step)
    ms_dev = max_over_ranks(ms_dev)
    ms_per_step = ms_dev / args.steps
    loss3 = step.loss3.tolist()

    # ---- (2) end to end through the public host-facing call: H2D inputs + D2H loss inside the timed region ------
    for _ in range(2):
        step.step_from_host(host["idx"], host["target"], host["col_target"])
    barrier()
    iso_s = 0.0                                    # isolated steps (L2 flushed, nothing overlapped): H2D -> run -> D2H, serial
    for _ in range(min(args.steps, 20)):
        flush.zero_()
        torch.cuda.synchronize(device)
        t0 = time.perf_counter()
        step.load_inputs(host["idx"], host["target"], host["col_target"])
        sharded.run()
        losses = step.loss3.tolist()              # D2H + sync (trainer.py:128-130 reads the losses every step)
        iso_s += time.perf_counter() - t0
    iso_s = max_over_ranks(iso_s) / min(args.steps, 20)
    # the training-loop form of the same API (TrainStep.iterate_from_host): batch k+1 crosses PCIe on a copy stream while step
    # k computes; every step still copies its own inputs from pinned memory and reads its losses back inside the timed region.
    # The working set of a step (> 0.6 GB of saved activations) is far larger than L2, so back-to-back steps start cold.
    batch = (host["idx"], host["target"], host["col_target"])
    for _ in step.iterate_from_host([batch] * 3, runner=sharded):
        pass
    barrier()
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for losses in step.iterate_from_host((batch for _ in range(args.steps)), runner=sharded):
        pass
    torch.cuda.synchronize(device)
    e2e_s = time.perf_counter() - t0
    barrier()
    e2e_s = max_over_ranks(e2e_s)
    h2d = host["idx"].numel() * 8 * 4 + host["target"].numel() * 4 + host["col_target"].numel() * 4
    clocks = sampler.stop() if rank == 0 else None

    # ---- (3) live per-kernel timing (library launch profiler, non-graph enqueue, L2 flushed between steps) -----
    n_prof = 5
    buf = ctypes.create_string_buffer(1 << 16)
    kernels, timing_mode = [], None
    if world == 1:
        # in-graph: the marks become event-record nodes of a freshly captured step graph, so a kernel's time is what it takes
        # inside the replayed graph (the form `value` is measured in), not a stream launch with the host's gaps
        try:
            torch.cuda.synchronize(device)
            pg = torch.cuda.CUDAGraph()
            lib.findb200_profile_enable(1)
            with torch.cuda.graph(pg):
                step._enqueue()
            lib.findb200_profile_enable(2)
            acc = {}
            for _ in range(n_prof):
                flush.zero_()
                pg.replay()
                torch.cuda.synchronize(device)
                nbytes = lib.findb200_profile_read(buf, len(buf))
                for ln in buf.value.decode().splitlines() if nbytes > 0 else []:
                    name, cnt, tot = ln.split(",")
                    a = acc.setdefault(name, [0, 0.0])
                    a[0] += int(cnt); a[1] += float(tot)
            kernels = [dict(name=k, launches_per_step=v[0] / n_prof, ms_per_step=v[1] / n_prof) for k, v in acc.items()]
            timing_mode = "CUDA events recorded as nodes of the replayed step graph (no host launch gaps)"
            del pg
        except Exception as e:                       # fall back to stream-order marks
            print(f"# in-graph kernel timing unavailable ({type(e).__name__}: {str(e)[:120]}); using stream-order marks", file=sys.stderr)
            kernels = []
        finally:
            lib.findb200_profile_enable(0)
    if not kernels:
        step.use_graph = False
        lib.findb200_profile_enable(1)
        for _ in range(n_prof):
            flush.zero_()
            sharded.run()
        torch.cuda.synchronize(device)
        nbytes = lib.findb200_profile_read(buf, len(buf))
        lib.findb200_profile_enable(0)
        step.use_graph = True
        for ln in buf.value.decode().splitlines() if nbytes > 0 else []:
            name, cnt, tot = ln.split(",")
            kernels.append(dict(name=name, launches_per_step=int(cnt) / n_prof, ms_per_step=float(tot) / n_prof))
        timing_mode = "CUDA events between stream launches (non-graph pass; includes the host's launch gaps)"
    launches = int(round(sum(k["launches_per_step"] for k in kernels)))
    # "side:" entries (the presort of the Chamfer: target sort + template order, csrc/chamfer_hooks.cuh) run on a side stream
    # beside the forward kernels: they are launches of the step, but their time is not part of the main stream's chain
    for k in kernels:
        if k["name"].startswith("side:"):
            k["overlapped"] = True
    prof_total = sum(k["ms_per_step"] for k in kernels if not k.get("overlapped")) or 1.0
    # The dW plan is launched in two parts (csrc/field.cu): the early CTAs on a side stream beside sum_feet and the trunk's
    # backward chain, the late CTAs after it.  The marks are stream-ordered on the main stream, so the three marks
    # sum_feet / trunk_bwd / dw_gemms_late tile the span from the fork to the join, and the dW kernel is running through all of
    # it: its duration is the SPAN (it shares the machine with the other two), which is what its rates are computed on.
    late = next((k for k in kernels if k["name"] == "dw_gemms_late"), None)
    if late is not None:
        span = sum(k["ms_per_step"] for k in kernels if k["name"] in ("sum_feet", "trunk_bwd", "dw_gemms_late"))
        merged = dict(name="dw_gemms", launches_per_step=1.0, ms_per_step=span, overlapped=True,
                      note="one plan = two launches (side:dw_gemms_early + dw_gemms_late, listed on their own and counted in "
                           "gpu_launches there); ms = span from the fork to the join = sum_feet + trunk_bwd + dw_gemms_late marks, "
                           "which run beside the early launch: shares of the overlapped entries sum to more than the step")
        early = next((k for k in kernels if k["name"] == "side:dw_gemms_early"), None)
        if early is not None:
            # SM-time of the plan: the early launch holds n_early SMs, the late one the rest (csrc/field.cu: n_early = the SMs the
            # trunk chain's CTA pairs leave free, even) — the duration the same SM-seconds would take on the whole GPU
            sms = torch.cuda.get_device_properties(device).multi_processor_count
            n_early = (sms - 2 * ((V + 255) // 256)) & ~1
            merged["early_ms"], merged["late_ms"], merged["early_ctas"], merged["late_ctas"] = (
                early["ms_per_step"], late["ms_per_step"], n_early, sms - n_early)
            merged["sm_time_ms"] = (n_early * early["ms_per_step"] + (sms - n_early) * late["ms_per_step"]) / sms
        kernels.append(merged)
    fp32_peak = None
    try:
        scratch = torch.zeros(16, device=device)
        blocks, iters = torch.cuda.get_device_properties(device).multi_processor_count * 8, 2048
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        lib.findb200_fp32_peak(1, blocks, iters, _lib.ptr(scratch), _lib.stream_ptr(device))
        e0.record()
        lib.findb200_fp32_peak(1, blocks, iters, _lib.ptr(scratch), _lib.stream_ptr(device))
        e1.record()
        torch.cuda.synchronize(device)
        fp32_peak = blocks * 256 * iters * 64 * 2 * 2 / (e0.elapsed_time(e1) * 1e-3) / 1e12
    except Exception:
        pass
    # pair evaluations the culled kNN kernel really executed in the last step (kernel counter, not an estimate)
    knn_stats = (ctypes.c_ulonglong * 4)()
    rc = lib.findb200_train_step_knn_stats(_lib.ptr(step.ws), feet, V, P2, knn_stats, _lib.stream_ptr(device))
    pairs_exec = int(knn_stats[0]) if rc == 0 and knn_stats[0] else None
    for k in kernels:
        k["share"] = k["ms_per_step"] / prof_total
        fl = kernel_flops(k["name"], wl)
        if fl is not None and k["ms_per_step"] > 0:
            if k["name"] == "knn_main":
                # only EXECUTED pair evaluations are credited against the FP32 peak; the brute-force-equivalent rate
                # (what the reference's knn_points would have to sustain for the same time) is reported beside it
                k["tflops_bruteforce_equiv"] = fl / (k["ms_per_step"] * 1e-3) / 1e12
                k["pairs_algorithmic"], k["pairs_executed"] = int(fl / FLOP_PER_PAIR), pairs_exec
                if pairs_exec is not None:
                    fl = float(FLOP_PER_PAIR) * pairs_exec
                k["bound"], k["peak"] = "fp32", fp32_peak
            else:
                k["bound"], k["peak"] = "tensor", peaks["bf16_sustained"]
            k["tflops"] = fl / (k["ms_per_step"] * 1e-3) / 1e12
            k["frac"] = k["tflops"] / k["peak"] if k["peak"] else None
    # dW GEMMs: every saved activation / dZ tile is read exactly once (fp16 tile images), which makes the
    # launch HBM-bound before it is tensor-bound: algorithmic bytes = 512 B per 256-wide row of each operand
    for k in kernels:
        if k["name"] == "dw_gemms" and not wl["frozen"] and k["ms_per_step"] > 0:
            fv_loc = feet * V
            by = fv_loc * 2 * (6 * 512 + 32) + V * (4 * 512 + 9 * 128 + 5 * 512)
            k["hbm_bytes_algorithmic"] = by
            k["hbm_gbs"] = by / (k["ms_per_step"] * 1e-3) / 1e9
            k["hbm_frac"] = k["hbm_gbs"] / peaks["hbm_gbs"]
    math_k = [k for k in kernels if "tflops" in k]
    dom = max(math_k, key=lambda k: k["ms_per_step"]) if math_k else None
    roofline = None
    traffic_all = {}
    try:       # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            traffic_all = json.load(fh).get(args.workload, {})
    except (OSError, ValueError):
        pass
    traffic = traffic_all.get(dom["name"] if dom else "", None)
    # whole-step fractions: EXECUTED tensor-core FLOPs of all GEMM kernels / step time / tensor peak, and the DRAM traffic of
    # all kernels (ncu capture, per launch) / step time / HBM peak.  Neither roof binds the step as a whole.
    tensor_flops_step = sum(kernel_flops(k["name"], wl) * k["launches_per_step"] for k in kernels
                            if k.get("bound") == "tensor")
    step_tensor_frac = tensor_flops_step / (ms_per_step * 1e-3) / 1e12 / peaks["bf16_sustained"]
    dram_step = sum(v for v in traffic_all.values() if isinstance(v, (int, float)))
    step_hbm_frac = (dram_step / (ms_per_step * 1e-3) / 1e9 / peaks["hbm_gbs"]) if dram_step else None
    if dom is not None:
        # SURVEY.md §8(d) puts the MLP GEMMs on the TENSOR roof (1.7 MFLOP per 36 B of unavoidable I/O): that fraction is the
        # primary `frac`.  The dW launch streams the saved fp16 tile images once (this design's traffic, not algorithmic I/O);
        # its HBM rate is reported beside it as `design_traffic_*`.
        roofline = dict(kernel=dom["name"], bound="tensor" if dom["bound"] == "tensor" else "hbm",
                        achieved=round(dom["tflops"], 2), peak=dom["peak"], unit="TFLOP/s",
                        frac=round(dom["frac"], 4) if dom["frac"] else None, traffic=traffic,
                        kernel_ms=round(dom["ms_per_step"], 4), share_of_step=round(dom["share"], 3),
                        step_tensor_frac=round(step_tensor_frac, 4),
                        step_hbm_frac=round(step_hbm_frac, 4) if step_hbm_frac is not None else None,
                        step_tensor_tflops=round(tensor_flops_step / (ms_per_step * 1e-3) / 1e12, 1),
                        peak_source=("findb200_fp32_peak FFMA2 probe, same run (FP32 pipe; reported under 'hbm' only because the "
                                     "contract has two labels)" if dom["bound"] == "fp32"
                                     else peaks["source"] + " bf16_tflops_sustained (fp16 operands, same tensor rate)"))
        if dom.get("hbm_frac") is not None:
            roofline.update(design_traffic_bytes=dom["hbm_bytes_algorithmic"], design_traffic_gbs=round(dom["hbm_gbs"], 1),
                            design_traffic_hbm_frac=round(dom["hbm_frac"], 4),
                            design_traffic_note="bytes of the saved fp16 tile images this design streams through the dW launch "
                                                "(not algorithmic I/O); HBM peak = " + str(peaks["hbm_gbs"]) + " GB/s measured")
        if dom.get("sm_time_ms"):
            # the plan shares the GPU with sum_feet / trunk_bwd during its span: `frac` above charges it the whole span on the
            # whole GPU; these charge it the SM-time it really holds (early CTAs x early duration + late CTAs x late duration)
            r = dom["ms_per_step"] / dom["sm_time_ms"]
            roofline.update(overlapped_with=["sum_feet", "trunk_bwd"], sm_time_ms=round(dom["sm_time_ms"], 4),
                            early_launch_ms=round(dom["early_ms"], 4), late_launch_ms=round(dom["late_ms"], 4),
                            early_ctas=dom["early_ctas"], late_ctas=dom["late_ctas"],
                            frac_on_sm_time=round(dom["frac"] * r, 4) if dom["frac"] else None)
            if dom.get("hbm_frac") is not None:
                roofline["design_traffic_hbm_frac_on_sm_time"] = round(dom["hbm_frac"] * r, 4)

    out = {
        "metric": METRIC, "value": fv_global / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "ms_per_step_rank0": {k: round(v, 5) for k, v in step_stats.items()}, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp16 tensor-core operands / fp32 accumulate (field), fp32 (Chamfer)",
        "data": "synthetic", "impl": "find_b200",
        "config": {"workload": f"{args.workload}: {wl['desc']}", "feet_per_gpu": feet, "global_feet": feet * world,
                   "verts": V, "target_points": P2, "frozen_mlp": wl["frozen"], "parallelism": f"dp{world} (foot-batch sharding)", "collective": sharded.collective,
                   "l2": "flushed between steps (256 MiB memset outside the event pairs)", "cuda_graph": True,
                   "loss": "1e4*chamfer(verts,target) + 1*mse(col,col_target)"},
        "e2e": {"value": fv_global / (e2e_s / args.steps), "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 12, "ms_per_step": e2e_s / args.steps * 1e3,
                "isolated_ms_per_step": iso_s * 1e3,
                "api": "find_b200.step.TrainStep.iterate_from_host (pinned host -> H2D of batch k+1 on a copy stream, straight into the input set that step k does not read (one captured graph per set, no device-to-device copy), while step k "
                       "replays -> losses of every step copied D2H asynchronously and read one step later); isolated_ms_per_step = step_from_host with nothing overlapped, L2 flushed"},
        "gpu_launches": launches * args.steps,
        "gpu_launches_per_step": launches,
        "roofline": roofline,
        "kernel_timing": timing_mode,
        "kernels": [{k2: (round(v, 4) if isinstance(v, float) else v) for k2, v in k.items()} for k in kernels],
        "loss": {"total": loss3[0], "chamfer": loss3[1], "colour_mse": loss3[2]},
        "wall_s_timed_region": t_wall,
        "clocks": clocks,
    }
    if world == 1 and not wl["frozen"]:
        out["shipped_3d_step"] = measure_shipped_step(model, step, wl, host, device, flush, min(args.steps, 50))
    if world == 1 and not args.skip_eager:
        out["gpu_eager_baseline"] = gpu_eager_baseline(wl, device, flush)
    # BASELINE.json's other configurations at this N (per-GPU shapes, weak scaling), a few steps each, so that the driver's
    # BENCH / SCALE records carry a value for every config
    if not args.skip_other_workloads and args.workload == "cfg2":
        del sharded, step, model
        torch.cuda.empty_cache()
        out["other_workloads"] = {}
        for name in ("cfg3", "cfg4", "cfg5"):
            out["other_workloads"][name] = measure_workload_brief(name, rank, world, device, flush, barrier, max_over_ranks,
                                                                  steps=min(args.steps, 10))
    if rank == 0 and world == 1 and not args.skip_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(wl, budget_s=args.cpu_budget)
    return out


def measure_workload_brief(name, rank, world, device, flush, barrier, max_over_ranks, steps=10, warmup=3):
    """Device-timed value of another BASELINE.json configuration at this N: same protocol as the headline (graph replay, CUDA
    events per step, L2 flushed between steps, max over ranks), fewer steps."""
    from find_b200.parallel import ShardedTrainStep
    wl = WORKLOADS[name]
    model, step, host = build_job(wl, rank, world, device)
    sharded = ShardedTrainStep(step, wl["feet"] * world, collective=os.environ.get("FINDB200_COLLECTIVE", "auto"))
    step.load_inputs(host["idx"], host["target"], host["col_target"])
    for _ in range(warmup):
        sharded.run()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush.zero_()
        a.record()
        sharded.run()
        b.record()
    barrier()
    ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev)) / steps
    loss3 = step.loss3.tolist()
    res = {"workload": f"{name}: {wl['desc']}", "feet_per_gpu": wl["feet"], "global_feet": wl["feet"] * world, "verts": wl["V"],
           "target_points": wl["P2"], "frozen_mlp": wl["frozen"], "ms_per_step": ms, "steps": steps, "warmup": warmup,
           "value": wl["feet"] * world * wl["V"] / (ms * 1e-3), "unit": UNIT, "collective": sharded.collective,
           "loss": {"total": loss3[0], "chamfer": loss3[1], "colour_mse": loss3[2]}}
    del sharded, step, model
    torch.cuda.empty_cache()
    return res


def gpu_eager_baseline(wl, device, flush, steps=5):
    """The comparators BASELINE.md §4 promises ("the kernel to beat"): the reference's own module arithmetic
    (model.py:393-453 + the registration transform 481-488) as torch-eager ops on THIS GPU — cuBLAS GEMMs + autograd — with
    an eager chunked Chamfer (the brute-force distance matrix PyTorch3D's knn kernel walks, here through torch.cdist-free
    broadcast + min), in fp32, with allow_tf32, and under bf16 autocast.  Library code end to end: none of this repo's
    kernels.  Same batch, same loss, forward + backward, optimiser excluded; median of `steps` after 2 warm-ups."""
    import torch.nn.functional as F
    from find_b200 import synthetic as syn
    from find_b200.field import NeuralDisplacementField
    feet, V, P2 = wl["feet"], wl["V"], wl["P2"]
    model = NeuralDisplacementField(use_shapevec=True, use_texvec=True, use_posevec=True, train_size=feet, val_size=1,
                                    shapevec_size=100, texvec_size=100, posevec_size=100, device=device)
    syn.randomise_last_disp_layer(model.mlp_disp[-1], seed=0)
    tv, _ = syn.make_template(V)
    tv = tv.to(device)
    lat = {k: v.to(device).requires_grad_(True) for k, v in syn.make_latents(feet, 0).items()}
    target = syn.make_targets(feet, P2, 0).to(device)
    colt = syn.make_col_target(feet, V, 0).to(device)
    Bm = model.encoder[0].B_on(device)
    params = [p for p in model.mlp_parameters()]
    for p in params:
        p.requires_grad_(not wl["frozen"])

    def mlp(seq, x):
        for m in seq:
            x = F.relu(x) if isinstance(m, torch.nn.ReLU) else F.linear(x, m.weight, m.bias)
        return x

    def euler(e):
        a, b, c = e.unbind(-1)
        one, zero = torch.ones_like(a), torch.zeros_like(a)
        ca, sa, cb, sb, cc, sc = a.cos(), a.sin(), b.cos(), b.sin(), c.cos(), c.sin()
        Rx = torch.stack([one, zero, zero, zero, ca, -sa, zero, sa, ca], -1).reshape(-1, 3, 3)
        Ry = torch.stack([cb, zero, sb, zero, one, zero, -sb, zero, cb], -1).reshape(-1, 3, 3)
        Rz = torch.stack([cc, -sc, zero, sc, cc, zero, zero, zero, one], -1).reshape(-1, 3, 3)
        return Rx @ Ry @ Rz

    def nn_sq(x, y, chunk=2048):
        """min_j |x_i - y_j|^2: the argmin over chunks of the exact-difference distance matrix under no_grad (what knn_points
        computes), then the differentiable distance to the gathered neighbours (what its autograd differentiates)."""
        with torch.no_grad():
            idx = []
            for i in range(0, x.shape[1], chunk):
                d = x[:, i:i + chunk, None, :] - y[:, None, :, :]
                idx.append((d * d).sum(-1).argmin(-1))
            idx = torch.cat(idx, 1)
        nb = torch.gather(y, 1, idx[..., None].expand(-1, -1, 3))
        return ((x - nb) ** 2).sum(-1)

    def step_fn():
        pos = tv[None].expand(feet, -1, -1)
        proj = 2 * torch.pi * (pos @ Bm)                                    # fourier_feature_transform.py:41-52
        x = torch.cat([pos, torch.sin(proj), torch.cos(proj)], -1)
        x = mlp(model.base, x)
        ex = lambda z: z[:, None, :].expand(-1, V, -1)
        disp = 0.1 * torch.tanh(mlp(model.mlp_disp, torch.cat([x, ex(lat["shapevec"]), ex(lat["posevec"])], -1)))
        col = 0.5 * (1 + torch.tanh(mlp(model.mlp_col, torch.cat([x, ex(lat["texvec"])], -1))))
        reg = lat["reg"]
        verts = ((pos + disp).float() * reg[:, None, 6:9]) @ euler(reg[:, 3:6]) + reg[:, None, 0:3]
        verts = verts.float()
        chamf = (nn_sq(verts, target).mean(1) + nn_sq(target, verts).mean(1)).mean()
        loss = 1e4 * chamf + F.mse_loss(col.float(), colt)
        for t in params + list(lat.values()):
            t.grad = None
        loss.backward()
        return loss

    res = {"what": "torch-eager (cuBLAS + autograd) field fwd+bwd + registration transform + chunked broadcast Chamfer on this GPU; "
                   "same batch / loss as the headline, optimiser excluded, L2 flushed between steps, median of %d steps" % steps,
           "feet": feet, "verts": V, "target_points": P2}
    old_tf32 = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    try:
        for mode in ("fp32", "tf32", "bf16_autocast"):
            torch.backends.cuda.matmul.allow_tf32 = mode == "tf32"
            torch.backends.cudnn.allow_tf32 = mode == "tf32"
            ctx = torch.autocast("cuda", dtype=torch.bfloat16) if mode == "bf16_autocast" else torch.autocast("cuda", enabled=False)

            def run():
                with ctx:
                    return step_fn()
            try:
                for _ in range(2):
                    run()
                torch.cuda.synchronize(device)
                times = []
                for _ in range(steps):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record(); loss = run(); e1.record()
                    torch.cuda.synchronize(device)
                    times.append(e0.elapsed_time(e1))
                ms = statistics.median(times)
                res[mode] = {"ms_per_step": ms, "vertices_per_s": feet * V / (ms * 1e-3), "loss": float(loss)}
            except Exception as e:                      # noqa: BLE001 — e.g. out of memory on the distance matrix chunks
                res[mode] = {"error": str(e)[:200]}
                torch.cuda.empty_cache()
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old_tf32
    for t in params:
        t.grad = None
        t.requires_grad_(True)
    del model
    torch.cuda.empty_cache()
    return res


def measure_shipped_step(model, base_step, wl, host, device, flush, steps):
    """Secondary figure (not the headline metric): the reference's shipped 3-D configuration (cfgs/train_3d.yaml: chamf on
    5000 surface samples + smooth + tex) through findb200_train_step_mesh, followed by the optimiser updates of
    train.py:161-168 (FusedAdam on the MLP and the latent tables, FusedSGD momentum on reg) — one full training iteration."""
    from find_b200.optim import FusedAdam, FusedSGD
    from find_b200.step import TrainStep
    del base_step
    step = TrainStep(model, wl["feet"], wl["P2"], with_colour=True, num_samples=5000, weight_smooth=1000.0, seed=1)
    step.load_inputs(host["idx"], host["target"], host["col_target"])
    tables = [t.data for t in step.tables.values() if t is not None]
    lat = [model.shapevec.data, model.posevec.data, model.texvec.data]
    opt_net = FusedAdam(step.mlp_params, lr=0.0)              # lr 0: the timed updates must not drift the benchmark weights
    opt_lat = FusedAdam(lat, lr=0.0)
    opt_reg = FusedSGD([model.reg.data], lr=0.0, momentum=0.9)

    def iteration():
        step.run()                      # incl. the dense latent-table gradient scatter (part of the enqueue)
        step.populate_grads()           # attaches the gradient buffers, no launch
        opt_net.step(); opt_lat.step(); opt_reg.step()

    for _ in range(3):
        iteration()
    torch.cuda.synchronize(device)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush.zero_()
        a.record(); iteration(); b.record()
    torch.cuda.synchronize(device)
    ms = sorted(a.elapsed_time(b) for a, b in ev)[len(ev) // 2]
    total, chamf, tex = step.loss3.tolist()
    edge, lap = step.loss_extra.tolist()
    for t in tables:
        t.grad = None
    for p_ in step.mlp_params:
        p_.grad = None
    return {"ms_per_iteration_median": ms, "vertices_per_s": wl["feet"] * wl["V"] / (ms * 1e-3),
            "what": "findb200_train_step_mesh (Chamfer on 5000 surface samples/foot + 1000*MeshSmoothnessLoss + colour) + "
                    "latent-grad scatter + FusedAdam(MLP) + FusedAdam(latents) + FusedSGD(reg); step is a graph replay, "
                    "optimiser launches are eager",
            "loss": {"total": total, "chamfer": chamf, "colour_mse": tex, "edge": edge, "laplacian": lap}}


# ---------------------------------------------------------------------------------------------------------------
def cpu_step_fn(wl, feet_sample):
    """-> (callable running one CPU step on `feet_sample` feet of the workload, threads used)."""
    from find_b200 import synthetic as syn
    from oracle import field_oracle as fo
    from oracle import step_oracle as so
    from oracle import chamfer_oracle as co
    V, P2 = wl["V"], wl["P2"]
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    Bm = fo.fourier_B()
    p = fo.random_params(0)
    tv, _ = syn.make_template(V)
    lat = syn.make_latents(feet_sample, 0)
    tables = {k: lat[k] for k in ("shapevec", "posevec", "texvec", "reg")}
    idx = torch.arange(feet_sample)
    target = syn.make_targets(feet_sample, P2, 0)
    colt = syn.make_col_target(feet_sample, V, 0)
    co.build()

    def run():
        return so.train_step(p, Bm, tv, tables, idx, target, colt, threads=threads, need_param_grads=not wl["frozen"])
    return run, threads


def cpu_feet_sample(wl):
    """Feet per CPU step: the WHOLE per-GPU batch when one step stays around a second (cfg2 / cfg3: 8 feet = 0.4 s on 16
    cores), otherwise a bounded sample so that the arm ends within minutes (cfg4: 1 of 4 feet; cfg5: 8 of 32)."""
    cost = wl["feet"] * (wl["V"] * 5.0e-6 + 2.0 * wl["V"] * wl["P2"] * 2.0e-11)      # rough seconds on ~16 cores
    if cost <= 2.0:
        return wl["feet"]
    return max(1, min(wl["feet"], int(wl["feet"] * 2.0 / cost)))


def cpu_baseline(wl, budget_s=20.0, steps=None, warmup=1):
    feet_sample = cpu_feet_sample(wl)
    run, threads = cpu_step_fn(wl, feet_sample)
    for _ in range(warmup):
        run()
    times = []
    t_start = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
        if steps is not None and len(times) >= steps:
            break
        if steps is None and (time.perf_counter() - t_start > budget_s or len(times) >= 30):
            break
    med = statistics.median(times)
    whole = feet_sample == wl["feet"]
    return {"value": feet_sample * wl["V"] / med, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": (f"the whole batch of the GPU arm ({feet_sample} feet x {wl['V']} verts x {wl['P2']} targets) per step"
                       if whole else
                       f"{feet_sample} of the {wl['feet']} feet of the {wl['V']}-vertex / {wl['P2']}-target batch per step "
                       f"(the step is linear in feet)")
                      + f", median of {len(times)} steps after {warmup} warm-up; "
                        f"oracle/step_oracle.py: torch CPU autograd (field) + OpenMP C Chamfer",
            "feet_per_step": feet_sample, "s_per_step": med}


def run_reference(args, wl, rank, world):
    """The reference arm: the CPU port of the reference's path on the host cores (rank 0 only)."""
    if rank != 0:
        return None
    base = cpu_baseline(wl, steps=args.steps, warmup=max(1, min(args.warmup, 2)))
    return {"metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": base["s_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"{args.workload}: {wl['desc']}", "feet_per_gpu": wl["feet"], "feet_per_step": base["feet_per_step"],
                       "verts": wl["V"], "target_points": wl["P2"], "frozen_mlp": wl["frozen"]},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference = pure Python on torch + PyTorch3D (absent, not installable offline): this arm times the "
                    "oracle PORT of its algorithm (cpu_baseline.kind = 'port') on the host cores, same batch per step"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="find_b200", choices=["find_b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-eager", action="store_true", help="skip the torch-eager GPU comparators (gpu_eager_baseline)")
    ap.add_argument("--skip-other-workloads", action="store_true", help="skip the brief cfg3 / cfg4 / cfg5 lines")
    ap.add_argument("--cpu-budget", type=float, default=20.0)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload is None:
        args.workload = "cfg2"        # same per-GPU shape at every N (weak scaling: 8 feet / GPU)
    wl = WORKLOADS[args.workload]

    if args.impl == "reference":
        out = run_reference(args, wl, rank, world)    # the same batch as the GPU arm per step (0.4 s for cfg2 on 16 cores)
        if out is not None:
            print(json.dumps(out), flush=True)
        return

    import torch.distributed as dist
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    rc = 0
    try:
        out = run_ours(args, wl, rank, world, local_rank)
        if rank == 0:
            print(json.dumps(out), flush=True)
    except BaseException:                       # noqa: BLE001 — report, then leave with a non-zero code
        import traceback
        traceback.print_exc()
        rc = 1
    if world > 1:
        # leave without tearing NCCL down: the communicator is referenced by a live CUDA graph (the captured allreduce) and
        # destroy_process_group() on such a communicator can block.  Everything is flushed and synchronised first.
        try:
            if rc == 0:
                dist.barrier()
                torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(rc)
    if rc:
        sys.exit(rc)


if __name__ == "__main__":
    main()
