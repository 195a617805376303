#!/usr/bin/env python
"""Headline benchmark: ResNet-50 batch-32 inference (BASELINE.json configs[1]) through Model.run on B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--math tf32_bf16x2|3xtf32|tf32]

One step = one Graph::run over one batch of 32 synthetic 3x224x224 images per GPU (weak scaling: the batch of
Graph::run is sharded per GPU, no data-path collective).  Prints ONE JSON line on rank 0.

  value      images/s with the input batch already resident in HBM (device-timed, CUDA events, L2 flushed
             between steps, max over ranks)
  e2e        the same through the public API with pinned HOST input and HOST output (H2D + D2H inside the timed
             region)
  roofline   the tcgen05 contraction kernel (all conv / fc launches): algorithmic flops / measured launch time vs
             the measured TF32 tensor peak (MEASURED_PEAKS.json bf16 / 2)
  cpu_baseline  the oracle's CPU restatement of the reference (same graph, same weights) on the host cores

`--impl reference` times only that CPU restatement (the reference is Rust; no Rust toolchain here — DESIGN.md).
"""
import argparse
import contextlib
import io
import json
import re
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 32
IMAGE = 224
CLASSES = 1000
FLOP_PER_IMAGE = None  # filled from the graph


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--math", default="tf32_bf16x2", choices=["tf32_bf16x2", "3xtf32", "tf32"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget for the bounded cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of one step before timing")
    ap.add_argument("--no-extras", action="store_true", help="skip the GEMM sweep / N-split / small-batch sections")
    return ap.parse_args()


def graph_flops(builder, batch):
    """Algorithmic flops of the contraction operators (2*M*N*K per Conv / Gemm / MatMul) for `batch` images."""
    shapes = {}
    for i, nd in enumerate(builder.nodes):
        if nd["kind"] == "const":
            shapes[i] = nd["data"].shape
    total = 0.0
    # walk with a shape-only evaluation: convs need the spatial size of their output
    hw = {builder.inputs[0]: (IMAGE, IMAGE)}
    for nd in builder.nodes:
        if nd["kind"] != "op":
            continue
        t, ins, out = nd["type"], nd["inputs"], nd["outputs"][0]
        attrs = {fid: val for fid, _k, val in nd["attrs"][1]} if nd["attrs"] else {}
        if t == "Conv":
            o, cg, kh, kw = shapes[ins[1]]
            h, w = hw[ins[0]]
            pads = attrs.get(1) or (0, 0, 0, 0)
            s = attrs.get(3) or (1, 1)
            oh = (h + pads[0] + pads[2] - kh) // s[0] + 1
            ow = (w + pads[1] + pads[3] - kw) // s[1] + 1
            hw[out] = (oh, ow)
            total += 2.0 * batch * oh * ow * o * cg * kh * kw
        elif t in ("MaxPool", "AveragePool"):
            h, w = hw[ins[0]]
            k = attrs[0]
            pads = attrs.get(2) or (0, 0, 0, 0)
            s = attrs.get(3) or (1, 1)
            hw[out] = ((h + pads[0] + pads[2] - k[0]) // s[0] + 1, (w + pads[1] + pads[3] - k[1]) // s[1] + 1)
        elif t == "Gemm":
            n, k = shapes[ins[1]]
            total += 2.0 * batch * n * k
        elif t == "GlobalAveragePool":
            hw[out] = (1, 1)
        else:
            if ins and ins[0] in hw:
                hw[out] = hw[ins[0]]
    return total


class ClockSampler:
    """Samples the SM clock and the clock-event (throttle) reasons during the timed region: NVML polled every ~2 ms from
    a thread (the timed region lasts tens of milliseconds, too short for `nvidia-smi -lms`), nvidia-smi as the
    fallback (B200_PROFILING.md clocks line)."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index, uuid=None):
        self.index, self.uuid = index, uuid
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self.thread, self.stop_flag, self.source = None, threading.Event(), None

    def _nvml_handle(self):
        import pynvml

        pynvml.nvmlInit()
        if self.uuid:
            try:
                return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(self.uuid)).encode())
            except Exception:
                pass
        return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def _poll_nvml(self, nv, h):
        names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                for bit, nm in names.items():
                    if mask & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.002)

    def _poll_smi(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
                r = [c.strip() for c in out.strip().split(",")]
                self.sm.append(float(r[0]))
                self.max_mhz = max(self.max_mhz or 0.0, float(r[1]))
                for nm, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        self.reasons.add(nm)
            except Exception:
                time.sleep(0.05)

    def start(self):
        try:
            nv, h = self._nvml_handle()
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.source = "nvml"
            self.thread = threading.Thread(target=self._poll_nvml, args=(nv, h), daemon=True)
        except Exception:
            self.source = "nvidia-smi"
            self.thread = threading.Thread(target=self._poll_smi, daemon=True)
        self.thread.start()

    def stop(self):
        if not self.thread:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["not sampled"]}
        self.stop_flag.set()
        self.thread.join(timeout=15)
        if not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "samples": 0, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(self.sm)), "sm_min_mhz": float(min(self.sm)), "sm_max_mhz": self.max_mhz,
                "samples": len(self.sm), "source": self.source, "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------- CPU restatement arm
def physical_cores():
    """The reference sizes its thread pool to the PHYSICAL cores (num_cpus::get_physical, src/threading.rs:41-63),
    RTEN_NUM_THREADS overriding it."""
    env = os.environ.get("RTEN_NUM_THREADS")
    if env and env.isdigit() and int(env) >= 1:
        return int(env)
    try:
        import psutil

        n = psutil.cpu_count(logical=False)
        if n:
            return int(n)
    except Exception:
        pass
    return os.cpu_count() or 1


def host_avx2_peak_gflops(threads):
    """cores x 2 FMA ports x 8 lanes x 2 flop x clock, the bound the reference quotes for its kernel
    (src/gemm.rs:1871-1886); clock = the highest 'cpu MHz' the kernel reports right now."""
    mhz = 0.0
    try:
        for l in open("/proc/cpuinfo"):
            if l.lower().startswith("cpu mhz"):
                mhz = max(mhz, float(l.split(":")[1]))
    except Exception:
        pass
    return threads * 2 * 8 * 2 * mhz / 1e3 if mhz else None, mhz


def cpu_resnet(builder, seconds, batch_cap=BATCH, steps=None, warmup=1, extras=False):
    """Times the oracle's CPU evaluation of the same ResNet-50 graph (same weights) the way the reference would run
    it: thread pool = physical cores (src/threading.rs:44), operator outputs from a buffer pool and Relu / Add in place
    (src/graph.rs:897-931, src/tensor_pool.rs) — `graph_oracle.run_graph(reuse=True)`.
    Returns (images/s, threads, sample description, ms per step, batch, extra lines)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import graph_oracle
    import oracle as o

    o.build()
    logical = o.num_threads()
    threads = max(1, min(physical_cores(), logical))
    o.set_num_threads(threads)
    rng = o.XorShiftRng(1234)
    pool = graph_oracle._Pool()
    run = lambda x: graph_oracle.run_graph(builder, [x], reuse=True, pool=pool)  # noqa: E731
    # calibrate on 2 images so that the bounded sample fits the budget
    x2 = rng.fill((2, 3, IMAGE, IMAGE))
    run(x2)
    t0 = time.perf_counter()
    run(x2)
    per_img = (time.perf_counter() - t0) / 2
    if steps is None:
        batch = int(max(1, min(batch_cap, seconds / 3 / per_img)))
        steps = int(max(1, min(5, seconds / (per_img * batch) - warmup)))
    else:
        batch = int(max(1, min(batch_cap, seconds / (steps + warmup) / per_img)))
    x = rng.fill((batch, 3, IMAGE, IMAGE))
    for _ in range(warmup):
        run(x)
    t0 = time.perf_counter()
    for _ in range(steps):
        run(x)
    dt = time.perf_counter() - t0
    ips = batch * steps / dt
    sample = (f"ResNet-50 full graph, batch {batch} x {steps} steps ({warmup} warm-up), CPU restatement of the reference "
              f"(AVX2 6x16 kernel, reference blocking, pooled buffers + in-place Relu/Add), {threads} threads = physical "
              f"cores ({logical} logical)")
    more = {}
    if extras:
        gflop_img = FLOP_PER_IMAGE / 1e9 if FLOP_PER_IMAGE else 8.178
        peak, mhz = host_avx2_peak_gflops(threads)
        more["gflops"] = ips * gflop_img
        if peak:
            more["fraction_of_avx2_fma_peak"] = ips * gflop_img / peak
            more["avx2_fma_peak_gflops"] = peak
            more["cpu_mhz"] = mhz
        # one thread (1 image), all logical CPUs, and a torch / oneDNN context line: informational
        o.set_num_threads(1)
        p1 = graph_oracle._Pool()
        x1 = rng.fill((1, 3, IMAGE, IMAGE))
        graph_oracle.run_graph(builder, [x1], reuse=True, pool=p1)
        t0 = time.perf_counter()
        graph_oracle.run_graph(builder, [x1], reuse=True, pool=p1)
        more["one_thread_images_per_s"] = 1.0 / (time.perf_counter() - t0)
        if logical != threads:
            o.set_num_threads(logical)
            run(x)
            t0 = time.perf_counter()
            run(x)
            more["all_logical_cpus_images_per_s"] = batch / (time.perf_counter() - t0)
        o.set_num_threads(threads)
        try:
            more["torch_cpu_context_images_per_s"] = torch_cpu_resnet(builder, x, threads)
        except Exception as exc:  # context only
            more["torch_cpu_context_error"] = str(exc)[:100]
    return ips, threads, sample, dt / steps * 1e3, batch, more


def torch_cpu_resnet(builder, x, threads):
    """Context line only: the same graph through torch's CPU kernels (oneDNN), same thread count."""
    import torch
    import torch.nn.functional as F

    torch.set_num_threads(threads)
    consts = {i: torch.from_numpy(nd["data"]) for i, nd in enumerate(builder.nodes) if nd["kind"] == "const"}

    def run(inp):
        vals = dict(consts)
        vals[builder.inputs[0]] = inp
        for nd in builder.nodes:
            if nd["kind"] != "op":
                continue
            a = {fid: val for fid, _k, val in nd["attrs"][1]} if nd["attrs"] else {}
            ins = [vals[i] if i is not None and i >= 0 else None for i in nd["inputs"]]
            t = nd["type"]
            if t == "Conv":
                pads = a.get(1) or (0, 0, 0, 0)
                y = F.conv2d(ins[0], ins[1], ins[2], stride=tuple(a.get(3) or (1, 1)), padding=(pads[0], pads[1]),
                             groups=a.get(2, 1))
            elif t == "Relu":
                y = F.relu(ins[0])
            elif t == "Add":
                y = ins[0] + ins[1]
            elif t == "MaxPool":
                pads = a.get(2) or (0, 0, 0, 0)
                y = F.max_pool2d(ins[0], tuple(a[0]), tuple(a.get(3) or (1, 1)), (pads[0], pads[1]))
            elif t == "GlobalAveragePool":
                y = ins[0].mean((2, 3), keepdim=True)
            elif t == "Flatten":
                y = ins[0].flatten(1)
            elif t == "Gemm":
                y = F.linear(ins[0], ins[1], ins[2])
            else:
                raise NotImplementedError(t)
            vals[nd["outputs"][0]] = y
        return vals[builder.outputs[0]]

    xt = torch.from_numpy(x)
    with torch.no_grad():
        run(xt)
        t0 = time.perf_counter()
        run(xt)
        return x.shape[0] / (time.perf_counter() - t0)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from rten_fork_b200 import synth

    builder = synth.resnet50(image=IMAGE, classes=CLASSES, seed=0)
    global FLOP_PER_IMAGE
    FLOP_PER_IMAGE = graph_flops(builder, 1)
    ips, threads, sample, ms, batch, more = cpu_resnet(builder, seconds=150.0, steps=args.steps,
                                                       warmup=max(1, min(args.warmup, 3)), extras=True)
    line = {
        "impl": "reference", "metric": "resnet50_b32_images_per_sec", "value": ips, "unit": "images/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ResNet-50 .rten, random-init weights, batch 32 3x224x224", "sample_batch": batch},
        "cpu_baseline": dict({"value": ips, "unit": "images/s", "cores": threads, "kind": "port", "sample": sample}, **more),
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------ GPU arm
def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tf32_burst": d["bf16_tflops"] / 2, "tf32_sustained": d["bf16_tflops_sustained"] / 2,
                "source": "MEASURED_PEAKS.json (bf16 cuBLAS / 2 for TF32)"}
    return {"hbm_gbs": 6650.0, "tf32_burst": 795.0, "tf32_sustained": 795.0, "source": "fallback (B200_PROFILING.md, bf16 1.59 PF / 2)"}


def timed_device_steps(torch, model, stream, x_dev, out_dev, flush, steps):
    """K steps, each bracketed by CUDA events on the launching stream, L2 flushed (256 MiB write) before each."""
    evs = []
    with torch.cuda.stream(stream):
        for _ in range(steps):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            model.run([x_dev], out=[out_dev])
            e1.record(stream)
            evs.append((e0, e1))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs)


def timed_e2e_steps(torch, model, stream, x_host, out_host, flush, steps):
    """K steps through the public API with pinned host input / output: H2D, graph, D2H and the final synchronise are
    inside each step's timed span; the L2 flush between steps is not."""
    total = 0.0
    with torch.cuda.stream(stream):
        for _ in range(steps):
            flush.zero_()
            stream.synchronize()
            t0 = time.perf_counter()
            model.run([x_host], out=[out_host])   # returns after the D2H copy has completed
            total += time.perf_counter() - t0
    return total * 1e3


def timed_e2e_pipelined(torch, model, stream, x_hosts, out_hosts, flush, steps):
    """K steps through Model.run_async (rtg_model_submit) with pinned host buffers, two submits in flight: every step's
    H2D upload, graph and D2H download are inside the window, the upload of step i + 1 overlapping the graph of step i.
    The L2 flush of each step is enqueued on the compute stream and therefore also inside the window."""
    with torch.cuda.stream(stream):
        stream.synchronize()
        t0 = time.perf_counter()
        prev = None
        for i in range(steps):
            flush.zero_()
            tk = model.run_async([x_hosts[i & 1]], [out_hosts[i & 1]])
            if prev is not None:
                prev.wait()
            prev = tk
        prev.wait()
        return (time.perf_counter() - t0) * 1e3


def gemm_section(torch, R, ops, ctx, stream, modes, peak_tf32):
    """The "f32 GEMM TFLOPS" half of BASELINE.json's metric on one GPU: C1 = rten-bench's 1024^3 case under bench_gemm's
    iteration rule (src/gemm.rs:1782-1903: iters = min(1000, 512^3 * 1000 / (m n k)) back-to-back calls, TFLOPS =
    2 m n k iters / t) and the C5 square sweep 512 ... 16384, per math mode.  Device-resident operands, CUDA events on
    the launching stream; `graph` = the same calls replayed from one CUDA graph (no host launch cost between them)."""
    dev = torch.device("cuda", ctx.device)
    out = {"unit": "TFLOP/s", "peak": peak_tf32, "rule": "bench_gemm: iters = min(1000, 512^3*1000/(m*n*k)), >= 3",
           "c1_1024": {}, "sweep": {}}
    gen = torch.Generator(device=dev).manual_seed(1234)
    with torch.cuda.stream(stream):
        for n in (512, 1024, 2048, 4096, 8192, 16384):
            a = torch.rand((n, n), device=dev, generator=gen)
            b = torch.rand((n, n), device=dev, generator=gen)
            c = torch.empty((n, n), device=dev)
            iters = int(max(3, min(1000, 512 ** 3 * 1000 // n ** 3)))
            for name, m in modes.items():
                ctx.set_math(m)
                for _ in range(3):
                    ops.matmul(a, b, out=c, ctx=ctx)
                stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(iters):
                    ops.matmul(a, b, out=c, ctx=ctx)
                e1.record(stream)
                stream.synchronize()
                us = e0.elapsed_time(e1) / iters * 1e3
                rec = {"us": round(us, 2), "tflops": round(2.0 * n ** 3 / us / 1e6, 1), "iters": iters,
                       "kernel": ctx.last_kernel}
                rec["frac"] = round(rec["tflops"] / peak_tf32, 4)
                if n <= 2048:  # launch-bound sizes: the same calls from one CUDA graph
                    try:
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, stream=stream):
                            for _ in range(iters):
                                ops.matmul(a, b, out=c, ctx=ctx)
                        g.replay()
                        stream.synchronize()
                        e0.record(stream)
                        g.replay()
                        e1.record(stream)
                        stream.synchronize()
                        gus = e0.elapsed_time(e1) / iters * 1e3
                        rec["graph_us"] = round(gus, 2)
                        rec["graph_tflops"] = round(2.0 * n ** 3 / gus / 1e6, 1)
                        del g
                    except Exception as exc:  # capture is an extra, never a reason to lose the line
                        rec["graph_error"] = str(exc)[:120]
                        torch.cuda.synchronize()
                out["sweep"].setdefault(str(n), {})[name] = rec
                if n == 1024:
                    out["c1_1024"][name] = rec
            del a, b, c
        # the memory-bound M <= 8 path (gemv kernels): bench_gemm's 1 x 4096 x 512 case in both B layouts
        # (src/gemm.rs:1790-1838) and the ResNet classifier head at batch 1 / 8, against the measured HBM copy rate
        hbm = load_peaks()["hbm_gbs"]
        out["gemv"] = {"unit": "GB/s", "peak": hbm, "cases": {}}
        ctx.set_math(list(modes.values())[0])
        for name, (m, n, k, tb) in {"1x4096x512": (1, 4096, 512, False), "1x4096x512_transposed_b": (1, 4096, 512, True),
                                    "fc_1x1000x2048": (1, 1000, 2048, True), "fc_8x1000x2048": (8, 1000, 2048, True),
                                    "1x16384x4096": (1, 16384, 4096, False)}.items():
            a = torch.rand((m, k), device=dev, generator=gen)
            b = torch.rand((n, k), device=dev, generator=gen).t() if tb else torch.rand((k, n), device=dev, generator=gen)
            c = torch.empty((m, n), device=dev)
            for _ in range(3):
                ops.matmul(a, b, out=c, ctx=ctx)
            try:
                g = torch.cuda.CUDAGraph()
                iters = 200
                with torch.cuda.graph(g, stream=stream):
                    for _ in range(iters):
                        ops.matmul(a, b, out=c, ctx=ctx)
                g.replay()
                stream.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                g.replay()
                e1.record(stream)
                stream.synchronize()
                us = e0.elapsed_time(e1) / iters * 1e3
                by = 4.0 * (k * n + m * k + m * n)
                out["gemv"]["cases"][name] = {"us": round(us, 2), "gbs": round(by / us / 1e3, 1),
                                              "frac": round(by / us / 1e3 / hbm, 3), "kernel": ctx.last_kernel,
                                              "note": "200 calls replayed from one CUDA graph; operands L2-resident "
                                                      "when they fit (8 MB), so small cases can exceed the HBM rate"}
                del g
            except Exception as exc:
                out["gemv"]["cases"][name] = {"error": str(exc)[:120]}
                torch.cuda.synchronize()
    return out


def hbm_section(torch, R, ops, ctx, stream):
    """The HBM-bound kernels either side of the contraction (SURVEY.md section 8 a7/a11/a12 and the depthwise branch of a4):
    algorithmic bytes (every operand read once, the result written once) / time against the measured HBM copy rate.
    Tensors are larger than L2 (126 MB) wherever the layer shape allows; CUDA events on the launching stream, 10 calls.
    Softmax and LayerNormalization have no operator-level entry point: one-operator models, replayed from their graph."""
    import numpy as np
    from rten_fork_b200.rten_builder import ModelBuilder

    dev = torch.device("cuda", ctx.device)
    hbm = load_peaks()["hbm_gbs"]
    out = {"unit": "GB/s", "peak": hbm, "cases": {}}
    gen = torch.Generator(device=dev).manual_seed(99)

    def record(name, us, nbytes, kernel):
        out["cases"][name] = {"us": round(us, 1), "gbs": round(nbytes / us / 1e3, 1),
                              "frac": round(nbytes / us / 1e3 / hbm, 3), "mbytes": round(nbytes / 1e6, 1), "kernel": kernel}

    def timed(fn, nbytes, name, kernel, iters=10):
        try:
            for _ in range(3):
                fn()
            stream.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(iters):
                fn()
            e1.record(stream)
            stream.synchronize()
            us = e0.elapsed_time(e1) / iters * 1e3
            record(name, us, nbytes, kernel)
        except Exception as exc:  # an extra: never a reason to lose the bench line
            out["cases"][name] = {"error": str(exc)[:160]}
            torch.cuda.synchronize()

    def nhwc(*shape):
        return torch.rand(shape, device=dev, generator=gen).contiguous(memory_format=torch.channels_last)

    with torch.cuda.stream(stream):
        # depthwise 3x3, the two largest MobileNetV2 layers at batch 64 (stride 1 and stride 2)
        for name, (c, hw, s) in {"depthwise3x3_s1_32ch_112": (32, 112, 1), "depthwise3x3_s2_96ch_112": (96, 112, 2)}.items():
            x = nhwc(64, c, hw, hw)
            w = torch.rand((c, 1, 3, 3), device=dev, generator=gen) - 0.5
            b = torch.rand((c,), device=dev, generator=gen)
            ohw = (hw + 2 - 3) // s + 1
            y = torch.empty((64, c, ohw, ohw), device=dev).contiguous(memory_format=torch.channels_last)
            timed(lambda: ops.conv(x, w, b, padding=(1, 1, 1, 1), groups=c, strides=(s, s), activation="clip",
                                   clip=(0.0, 6.0), out=y, ctx=ctx), 4.0 * (x.numel() + y.numel() + w.numel()), name,
                  f"depthwise3x3_tma<s{s}>")
            del x, y
        # elementwise: Relu in place / out of place, Add (ResNet's residual form: same shapes)
        x = torch.rand((64 << 20,), device=dev, generator=gen) - 0.5
        z = torch.rand((64 << 20,), device=dev, generator=gen)
        timed(lambda: ops.relu(x, in_place=True, ctx=ctx), 8.0 * x.numel(), "relu_in_place_256MB", "unary_kernel")
        y = torch.empty_like(x)
        timed(lambda: ops.add(x, z, out=y, ctx=ctx), 12.0 * x.numel(), "add_256MB", "binary_kernel")
        del x, y, z
        # MaxPool 3x3/2 of ResNet-50's stem at batch 32 (NHWC, as Graph::run sees it)
        x = nhwc(32, 64, 112, 112)
        y = torch.empty((32, 64, 56, 56), device=dev).contiguous(memory_format=torch.channels_last)
        timed(lambda: ops.max_pool(x, (3, 3), (2, 2), (1, 1, 1, 1), ctx=ctx, out=y), 4.0 * (x.numel() + y.numel()),
              "maxpool3x3_s2_64ch_112", "maxpool_nhwc4_rows2_kernel<3,3,2>")
        del x, y
        # Softmax over BERT's score rows and LayerNormalization over its hidden rows, scaled past L2
        for name, op, cols, attrs, extra in (
                ("softmax_rows128", "Softmax", 128, ("SoftmaxAttrs", [(0, "i32", -1)]), 0),
                ("layer_norm_rows768", "LayerNormalization", 768,
                 ("LayerNormalizationAttrs", [(0, "i32", -1), (1, "f32", 1e-12)]), 2)):
            try:
                rows = (200 << 20) // (4 * cols)
                gb = ModelBuilder()
                xin = gb.add_value("x", [rows, cols])
                ins = [xin]
                if extra:
                    ins += [gb.add_constant("gamma", np.ones(cols, np.float32)), gb.add_constant("beta", np.zeros(cols, np.float32))]
                yv = gb.op(op, ins, attrs, name="op")
                gb.inputs, gb.outputs = [xin], [yv]
                m = R.Model.load(gb.build(), ctx=ctx)
                xd = torch.rand((rows, cols), device=dev, generator=gen)
                yd = torch.empty((rows, cols), device=dev)
                # the operator's own device time (RunOptions timing: events around the operator), not the run's: a
                # one-operator model also copies its input in and its output out
                m.set_timing(True)
                best = None
                for _ in range(5):
                    m.run([xd], out=[yd])
                    hit = re.search(op + r"\s+([0-9.]+)ms", m.timing_report())
                    if hit:
                        best = float(hit.group(1)) if best is None else min(best, float(hit.group(1)))
                if best is None:
                    raise RuntimeError("no timing row for " + op)
                record(name, best * 1e3, 8.0 * rows * cols, "softmax_rows_kernel" if op == "Softmax" else "layer_norm_rows_kernel")
                out["cases"][name]["timed"] = "operator time from the model's timing report, best of 5"
                m.close()
                del xd, yd
            except Exception as exc:
                out["cases"][name] = {"error": str(exc)[:160]}
                torch.cuda.synchronize()
    return out


def nsplit_section(torch, dist, R, ctx, stream, math_name, world, rank):
    """BASELINE.json configs[4] at N GPUs: square GEMM, B column-split over the ranks, result on every rank.  local = this
    rank's block alone, nccl = block + one NCCL all-gather + strided copy, fused = tiles TMA-stored into every rank's
    buffer from the GEMM epilogue over NVLink (rtg_gemm_scatter) + 4-byte all-reduces.  Max over ranks."""
    from rten_fork_b200 import ops
    from rten_fork_b200.parallel import SymmetricMatrix, column_blocks, gemm_nsplit, gemm_nsplit_fused

    dev = torch.device("cuda", ctx.device)
    res = {}

    def timed(fn, iters):
        for _ in range(2):
            fn()
        stream.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(iters):
            fn()
        e1.record(stream)
        stream.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    with torch.cuda.stream(stream):
        for n, iters in ((4096, 10), (8192, 5), (16384, 3)):
            g = torch.Generator(device=dev).manual_seed(1234)
            a = torch.rand((n, n), device=dev, generator=g)
            b = torch.rand((n, n), device=dev, generator=g)
            lo, hi = column_blocks(n, world)[rank]
            blk = torch.empty((n, max(hi - lo, 1)), device=dev)
            ctx.set_stream(stream.cuda_stream)
            t_local = timed(lambda: ops.matmul(a, b[:, lo:hi], out=blk[:, : hi - lo], ctx=ctx) if hi > lo else None, iters)
            t_nccl = timed(lambda: gemm_nsplit(a, b, ctx=ctx), iters)
            sym = SymmetricMatrix(n, n, ctx=ctx)
            t_fused = timed(lambda: gemm_nsplit_fused(a, b, sym), iters)
            ref = gemm_nsplit(a, b, ctx=ctx)
            stream.synchronize()
            same = torch.tensor([1 if torch.equal(ref, sym.tensor) else 0], device=dev)
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            sym.close()
            fl = 2.0 * n ** 3
            gather = (world - 1) * 4 * n * n // world   # bytes every GPU receives from its peers
            res[str(n)] = {"local_ms": round(t_local, 4), "nccl_ms": round(t_nccl, 4), "fused_ms": round(t_fused, 4),
                           "nccl_tflops": round(fl / t_nccl / 1e9, 1), "fused_tflops": round(fl / t_fused / 1e9, 1),
                           "gather_bytes_per_gpu": gather,
                           "fused_nvlink_gbs_in": round(gather / (t_fused * 1e-3) / 1e9, 1),
                           "fused_equals_nccl_bitwise": bool(same.item())}
            del a, b, blk, ref
    return {"math": math_name, "gpus": world, "sizes": res,
            "note": "n x n x n; NVLink figure = bytes received per GPU / fused time (the stores overlap the contraction)"}


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import rten_fork_b200 as R
    from rten_fork_b200 import synth

    ctx = R.Context(local)
    stream = torch.cuda.Stream(device=local)
    ctx.set_stream(stream.cuda_stream)
    modes = {"tf32_bf16x2": R.MATH_TF32_BF16X2, "3xtf32": R.MATH_3XTF32, "tf32": R.MATH_TF32}
    math = modes[args.math]
    ctx.set_math(math)
    builder = synth.resnet50(image=IMAGE, classes=CLASSES, seed=0)
    flops_step = graph_flops(builder, args.batch)
    model = R.Model.load(builder.build(), ctx=ctx)

    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    x_host = torch.rand((args.batch, 3, IMAGE, IMAGE), generator=gen, dtype=torch.float32).pin_memory()
    out_host = torch.empty((args.batch, CLASSES), dtype=torch.float32).pin_memory()
    x_dev = x_host.to(f"cuda:{local}")
    out_dev = torch.empty((args.batch, CLASSES), dtype=torch.float32, device=f"cuda:{local}")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # parity of the benchmarked configuration itself, before (outside) any timed region: one step's logits against the
    # oracle's CPU evaluation of the same graph and input (tests/test_fullsize_gpu.py holds the per-layer version)
    parity = {"parity_checked": False}
    if rank == 0 and not args.no_parity:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import graph_oracle
        import oracle as o

        o.build()
        with torch.cuda.stream(stream):
            model.run([x_dev], out=[out_dev])
        stream.synchronize()
        want = graph_oracle.run_graph(builder, [x_host.numpy()])[0]
        got = out_dev.cpu().numpy()
        scale = float(np.abs(want).max())
        err = float(np.abs(got.astype(np.float64) - want).max())
        rtol = 2e-3 if args.math == "tf32" else 1e-5
        parity = {"parity_checked": True, "parity_ok": bool(err <= 20 * rtol * scale), "parity_max_err_over_scale": err / scale,
                  "parity_rule": f"max |logits - oracle| <= 20 x {rtol} x max|oracle| (53 convolutions deep; per-layer "
                                 "per-element checks are tests/test_fullsize_gpu.py)"}
        if not parity["parity_ok"]:
            raise SystemExit(f"bench.py: parity check failed before timing: {parity}")

    # warm-up (also fills the weight cache, tensor maps and the pool)
    timed_device_steps(torch, model, stream, x_dev, out_dev, flush, max(args.warmup, 3))
    sampler = ClockSampler(local, getattr(torch.cuda.get_device_properties(local), "uuid", None))
    if rank == 0:
        sampler.start()
    barrier()
    l0 = ctx.launch_count
    ms_total = timed_device_steps(torch, model, stream, x_dev, out_dev, flush, args.steps)
    headline_out = out_dev.cpu().numpy().copy()   # out_dev is reused by the other math modes below
    launches = ctx.launch_count - l0
    barrier()
    ms_total = max_over_ranks(ms_total)

    # end-to-end (host buffers)
    timed_e2e_steps(torch, model, stream, x_host, out_host, flush, 2)
    barrier()
    e2e_ms = timed_e2e_steps(torch, model, stream, x_host, out_host, flush, args.steps)
    barrier()
    e2e_ms = max_over_ranks(e2e_ms)
    same = bool(torch.equal(out_host.to(out_dev.device), out_dev))
    # the same, pipelined: what a serving loop calls (two host buffer pairs, both holding this rank's batch)
    x_hosts = [x_host, x_host.clone().pin_memory()]
    out_hosts = [out_host, torch.empty_like(out_host).pin_memory()]
    timed_e2e_pipelined(torch, model, stream, x_hosts, out_hosts, flush, 3)
    barrier()
    e2e_pipe_ms = timed_e2e_pipelined(torch, model, stream, x_hosts, out_hosts, flush, args.steps)
    barrier()
    e2e_pipe_ms = max_over_ranks(e2e_pipe_ms)
    same = same and bool(torch.equal(out_hosts[1].to(out_dev.device), out_dev)) and \
        bool(torch.equal(out_hosts[0].to(out_dev.device), out_dev))
    clocks = sampler.stop() if rank == 0 else None

    # dominant kernel: every launch of the tcgen05 contraction kernel bracketed by events on its own stream
    ctx.set_kernel_timing(True)
    kt_step_ms = timed_device_steps(torch, model, stream, x_dev, out_dev, flush, 3) / 3   # the step of this same pass
    kt = ctx.kernel_timing()
    ctx.set_kernel_timing(False)

    # the other contraction modes, for the record
    others = {}
    for name, m in modes.items():
        if m == math:
            continue
        ctx.set_math(m)
        timed_device_steps(torch, model, stream, x_dev, out_dev, flush, 3)
        o_ms = max_over_ranks(timed_device_steps(torch, model, stream, x_dev, out_dev, flush, args.steps))
        ctx.set_kernel_timing(True)
        timed_device_steps(torch, model, stream, x_dev, out_dev, flush, 3)
        others[name] = (o_ms, ctx.kernel_timing())
        ctx.set_kernel_timing(False)
    ctx.set_math(math)

    extras = {}
    if not args.no_extras:
        from rten_fork_b200 import ops

        if world > 1:
            extras["nsplit"] = nsplit_section(torch, dist, R, ctx, stream, args.math, world, rank)
        else:
            extras["gemm"] = gemm_section(torch, R, ops, ctx, stream, modes, load_peaks()["tf32_sustained"])
            extras["hbm_kernels"] = hbm_section(torch, R, ops, ctx, stream)
            ctx.set_math(math)
            # small batches, where the per-launch host cost weighs most (one GPU)
            small = {}
            for bsz in (1, 4):
                xs = x_dev[:bsz].contiguous()
                os_ = torch.empty((bsz, CLASSES), dtype=torch.float32, device=x_dev.device)
                timed_device_steps(torch, model, stream, xs, os_, flush, 3)
                ms = timed_device_steps(torch, model, stream, xs, os_, flush, args.steps) / args.steps
                small[str(bsz)] = {"ms_per_step": ms, "images_per_s": bsz / (ms * 1e-3)}
            extras["small_batch"] = small
        ctx.set_math(math)

    # strong scaling of ONE call: a batch of 32 host images through rtg_sharded_model_run over all N GPUs of this box
    # (rank 0's process drives them: a context, worker thread and weight replica per GPU; the other ranks wait)
    if not args.no_extras and world > 1:
        barrier()
        # the other ranks must wait on the HOST: an NCCL barrier would spin a kernel on their GPUs, which are shards here
        store = dist.distributed_c10d._get_default_store()
        if rank != 0:
            store.wait(["strong_done"])
        if rank == 0:
            try:
                sm = R.ShardedModel(builder.build(), list(range(world)), math=math)
                xs_h = x_host.numpy()
                os_h = np.empty((args.batch, CLASSES), np.float32)
                for _ in range(4):
                    sm.run([xs_h], [os_h])
                t0 = time.perf_counter()
                reps = max(args.steps, 10)
                for _ in range(reps):
                    sm.run([xs_h], [os_h])
                dt = (time.perf_counter() - t0) / reps
                extras["strong"] = {"gpus": world, "batch": args.batch, "ms_per_call": dt * 1e3,
                                    "images_per_s": args.batch / dt,
                                    "matches_one_gpu": bool(np.abs(os_h - headline_out).max() <= 2e-4 * np.abs(os_h).max()),
                                    "api": "rtg_sharded_model_run: pageable host input -> dim-0 chunks on N GPUs of one "
                                           "process -> host output; wall clock, copies inside"}
                sm.close()
            except Exception as exc:
                extras["strong"] = {"error": str(exc)[:200]}
            store.set("strong_done", "1")
        barrier()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = load_peaks()
    images = args.batch * world * args.steps
    tf = lambda k: k["flops"] / (k["ms"] * 1e-3) / 1e12 if k["ms"] > 0 else 0.0
    traffic = traffic_note = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic, traffic_note = tj.get(args.math), tj.get("note")
    line = {
        "metric": "resnet50_b32_images_per_sec", "value": images / (ms_total * 1e-3), "unit": "images/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"tf32_bf16x2": "f32 in/out; split a = hi + lo: hi*hi as TF32 plus one bf16 MMA for both correction "
                                 "products on tcgen05, fp32 accumulate (rtol 1e-5 vs the reference)",
                  "3xtf32": "f32 in/out; 3xTF32 split on tcgen05, fp32 accumulate (rtol 1e-5 vs the reference)",
                  "tf32": "tf32 (rtol 2e-3)"}[args.math],
        "data": "synthetic",
        "config": {"workload": "ResNet-50 .rten, random-init weights, batch 32 3x224x224 per GPU (Conv+MatMul)",
                   "batch_per_gpu": args.batch, "math": args.math, "l2": "256 MiB flush write before every step",
                   "gflop_per_step_per_gpu": flops_step / 1e9, "parallelism": f"batch-sharded x{world}, no collective"},
        "tflops": flops_step * world * args.steps / (ms_total * 1e-3) / 1e12,
        "clocks": clocks,
        "e2e": {"value": images / (e2e_pipe_ms * 1e-3), "unit": "images/s", "ms_per_step": e2e_pipe_ms / args.steps,
                "h2d_bytes_per_step": x_host.numel() * 4, "d2h_bytes_per_step": out_host.numel() * 4,
                "matches_device_run": same,
                "api": "Model.run_async / rtg_model_submit, pinned host buffers, two submits in flight: each step's "
                       "H2D, L2 flush, graph and D2H are inside the wall-clock window",
                "one_call_at_a_time": {"value": images / (e2e_ms * 1e-3), "unit": "images/s",
                                       "ms_per_step": e2e_ms / args.steps,
                                       "api": "Model.run / rtg_model_run, H2D -> graph -> D2H serialised per call"}},
        "gpu_launches": int(launches),
        "roofline": {"bound": "tensor", "kernel": "gemm_tcgen05_kernel (all Conv / Gemm launches of a step)",
                     "achieved": tf(kt), "peak": peaks["tf32_sustained"], "unit": "TFLOP/s",
                     "frac": tf(kt) / peaks["tf32_sustained"], "traffic": traffic, "traffic_note": traffic_note,
                     "launches_per_step": kt["launches"] // 3, "kernel_ms_per_step": kt["ms"] / 3,
                     # share of the SAME pass (step timed around the per-launch events): cannot exceed 1.  The plain
                     # step is shorter than this pass because programmatic dependent launch overlaps kernel heads
                     # and tails when no event sits between launches.
                     "kernel_share_of_step": (kt["ms"] / 3) / kt_step_ms,
                     "event_pass_ms_per_step": kt_step_ms,
                     "achieved_in_stream": flops_step / ((ms_total / args.steps) * 1e-3 * (kt["ms"] / 3) / kt_step_ms) / 1e12,
                     "algorithmic_gbytes_per_step": kt["bytes"] / 3 / 1e9,
                     "peak_source": peaks["source"] + ", sustained",
                     "note": "achieved = algorithmic flops (counted once; the fp32-accurate modes issue 2 (TF32 + bf16 "
                             "correction) or 3 (3xTF32) tensor-core passes per product) / the summed launch durations "
                             "of a pass with an event pair around every launch.  That pass serialises launches which "
                             "otherwise overlap head and tail (programmatic dependent launch), so it is slower than "
                             "the plain step: kernel_share_of_step is taken within that pass, achieved_in_stream "
                             "applies the share to the plain step (flops / (step x share)); the ncu launch list in "
                             "profiles/ gives the share under serialisation"},
        "other_modes": {name: {"value": images / (o_ms * 1e-3), "ms_per_step": o_ms / args.steps,
                               "kernel_tflops": tf(k), "kernel_frac": tf(k) / peaks["tf32_sustained"]}
                        for name, (o_ms, k) in others.items()},
    }
    line.update(parity)
    line.update(extras)
    if world == 1 and not args.no_cpu_baseline:
        global FLOP_PER_IMAGE
        FLOP_PER_IMAGE = flops_step / args.batch
        ips, threads, sample, _ms, _b, more = cpu_resnet(builder, seconds=args.cpu_seconds, extras=True)
        line["cpu_baseline"] = dict({"value": ips, "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
                                    **more)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    # stdout carries exactly ONE JSON line: libraries that write to file descriptor 1 behind Python's back (NCCL prints
    # its version banner there) are pointed at stderr until the line is printed
    sys.stdout.flush()
    _real_stdout = os.dup(1)
    os.dup2(2, 1)
    _buf = io.StringIO()
    with contextlib.redirect_stdout(_buf):
        if a.impl == "reference":
            run_reference(a)
        else:
            run_ours(a)
    sys.stdout.flush()
    os.dup2(_real_stdout, 1)
    lines = [l for l in _buf.getvalue().splitlines() if l.strip()]
    for l in lines[:-1]:
        print(l, file=sys.stderr)
    if lines:
        print(lines[-1], flush=True)
