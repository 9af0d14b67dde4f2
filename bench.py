#!/usr/bin/env python
"""bench.py -- measures the deePore hot path on B200 (driver contract: see DESIGN.md "measurement").

One STEP = one interval shard (`--windows` consecutive windows of BASELINE.json configs[1]: synthetic 30x
nanopore-length reads over a 9 Mb reference, window centres every 100 bp) pushed through the hot path:
pileup encode (+ model forward and vote when `--stage e2e`).  Successive steps take successive intervals,
so every step's inputs and outputs are new (outputs of one step, 120 kB/window, exceed L2).

  value   genomic positions/s with the step's inputs already resident in HBM
  e2e     the same metric through the public host-buffer API: pinned host shard -> one H2D copy -> kernels
          -> D2H of the per-window result, all inside the timed region
  roofline / cpu_baseline / clocks / gpu_launches: as the contract asks

`--impl reference` times the CPU implementation of the same path (oracle port of the reference's Pileup.py;
the reference itself is Python + pysam and cannot travel to the GPU box) on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "genomic positions/sec (pileup encode + CNN/GRU infer)"
UNIT = "positions/s"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        try:
            d = json.load(open(p))
            return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                    "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    "source": "measured"}
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------------
def cpu_encode_rate(shard, seconds_budget=12.0, threads=None):
    """Oracle port (oracle/pileup_oracle.c, a C restatement of Pileup.py) on the host cores: windows/s.
    ctypes releases the GIL, so one Python thread per core drives one slice of the sample."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pileup_oracle as po
    po.lib()
    threads = threads or os.cpu_count() or 1
    # calibrate on a few windows, then size the sample to the budget
    n0 = min(shard.n_windows, 16)
    t = time.perf_counter()
    po.encode_windows(shard.reads, shard.win_qs[:n0], shard.win_lo[:n0], shard.win_hi[:n0], shard.ref, shard.ref_base)
    per = max((time.perf_counter() - t) / n0, 1e-6)
    n = int(min(shard.n_windows, max(threads * 8, seconds_budget / per * threads)))
    idx = np.array_split(np.arange(n), threads)

    def work(ix):
        if ix.size:
            po.encode_windows(shard.reads, shard.win_qs[ix], shard.win_lo[ix], shard.win_hi[ix], shard.ref,
                              shard.ref_base)
    ths = [threading.Thread(target=work, args=(ix,)) for ix in idx]
    t = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    dt = time.perf_counter() - t
    return n / dt, n, threads, dt


def run_reference(args, rank, world):
    """--impl reference: CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from deepore_b200.workload import SyntheticWorkload
    wl = SyntheticWorkload(stride=args.stride)
    nsh = 1
    sh = wl.shards(0, min(args.windows, 2048), nsh)[0]
    threads = os.cpu_count() or 1
    rates = []
    budget = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    n = 0
    for i in range(args.warmup + args.steps):
        wps, n, threads, dt = cpu_encode_rate(sh, seconds_budget=budget, threads=threads)
        if i >= args.warmup:
            rates.append(wps)
    wps = float(np.mean(rates))
    val = wps * args.stride
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n / wps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: synthetic 30x nanopore-length reads over 9Mb reference, stride %d"
                       % args.stride, "stage": args.stage, "windows_per_step": n},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d windows of the first shard per step, oracle/pileup_oracle.c on %d threads"
                             % (n, threads)},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--stage", default="encode", choices=["encode", "e2e"])
    ap.add_argument("--windows", type=int, default=8192, help="windows per step (interval shard size)")
    ap.add_argument("--stride", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    from deepore_b200 import _lib
    from deepore_b200.pileup import Encoder, ShardBuffers
    from deepore_b200.workload import SyntheticWorkload, algorithmic_bytes

    _lib.require_cuda()
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- workload: rank r owns a contiguous run of shards (weak scaling: fixed work per GPU)
    nsteps = args.steps + args.warmup
    wl = SyntheticWorkload(stride=args.stride)
    first = rank * nsteps * args.windows
    shards = wl.shards(first, args.windows, nsteps)
    while len(shards) < nsteps:                                   # wrapped past the contig end
        shards += wl.shards(0, args.windows, nsteps - len(shards))
    bufs = [ShardBuffers(sh, dev) for sh in shards]
    enc = Encoder(device=dev)
    nw = args.windows
    H, W = enc.H, enc.W
    # two rotating output sets (the model consumes one while the next is encoded)
    outs = [dict(img=torch.empty((nw, H, W, 4), dtype=torch.uint8, device=dev),
                 rowpos=torch.empty((nw, H), dtype=torch.int32, device=dev),
                 rowins=torch.empty((nw, H), dtype=torch.int32, device=dev),
                 status=torch.empty((nw,), dtype=torch.int32, device=dev)) for _ in range(2)]
    host_status = torch.empty((nw,), dtype=torch.int32).pin_memory()
    host_rowpos = torch.empty((nw, H), dtype=torch.int32).pin_memory()

    def step_device(i):
        sb = bufs[i]
        o = outs[i & 1]
        n = sb.n_windows
        enc.encode(sb, o["img"][:n], o["rowpos"][:n], o["rowins"][:n], o["status"][:n])
        return 2  # launches: prep_reads_kernel + encode_windows_kernel

    def step_e2e(i):
        sb = bufs[i]
        sb.upload()
        k = step_device(i)
        o = outs[i & 1]
        n = sb.n_windows
        host_status[:n].copy_(o["status"][:n], non_blocking=True)
        host_rowpos[:n].copy_(o["rowpos"][:n], non_blocking=True)
        return k

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, sampler=None):
        for b in bufs:
            b.upload()
        for i in range(args.warmup):
            fn(i)
        barrier()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches = 0
        per_step = []
        evs = []
        e0.record()
        for i in range(args.warmup, nsteps):
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record()
            launches += fn(i)
            b.record()
            evs.append((a, b))
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = e0.elapsed_time(e1)
        per_step = [a.elapsed_time(b) for a, b in evs]
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, clocks, per_step

    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms_dev, launches, clocks, per_step = timed(step_device, sampler)
    ms_e2e, _, _, _ = timed(step_e2e)

    win_timed = sum(b.n_windows for b in bufs[args.warmup:])
    tot_windows = win_timed * world
    value = tot_windows * args.stride / (ms_dev * 1e-3)
    e2e_value = tot_windows * args.stride / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = load_peaks()
        alg = sum(algorithmic_bytes(sh) for sh in shards[args.warmup:])
        # the encode kernel pair is the whole device step at --stage encode: its mean duration per launch
        kern_ms = float(np.mean(per_step))
        achieved = alg / args.steps / (kern_ms * 1e-3) / 1e9
        h2d = int(np.mean([b.nbytes for b in bufs[args.warmup:]]))
        d2h = int(nw * 4 + nw * H * 4)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": "configs[1]: synthetic 30x nanopore-length reads over 9Mb reference, "
                           "window centres every %d bp" % args.stride, "stage": args.stage,
                           "windows_per_step": nw, "l2": "each step reads a new shard and writes %.0f MB (> L2)"
                           % (nw * H * W * 4 / 1e6)},
                "windows_per_s": tot_windows / (ms_dev * 1e-3),
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": achieved / peaks["hbm_gbs"], "traffic": None, "peak_source": peaks["source"],
                             "kernel": "prep_reads_kernel+encode_windows_kernel",
                             "algorithmic_bytes_per_window": alg / max(1, win_timed)},
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": launches, "clocks": clocks}
        if not args.no_cpu_baseline:
            wps, n, threads, dt = cpu_encode_rate(shards[args.warmup], seconds_budget=12.0)
            line["cpu_baseline"] = {"value": wps * args.stride, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": "%d windows of the first timed shard, oracle/pileup_oracle.c "
                                              "(C restatement of Pileup.py) on %d threads, %.1f s" % (n, threads, dt)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
