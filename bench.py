#!/usr/bin/env python
"""bench.py -- measures the deePore hot path on B200 (driver contract: see DESIGN.md "measurement").

One STEP = one interval shard (`--windows` consecutive windows of BASELINE.json configs[1]: synthetic 30x
nanopore-length reads over a 9 Mb reference, window centres every 100 bp) pushed through the whole hot path:
pileup encode -> CNN/GRU forward over every sliding 3-row window -> argmax -> 3-way vote (`--stage encode` stops after
the encoder).  Successive steps take successive intervals, so every step's inputs and outputs are new (the pileup
tensors of one step, 120 kB/window, exceed L2).

  value   genomic positions/s with the step's inputs already resident in HBM
  e2e     the same metric through the public host-buffer API: pinned host shard -> one H2D copy -> kernels
          -> D2H of the per-window result, all inside the timed region
  roofline / cpu_baseline / clocks / gpu_launches: as the contract asks

`--impl reference` times the CPU implementation of the same path (oracle ports of the reference's Pileup.py and
model.py; the reference itself is Python + pysam + /root/reference and cannot travel to the GPU box) on all host cores.
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


def traffic_of(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the named kernel, from the committed `ncu --set full`
    capture of this bench command (profiles/r1_traffic.json); None when no capture is recorded."""
    p = os.path.join(ROOT, "profiles", "r1_traffic.json")
    try:
        return json.load(open(p))[kernel]["dram_bytes_per_launch"]
    except Exception:
        return None


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
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML is polled
    from a thread of this process (pynvml, every 20 ms; an `nvidia-smi -lms` child is the fallback when pynvml is
    missing).  start() long before the timed region -- the start-up (NVML initialisation, or fork + exec of nvidia-smi)
    must not fall inside it -- and mark() when it begins: only samples taken after mark() are used."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    # nvmlClocksEventReason* bits
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.nvml = None
        self.samples = []          # (time, sm_mhz, max_mhz, set of reasons)
        self.t0 = 0.0
        self.stop_flag = False

    def mark(self):
        self.t0 = time.time()

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nvml
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((time.time(), mhz, self.max_mhz, {n for n, b in self.BITS.items() if mask & b}))
            except Exception:
                pass
            time.sleep(0.02)

    def _pump(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.proc.stdout:
            f = [x.strip() for x in ln.strip().split(",")]
            if len(f) < 6:
                continue
            try:
                self.samples.append((time.time(), float(f[0]), float(f[1]),
                                     {n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")}))
            except ValueError:
                continue

    def stop(self):
        if self.nvml is None and self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["NVML / nvidia-smi unavailable"]}
        time.sleep(0.03)
        self.stop_flag = True
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        else:
            self.t.join(timeout=1)
        use = [x for x in self.samples if x[0] >= self.t0] or self.samples[-2:]
        sm = [x[1] for x in use]
        mx = [x[2] for x in use]
        reasons = set()
        for x in use:
            reasons |= x[3]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ----------------------------------------------------------------------------------------------------
WORKLOAD = "configs[1]: synthetic 30x nanopore-length reads over 9Mb reference, window centres every %d bp"
MODEL_SEED = 7
# algorithmic FLOPs (2*MAC, unpadded sizes) of the schedule the CUDA path executes -- DESIGN.md "model roofline"
FLOP_FOLD_PER_ROW = 2.0 * 2400 * 1500                 # folded CNN+Linear: K = 3 rows * 200 * 4, N = 3 * 500
FLOP_GRU_STEP_PER_ROW = 2 * (2.0 * 500 * 1500) * 2    # one launch: both directions, input + hidden products
FLOP_GRU_STEP0_PER_ROW = 2 * (2.0 * 500 * 1500)       # layer 0, first step: h_prev == 0, hidden product skipped


def cpu_encode_rate(shard, seconds_budget=12.0, threads=None):
    """Oracle port (oracle/pileup_oracle.c, a C restatement of Pileup.py) on the host cores: windows/s.
    ctypes releases the GIL, so one Python thread per core drives one slice of the sample."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pileup_oracle as po
    po.lib()
    threads = threads or os.cpu_count() or 1
    n0 = min(shard.n_windows, 16)
    t = time.perf_counter()
    po.encode_windows(shard.reads, shard.win_qs[:n0], shard.win_lo[:n0], shard.win_hi[:n0], shard.ref, shard.ref_base)
    per = max((time.perf_counter() - t) / n0, 1e-6)
    n = int(min(shard.n_windows, max(threads * 8, seconds_budget / per * threads)))
    idx = np.array_split(np.arange(n), threads)
    out = [None] * threads

    def work(k, ix):
        if ix.size:
            out[k] = po.encode_windows(shard.reads, shard.win_qs[ix], shard.win_lo[ix], shard.win_hi[ix], shard.ref,
                                       shard.ref_base)[0]
    ths = [threading.Thread(target=work, args=(k, ix)) for k, ix in enumerate(idx)]
    t = time.perf_counter()
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    dt = time.perf_counter() - t
    first = next(o for o in out if o is not None)
    return n / dt, n, threads, dt, first


def cpu_model_rate(images, n_windows=2):
    """Oracle port of model.py + the test.py decode (oracle/model_oracle.py, numpy float32 on the BLAS threads):
    windows/s for whole-image inference (CNN once per image row, the cheapest correct CPU schedule)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import model_oracle as mo
    from deepore_b200 import modelio
    st = modelio.seeded_state(MODEL_SEED)
    n = min(n_windows, images.shape[0])
    try:                                      # torchrun exports OMP_NUM_THREADS=1: give the BLAS every host core back
        from threadpoolctl import threadpool_limits
        limits = threadpool_limits(limits=os.cpu_count() or 1)
    except Exception:
        limits = None
    t = time.perf_counter()
    for i in range(n):
        lg = mo.forward_image(st, images[i], dtype=np.float32)
        mo.vote(mo.argmax_first(lg))
    dt = time.perf_counter() - t
    if limits is not None:
        limits.restore_original_limits()
    return n / dt, n, dt


def cpu_path_rate(shard, stage, enc_budget=8.0, model_windows=2):
    """positions/s of the CPU port for encode (+ model): per-window times of the two stages add up."""
    wps_e, n_e, threads, dt_e, imgs = cpu_encode_rate(shard, seconds_budget=enc_budget)
    sec_per_window = 1.0 / wps_e
    sample = "%d windows encoded by oracle/pileup_oracle.c on %d threads (%.1f s)" % (n_e, threads, dt_e)
    if stage != "encode":
        wps_m, n_m, dt_m = cpu_model_rate(imgs, model_windows)
        sec_per_window += 1.0 / wps_m
        sample += "; %d of them through oracle/model_oracle.py (numpy fp32, BLAS threads, %.1f s)" % (n_m, dt_m)
    return 1.0 / sec_per_window, threads, sample


def run_reference(args, rank, world):
    """--impl reference: CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from deepore_b200.workload import SyntheticWorkload
    wl = SyntheticWorkload(stride=args.stride)
    sh = wl.shards(0, min(args.windows, 2048), 1)[0]
    total = max(1, args.steps + args.warmup)
    enc_budget = max(1.0, min(8.0, 60.0 / total))
    mw = 2 if total <= 6 else 1
    rates, threads, sample = [], 1, ""
    t0 = time.perf_counter()
    for i in range(args.warmup + args.steps):
        wps, threads, sample = cpu_path_rate(sh, args.stage, enc_budget, mw)
        if i >= args.warmup:
            rates.append(wps)
    wps = float(np.mean(rates))
    val = wps * args.stride
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (time.perf_counter() - t0) / total,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8" if args.stage == "encode" else "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD % args.stride, "stage": args.stage,
                       "note": "bounded sample per step; CPU ports of Pileup.py / model.py (the reference is Python and "
                               "needs /root/reference + pysam, absent on the GPU box)"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--stage", default="e2e", choices=["e2e", "encode"],
                    help="e2e = encode + model + vote (the metric); encode = pileup encoder only")
    ap.add_argument("--windows", type=int, default=2048, help="windows per step (interval shard size)")
    ap.add_argument("--model-chunk", type=int, default=512, help="windows per model forward call")
    ap.add_argument("--precision", default="fp16", choices=["bf16", "bf16x3", "fp16", "fp16x3"])
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
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()            # long before the timed region: its start-up must not fall inside it
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # ---- workload: rank r owns a contiguous run of interval shards (weak scaling: fixed work per GPU, no collective)
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
    with_model = args.stage == "e2e"
    mchunk = min(args.model_chunk, nw)
    model = eng = None
    if with_model:
        from deepore_b200 import model as dmodel, modelio
        model = dmodel.from_state(modelio.seeded_state(MODEL_SEED), dev, args.precision)
        eng = model.engine(mchunk * H, dev)
    # two rotating output sets (step i+1 may be encoded while step i's results drain to the host)
    outs = [dict(img=torch.empty((nw, H, W, 4), dtype=torch.uint8, device=dev),
                 rowpos=torch.empty((nw, H), dtype=torch.int32, device=dev),
                 rowins=torch.empty((nw, H), dtype=torch.int32, device=dev),
                 status=torch.empty((nw,), dtype=torch.int32, device=dev),
                 pred=torch.empty((nw * H, 3), dtype=torch.uint8, device=dev),
                 calls=torch.empty((nw, H), dtype=torch.int8, device=dev)) for _ in range(2)]
    host_status = torch.empty((nw,), dtype=torch.int32).pin_memory()
    host_rowpos = torch.empty((nw, H), dtype=torch.int32).pin_memory()
    host_calls = torch.empty((nw, H), dtype=torch.int8).pin_memory()
    import ctypes as C

    def step_device(i):
        sb = bufs[i]
        o = outs[i & 1]
        n = sb.n_windows
        enc.encode(sb, o["img"][:n], o["rowpos"][:n], o["rowins"][:n], o["status"][:n])
        launches = 4                                   # prep_reads, base_codes, encode_windows_v3 (fast), encode_windows (exact, listed windows)
        if with_model:
            rows = o["img"].view(nw * H, W, 4)
            for c0 in range(0, n, mchunk):
                c1 = min(n, c0 + mchunk)
                eng.forward_rows(rows[c0 * H:c1 * H], want_logits=False, want_pred=True, pred=o["pred"][c0 * H:c1 * H])
                launches += 12                         # u8->bf16, folded GEMM, 9 GRU steps, fc1+argmax
            st = torch.cuda.current_stream(dev)
            _lib.check(_lib.lib().dp_vote(o["pred"].data_ptr(), n, H, 3, o["calls"].data_ptr(), None, None,
                                          C.c_void_p(st.cuda_stream)), "dp_vote")
            launches += 1
        return launches

    def step_e2e(i):
        sb = bufs[i]
        sb.upload()
        k = step_device(i)
        o = outs[i & 1]
        n = sb.n_windows
        host_status[:n].copy_(o["status"][:n], non_blocking=True)
        host_rowpos[:n].copy_(o["rowpos"][:n], non_blocking=True)
        if with_model:
            host_calls[:n].copy_(o["calls"][:n], non_blocking=True)
        return k

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, sampler=None, profile=False):
        for b in bufs:
            b.upload()
        for i in range(args.warmup):
            fn(i)
        barrier()
        if profile and eng is not None:
            eng.profile(True)
        if sampler:
            sampler.mark()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches = 0
        enc_evs = []
        e0.record()
        for i in range(args.warmup, nsteps):
            launches += fn(i)
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        prof = None
        if profile and eng is not None:
            prof = eng.profile_read()
            eng.profile(False)
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, clocks, prof

    def time_encoder_alone():
        """Average duration of the encoder launch pair over the timed shards (CUDA events on the launching stream)."""
        evs = []
        for i in range(args.warmup, nsteps):
            sb, o = bufs[i], outs[i & 1]
            n = sb.n_windows
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            enc.encode(sb, o["img"][:n], o["rowpos"][:n], o["rowins"][:n], o["status"][:n])
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize(dev)
        return float(np.mean([a.elapsed_time(b) for a, b in evs]))

    # instrumented pass (CUDA events around every model kernel -> kernel shares and the roofline's launch time), then the
    # pass `value` is taken from: the same K steps without the per-kernel events, clocks sampled during it
    prof = ms_prof = None
    if with_model:
        ms_prof, _, _, prof = timed(step_device, None, profile=True)
    ms_dev, launches, clocks, _ = timed(step_device, sampler, profile=False)
    if ms_prof is None:
        ms_prof = ms_dev
    ms_e2e, _, _, _ = timed(step_e2e)
    enc_ms = time_encoder_alone()

    win_timed = sum(b.n_windows for b in bufs[args.warmup:])
    tot_windows = win_timed * world
    value = tot_windows * args.stride / (ms_dev * 1e-3)
    e2e_value = tot_windows * args.stride / (ms_e2e * 1e-3)

    if rank == 0:
        peaks = load_peaks()
        alg = sum(algorithmic_bytes(sh) for sh in shards[args.warmup:])
        enc_achieved = alg / args.steps / (enc_ms * 1e-3) / 1e9
        enc_roof = {"bound": "hbm", "achieved": enc_achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": enc_achieved / peaks["hbm_gbs"], "peak_source": peaks["source"],
                    "kernel": "prep_reads_kernel+base_codes_kernel+encode_windows_v3_kernel+encode_windows_kernel",
                    "ms_per_launch": enc_ms, "traffic": traffic_of("encoder"),
                    "algorithmic_bytes_per_window": alg / max(1, win_timed)}
        h2d = int(np.mean([b.nbytes for b in bufs[args.warmup:]]))
        d2h = int(nw * 4 + nw * H * 4 + (nw * H if with_model else 0))
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": args.precision if with_model else "u8",
                "data": "synthetic",
                "config": {"workload": WORKLOAD % args.stride, "stage": args.stage, "windows_per_step": nw,
                           "model_chunk_windows": mchunk if with_model else None,
                           "weights": "random init, modelio.seeded_state(%d)" % MODEL_SEED if with_model else None,
                           "l2": "every step reads a new shard and writes %.0f MB of pileup tensors (> L2)"
                                 % (nw * H * W * 4 / 1e6)},
                "windows_per_s": tot_windows / (ms_dev * 1e-3),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": launches, "clocks": clocks}
        if with_model and prof:
            g_ms, g_n = prof["gru_step"]
            rows_per_launch = mchunk * H
            n_layers = 3
            flop_per_launch = rows_per_launch * ((n_layers * 3 - 1) * FLOP_GRU_STEP_PER_ROW + FLOP_GRU_STEP0_PER_ROW) / (n_layers * 3)
            ach = flop_per_launch / (g_ms / max(1, g_n) * 1e-3) / 1e12
            line["roofline"] = {"bound": "tensor", "achieved": ach, "peak": peaks["bf16_tflops_sustained"],
                                "unit": "TFLOP/s", "frac": ach / peaks["bf16_tflops_sustained"], "traffic": traffic_of("gru_step"),
                                "peak_source": peaks["source"] + " (sustained bf16)", "kernel": "gru_step2_kernel",
                                "ms_per_launch": g_ms / max(1, g_n), "launches": g_n,
                                "algorithmic_flop_per_launch": flop_per_launch,
                                "share_of_step": g_ms / ms_dev,
                                "share_of_instrumented_pass": g_ms / ms_prof}
            f_ms, f_n = prof["fold_gemm"]
            fach = rows_per_launch * FLOP_FOLD_PER_ROW / (f_ms / max(1, f_n) * 1e-3) / 1e12
            line["kernels"] = {k: {"ms_total": v[0], "launches": v[1], "share_of_step": v[0] / ms_dev} for k, v in prof.items()}
            line["kernels"]["fold_gemm"]["achieved_tflops"] = fach
            line["kernels"]["encoder"] = {"ms_per_launch_pair": enc_ms, "share_of_step": enc_ms * args.steps / ms_dev}
            line["roofline_encoder"] = enc_roof
        else:
            line["roofline"] = enc_roof
        if not args.no_cpu_baseline and world == 1:            # rank 0 at N=1 only (torchrun pins OMP_NUM_THREADS=1)
            val, threads, sample = cpu_path_rate(shards[args.warmup], args.stage)
            line["cpu_baseline"] = {"value": val * args.stride, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": sample}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
