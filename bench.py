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


def traffic_of(kernel, units):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the named kernel from the committed `ncu --set full`
    captures (profiles/r2_traffic.json, else profiles/r1_traffic.json), for the launch size (`units` = windows per encoder
    call / image rows per GRU launch) this run uses.  -> (bytes per launch or None, name of the capture or None)."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            d = json.load(open(os.path.join(ROOT, "profiles", name)))[kernel]
        except Exception:
            continue
        by = d.get("by_launch_size")
        if by is not None:
            e = by.get(str(int(units)))
            if e is None:
                continue
            return e["dram_bytes_per_launch"], e.get("source", name)
        if d.get("launch_size") == int(units):             # a capture that does not name its launch size is not used
            return d["dram_bytes_per_launch"], d.get("source", name)
    return None, None


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


def _all_blas_threads():
    try:                                      # torchrun exports OMP_NUM_THREADS=1: give the BLAS every host core back
        from threadpoolctl import threadpool_limits
        return threadpool_limits(limits=os.cpu_count() or 1)
    except Exception:
        return None


def cpu_model_rate(images, n_windows=2):
    """Oracle port of model.py + the test.py decode (oracle/model_oracle.py, numpy float32 on the BLAS threads):
    windows/s for whole-image inference with the DE-DUPLICATED schedule (CNN once per image row, the cheapest correct CPU
    schedule -- the reference itself runs the CNN three times per row, see cpu_torch_rate)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import model_oracle as mo
    from deepore_b200 import modelio
    st = modelio.seeded_state(MODEL_SEED)
    n = min(n_windows, images.shape[0])
    limits = _all_blas_threads()
    t = time.perf_counter()
    for i in range(n):
        lg = mo.forward_image(st, images[i], dtype=np.float32)
        mo.vote(mo.argmax_first(lg))
    dt = time.perf_counter() - t
    if limits is not None:
        limits.restore_original_limits()
    return n / dt, n, dt


_TORCH_MODEL = None


def cpu_torch_rate(images, seconds_budget=10.0, call_batch=32, min_calls=4):
    """What the reference's own test.py executes, on the host cores: nn.Conv2d / nn.BatchNorm2d / nn.Linear / nn.GRU modules
    filled from the same seeded state (oracle/torch_model.py, forward restated from model.py:93-156), torch CPU fp32,
    torch.set_num_threads(all cores), call batch 32, the reference's NAIVE schedule (every forward call runs the
    CNN on its three rows, test.py:59-66).  Runs forward calls for `seconds_budget` seconds (at least `min_calls`) and
    returns windows/s = call_batch / (148 * seconds per call)."""
    global _TORCH_MODEL
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch
    import torch_model as tm
    from deepore_b200 import modelio
    torch.set_num_threads(os.cpu_count() or 1)
    if _TORCH_MODEL is None:
        _TORCH_MODEL = tm.build(modelio.seeded_state(MODEL_SEED))
    m = _TORCH_MODEL
    reps = -(-call_batch // images.shape[0])
    batch = np.concatenate([images] * reps, axis=0)[:call_batch]
    x = (torch.from_numpy(np.ascontiguousarray(batch)).permute(0, 3, 1, 2).float() / 255.0).contiguous()
    hidden = torch.zeros(2, call_batch, m.hidden_size)
    calls, row = 0, 0
    with torch.no_grad():
        tm.forward(m, x[:, :, 0:3, :], hidden)                 # untimed: thread pool / primitive cache warm-up
        t = time.perf_counter()
        while calls < min_calls or time.perf_counter() - t < seconds_budget:
            tm.forward(m, x[:, :, row:row + 3, :], hidden)
            row = (row + 1) % 148
            calls += 1
        dt = time.perf_counter() - t
    return call_batch / (148.0 * dt / calls), calls, dt, torch.get_num_threads()


def cpu_path_rate(shard, stage, enc_budget=8.0, model_windows=2, torch_budget=10.0):
    """positions/s of the CPU implementation for encode (+ model): per-window times of the two stages add up.  The model
    stage is timed twice -- the numpy port with the de-duplicated schedule and the torch-CPU modules with the reference's
    naive schedule -- and the FASTER of the two enters the value.  -> (windows/s, threads, sample text, legs dict)."""
    wps_e, n_e, threads, dt_e, imgs = cpu_encode_rate(shard, seconds_budget=enc_budget)
    sec_per_window = 1.0 / wps_e
    sample = "%d windows encoded by oracle/pileup_oracle.c on %d threads (%.1f s)" % (n_e, threads, dt_e)
    legs = {"encoder_c_port_windows_per_s": wps_e}
    if stage != "encode":
        wps_m, n_m, dt_m = cpu_model_rate(imgs, model_windows)
        wps_t, n_t, dt_t, nth = cpu_torch_rate(imgs, torch_budget)
        legs.update({"model_numpy_port_windows_per_s": wps_m, "model_numpy_port_schedule": "de-duplicated (CNN once per row)",
                     "model_torch_cpu_windows_per_s": wps_t, "model_torch_cpu_schedule": "naive 3-row calls, call batch 32",
                     "model_torch_cpu_threads": nth, "model_leg_used": "numpy_port" if wps_m >= wps_t else "torch_cpu"})
        sec_per_window += 1.0 / max(wps_m, wps_t)
        sample += ("; model: %d windows through oracle/model_oracle.py (numpy fp32, de-duplicated schedule, %.1f s) = %.3g "
                   "windows/s; %d forward calls of batch 32 through oracle/torch_model.py (torch CPU fp32 nn modules, the "
                   "reference's naive schedule, %d threads, %.1f s) = %.3g windows/s; faster leg used"
                   % (n_m, dt_m, wps_m, n_t, nth, dt_t, wps_t))
    return 1.0 / sec_per_window, threads, sample, legs


def run_reference(args, rank, world):
    """--impl reference: CPU implementation of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from deepore_b200.workload import SyntheticWorkload
    wl = SyntheticWorkload(stride=args.stride)
    sh = wl.shards(0, min(args.windows, 2048), 1)[0]
    total = max(1, args.steps + args.warmup)
    enc_budget = max(1.0, min(8.0, 50.0 / total))
    torch_budget = max(1.5, min(8.0, 50.0 / total))
    mw = 2 if total <= 6 else 1
    rates, threads, sample, legs = [], 1, "", {}
    t0 = time.perf_counter()
    for i in range(args.warmup + args.steps):
        wps, threads, sample, legs = cpu_path_rate(sh, args.stage, enc_budget, mw, torch_budget)
        if i >= args.warmup:
            rates.append(wps)
    wps = float(np.mean(rates))
    val = wps * args.stride
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * (time.perf_counter() - t0) / total,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8" if args.stage == "encode" else "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD % args.stride, "stage": args.stage,
                       "note": "bounded sample per step; CPU ports of Pileup.py (C) and model.py (numpy port and torch-CPU "
                               "nn modules); the reference itself is Python + pysam + /root/reference, absent on the GPU box"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample, "legs": legs},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def bam_path(dev, model, precision, rank, stride, n_chunks=3, windows_per_chunk=8192, level=6):
    """positions/s from COMPRESSED BAM BYTES IN HOST MEMORY to calls in host memory (deepore_b200/pipeline.py): per chunk one
    H2D copy of the BGZF image on a side stream, dp_bgzf_inflate, dp_bam_scan / dp_bam_extract, dp_encode_prepare,
    dp_window_ranges, then batches of dp_encode_prepared -> model -> dp_vote, and one D2H copy per result array.  The ingest
    of chunk i+1 overlaps the compute of chunk i.  The first chunk is the warm-up; the rest are timed with the host clock
    (the path spans two streams and a host thread), device idle on both sides."""
    import torch
    from deepore_b200 import pipeline, synth
    from deepore_b200.workload import SyntheticWorkload
    wl = SyntheticWorkload(stride=stride)
    chunks, comp, raw = [], 0, 0
    for k in range(n_chunks):
        w0 = ((rank * n_chunks + k) * windows_per_chunk) % max(1, wl.n_windows - windows_per_chunk)
        wq = wl.all_win_qs[w0:w0 + windows_per_chunk]
        region = (int(wq[0]), int(wq[-1]) + 2 * wl.flank + 2)
        rb = wl.synth.simulate_reads(wl.ref, coverage=wl.coverage, seed=wl.read_seed, region=region)
        stream = synth.bam_stream(rb, "chr3", wl.ref_len)
        image = synth.bgzf_image(stream, level=level, threads=os.cpu_count() or 1)
        lo, hi = max(0, region[0] - 1), min(wl.ref_len, region[1] + 1)
        chunks.append(pipeline.ContigChunk(image, "chr3", wl.ref[lo:hi], lo, int(wq[0]), int(wq.shape[0]), stride, tag=k))
        if k > 0:
            comp += int(image.size)
            raw += len(stream)
    caller = pipeline.ContigCaller(model, dev, windows_per_call=4096, model_chunk=512, precision=precision)
    list(caller.run(chunks[:1]))                                   # warm-up (allocations, kernel attributes)
    caller.timing = {"ingest_ms": 0.0, "compute_ms": 0.0, "chunks": 0}
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    outs = list(caller.run(chunks[1:]))
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    nwin = sum(int(o.calls.shape[0]) for o in outs)
    ok = all(((o.calls[:, 2:148] >= 0) & (o.calls[:, 2:148] < 3)).all() for o in outs)
    return {"seconds": dt, "windows": nwin, "chunks": len(outs), "compressed_bytes": comp, "inflated_bytes": raw,
            "ingest_ms_on_side_stream": caller.timing["ingest_ms"], "compute_ms_on_compute_stream": caller.timing["compute_ms"],
            "fallback_windows": int(sum(o.fallback_windows for o in outs)), "calls_in_range": bool(ok)}


def parity_block(dev, precision, trained=False):
    """Parity of the precision this run is timed in, on the committed goldens (outputs of the UNMODIFIED reference model over
    the 260 bundled NA12878 pileup images; no oracle/ import): logits max-abs / rms, top-1 classes and votes that differ,
    counted over all of them.  trained: the golden of the checkpoint that took SGD steps on these windows
    (oracle/make_trained_golden.py; seeded state + the tensors the steps changed, stored in the fixture).  Outside every
    timed region."""
    import torch
    from deepore_b200 import model as dmodel, modelio
    name = "bundled_calls_trained.npz" if trained else "bundled_calls_seed7.npz"
    path = os.path.join(ROOT, "tests", "golden", name)
    if not os.path.isfile(path):
        return None
    gc = np.load(path)
    state = dict(modelio.seeded_state(int(gc["seed"])))
    for k in gc.files:
        if k.startswith("state/"):
            state[k[len("state/"):]] = gc[k]
    img = torch.from_numpy(np.load(os.path.join(ROOT, "tests", "golden", "bundled_na12878.npz"))["img"]).to(dev)
    m = dmodel.from_state(state, dev, precision)
    logits, pred = m.forward_windows(img)
    nw, H = int(img.shape[0]), int(img.shape[1])
    votes, _ = dmodel.vote(pred.view(-1, 3), nw, H, 3)
    torch.cuda.synchronize(dev)
    lg = logits.cpu().numpy()[:, :148]
    pr = pred.cpu().numpy()[:, :148]
    v = votes.cpu().numpy()[:, 2:148]
    d = lg - gc["logits"]
    what = ("after %d SGD steps of train.py's loop on these windows" % int(gc["steps"])) if trained else "seeded weights"
    out = {"precision": precision, "golden": "tests/golden/%s (unmodified reference model, torch CPU fp32, %s, "
           "260 bundled NA12878 windows x 148 calls)" % (name, what), "logits_max_abs": float(np.abs(d).max()),
           "logits_rms": float(np.sqrt((d ** 2).mean())), "logit_scale": float(np.abs(gc["logits"]).max()),
           "n_argmax": int(pr.size), "n_argmax_diff": int((pr != gc["pred"]).sum()),
           "n_votes": int(v.size), "n_vote_diff": int((v != gc["calls"][:, 2:148]).sum())}
    del m
    torch.cuda.empty_cache()
    return out


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
    ap.add_argument("--no-x3", action="store_true", help="skip the extra pass in the split (x3) precision")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity block (golden windows through the model)")
    ap.add_argument("--no-bam", action="store_true", help="skip the e2e_from_bam measurement (file image -> calls pipeline)")
    ap.add_argument("--bam-chunks", type=int, default=4, help="timed contig chunks of the e2e_from_bam measurement")
    ap.add_argument("--bam-windows", type=int, default=8192, help="windows per contig chunk of the e2e_from_bam measurement")
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

    # ---- workload: rank r owns a contiguous run of interval shards (weak scaling: fixed work per GPU, no data-path
    # collective; deepore_b200/sharding.py cuts the window axis, results are gathered on the host)
    from deepore_b200 import sharding
    nsteps = args.steps + args.warmup
    wl = SyntheticWorkload(stride=args.stride)
    first, last = sharding.rank_range(world * nsteps * args.windows, world, rank)
    assert last - first == nsteps * args.windows
    shards = wl.shards(first, args.windows, nsteps)
    while len(shards) < nsteps:                                   # wrapped past the contig end
        shards += wl.shards(0, args.windows, nsteps - len(shards))
    bufs = [ShardBuffers(sh, dev) for sh in shards]
    enc = Encoder(device=dev)
    nw = args.windows
    H, W = enc.H, enc.W
    with_model = args.stage == "e2e"
    mchunk = min(args.model_chunk, nw)
    engines = {}
    if with_model:
        from deepore_b200 import model as dmodel, modelio
        state = modelio.seeded_state(MODEL_SEED)

        def engine_for(precision):
            if precision not in engines:
                m = dmodel.from_state(state, dev, precision)
                engines[precision] = (m, m.engine(mchunk * H, dev))
            return engines[precision][1]
        engine_for(args.precision)
    # two rotating output sets (step i+1 may be encoded while step i's results drain to the host)
    outs = [dict(img=torch.empty((nw, H, W, 4), dtype=torch.uint8, device=dev),
                 rowpos=torch.empty((nw, H), dtype=torch.int32, device=dev),
                 rowins=torch.empty((nw, H), dtype=torch.int32, device=dev),
                 status=torch.empty((nw,), dtype=torch.int32, device=dev),
                 pred=torch.empty((nw * H, 3), dtype=torch.uint8, device=dev),
                 calls=torch.empty((nw, H), dtype=torch.int8, device=dev)) for _ in range(2)]
    host_status = torch.zeros((nsteps, nw), dtype=torch.int32).pin_memory()
    host_rowpos = torch.empty((nw, H), dtype=torch.int32).pin_memory()
    host_calls = torch.full((nsteps, nw, H), -1, dtype=torch.int8).pin_memory()
    host_group = dist.new_group(backend="gloo") if world > 1 else None      # host path of the result gather
    import ctypes as C

    def make_steps(eng):
        def step_device(i):
            sb = bufs[i]
            o = outs[i & 1]
            n = sb.n_windows
            enc.encode(sb, o["img"][:n], o["rowpos"][:n], o["rowins"][:n], o["status"][:n])
            launches = 4                               # prep_reads, base_codes, encode_windows_v3 (fast), encode_windows (exact, listed windows)
            if with_model:
                rows = o["img"].view(nw * H, W, 4)
                for c0 in range(0, n, mchunk):
                    c1 = min(n, c0 + mchunk)
                    eng.forward_rows(rows[c0 * H:c1 * H], want_logits=False, want_pred=True, pred=o["pred"][c0 * H:c1 * H], window_rows=H)
                    launches += 12                     # u8->16-bit, folded GEMM, 9 GRU steps, fc1+argmax
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
            host_status[i, :n].copy_(o["status"][:n], non_blocking=True)
            host_rowpos[:n].copy_(o["rowpos"][:n], non_blocking=True)
            if with_model:
                host_calls[i, :n].copy_(o["calls"][:n], non_blocking=True)
            return k
        return step_device, step_e2e

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, sampler=None, profile_eng=None):
        for b in bufs:
            b.upload()
        for i in range(args.warmup):
            fn(i)
        barrier()
        if profile_eng is not None:
            profile_eng.profile(True)
        if sampler:
            sampler.mark()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches = 0
        e0.record()
        for i in range(args.warmup, nsteps):
            launches += fn(i)
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        prof = None
        if profile_eng is not None:
            prof = profile_eng.profile_read()
            profile_eng.profile(False)
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, clocks, prof

    def gather_results():
        """Host-side gather of the timed steps' call records on rank 0 (interval order = rank order), timed with the host
        clock, max over ranks: the one piece of the multi-GPU path that is not device work."""
        if world == 1:
            return 0.0, None
        local = host_calls[args.warmup:].numpy().reshape(-1, H).copy()      # later passes reuse the pinned buffer
        dist.barrier(group=host_group)
        t0 = time.perf_counter()
        allc = sharding.gather_on_host(local, group=host_group, dst=0)
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX, group=host_group)
        return float(dt.item()) * 1e3, (allc, local)

    def time_encoder_alone():
        """Average duration of the encoder launch group over the timed shards (CUDA events on the launching stream)."""
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
    eng = engines[args.precision][1] if with_model else None
    step_device, step_e2e = make_steps(eng)
    prof = ms_prof = None
    if with_model:
        ms_prof, _, _, prof = timed(step_device, None, profile_eng=eng)
    ms_dev, launches, clocks, _ = timed(step_device, sampler)
    if ms_prof is None:
        ms_prof = ms_dev
    ms_e2e, _, _, _ = timed(step_e2e)
    gather_ms, gathered = gather_results()
    fallbacks = enc.fallback_count()                               # of the last timed shard (synchronises)
    st_host = host_status[args.warmup:].numpy()
    assert (st_host == 0).all(), "a timed shard returned a non-zero window status: %s" % np.unique(st_host)
    if with_model:
        hc = host_calls[args.warmup:].numpy()
        for k, b in enumerate(bufs[args.warmup:]):
            assert ((hc[k, :b.n_windows, 2:148] >= 0) & (hc[k, :b.n_windows, 2:148] < 3)).all(), "calls out of range"
        assert (hc[:, :, :2] == -1).all()
    enc_ms = time_encoder_alone()
    # the same path in the parity-grade precision (operands split hi + lo, three passes), fewer words: value / e2e only
    x3 = None
    if with_model and not args.precision.endswith("x3") and not args.no_x3:
        px3 = args.precision + "x3"
        sd3, se3 = make_steps(engine_for(px3))
        ms3, _, _, _ = timed(sd3)
        ms3e, _, _, _ = timed(se3)
        x3 = (px3, ms3, ms3e)

    bam = None
    if with_model and not args.no_bam:
        bam = bam_path(dev, engines[args.precision][0], args.precision, rank, args.stride, n_chunks=1 + args.bam_chunks,
                       windows_per_chunk=args.bam_windows)
        if world > 1:
            t = torch.tensor([bam["seconds"]], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            bam["seconds"] = float(t.item())

    win_timed = sum(b.n_windows for b in bufs[args.warmup:])
    tot_windows = win_timed * world
    value = tot_windows * args.stride / (ms_dev * 1e-3)
    e2e_value = tot_windows * args.stride / ((ms_e2e + gather_ms) * 1e-3)

    if rank == 0:
        if gathered is not None:
            allc, local = gathered
            assert allc.shape == (tot_windows, H) and np.array_equal(allc[:win_timed], local), "host gather lost rank 0's rows"
        peaks = load_peaks()
        alg = sum(algorithmic_bytes(sh) for sh in shards[args.warmup:])
        enc_achieved = alg / args.steps / (enc_ms * 1e-3) / 1e9
        enc_traffic, enc_tsrc = traffic_of("encoder", nw)
        enc_roof = {"bound": "hbm", "achieved": enc_achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": enc_achieved / peaks["hbm_gbs"], "peak_source": peaks["source"],
                    "kernel": "prep_reads_kernel+base_codes_kernel+encode_windows_v3_kernel+encode_windows_kernel",
                    "ms_per_launch": enc_ms, "traffic": enc_traffic, "traffic_source": enc_tsrc,
                    "algorithmic_bytes_per_window": alg / max(1, win_timed), "windows_per_launch": nw,
                    "fallback_windows_last_shard": fallbacks}
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
                                 % (nw * H * W * 4 / 1e6),
                           "sharding": "deepore_b200.sharding.rank_range over %d windows" % (world * nsteps * nw)},
                "windows_per_s": tot_windows / (ms_dev * 1e-3),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "host_gather_ms": gather_ms},
                "gpu_launches": launches, "clocks": clocks, "status_checked": "all timed windows returned status 0; calls in range"}
        if x3 is not None:
            line["value_" + x3[0]] = tot_windows * args.stride / (x3[1] * 1e-3)
            line["e2e_" + x3[0]] = tot_windows * args.stride / ((x3[2] + gather_ms) * 1e-3)
        if with_model and prof:
            g_ms, g_n = prof["gru_step"]
            n_layers = 3
            # the calls test.py makes: 148 of the 150 rows of a window (no 3-row window starts at rows 148, 149;
            # dp_model_forward_windows drops them behind the folded GEMM, so a GRU launch works on exactly these rows);
            # FLOPs at unpadded sizes, hidden product skipped where h_prev == 0
            calls_per_launch = mchunk * (H - 2)
            rows_per_launch = calls_per_launch
            flop_per_launch = calls_per_launch * ((n_layers * 3 - 1) * FLOP_GRU_STEP_PER_ROW + FLOP_GRU_STEP0_PER_ROW) / (n_layers * 3)
            ach = flop_per_launch / (g_ms / max(1, g_n) * 1e-3) / 1e12
            g_traffic, g_tsrc = traffic_of("gru_step", rows_per_launch)
            line["roofline"] = {"bound": "tensor", "achieved": ach, "peak": peaks["bf16_tflops_sustained"],
                                "unit": "TFLOP/s", "frac": ach / peaks["bf16_tflops_sustained"], "traffic": g_traffic,
                                "traffic_source": g_tsrc,
                                "peak_source": peaks["source"] + " (sustained bf16; fp16 runs at the same tensor-pipe rate)",
                                "kernel": "gru_step2_kernel",
                                "ms_per_launch": g_ms / max(1, g_n), "launches": g_n,
                                "algorithmic_flop_per_launch": flop_per_launch, "rows_per_launch": rows_per_launch,
                                "calls_per_launch": calls_per_launch,
                                "share_of_step": g_ms / ms_dev,
                                "share_of_instrumented_pass": g_ms / ms_prof}
            f_ms, f_n = prof["fold_gemm"]
            fach = calls_per_launch * FLOP_FOLD_PER_ROW / (f_ms / max(1, f_n) * 1e-3) / 1e12
            line["kernels"] = {k: {"ms_total": v[0], "launches": v[1], "share_of_step": v[0] / ms_dev} for k, v in prof.items()}
            line["kernels"]["fold_gemm"]["achieved_tflops"] = fach
            line["kernels"]["encoder"] = {"ms_per_launch_pair": enc_ms, "share_of_step": enc_ms * args.steps / ms_dev}
            line["roofline_encoder"] = enc_roof
        else:
            line["roofline"] = enc_roof
        if bam is not None:
            line["e2e_from_bam"] = {
                "value": bam["windows"] * world * args.stride / bam["seconds"], "unit": UNIT,
                "what": "compressed BAM bytes in host memory -> H2D on a side stream -> inflate, record split, window index, encode, "
                        "model, vote on the device -> calls + row map in host memory; ingest of chunk i+1 overlaps compute of chunk i",
                "timer": "host clock around the timed chunks, device idle on both sides, max over ranks",
                "chunks": bam["chunks"], "windows_per_chunk": bam["windows"] // max(1, bam["chunks"]),
                "h2d_compressed_bytes_per_chunk": bam["compressed_bytes"] // max(1, bam["chunks"]),
                "inflated_bytes_per_chunk": bam["inflated_bytes"] // max(1, bam["chunks"]),
                "ingest_share_of_device_time": bam["ingest_ms_on_side_stream"] / max(1e-9, bam["ingest_ms_on_side_stream"] + bam["compute_ms_on_compute_stream"]),
                "ingest_ms_per_chunk": bam["ingest_ms_on_side_stream"] / max(1, bam["chunks"]),
                "compute_ms_per_chunk": bam["compute_ms_on_compute_stream"] / max(1, bam["chunks"]),
                "fallback_windows": bam["fallback_windows"], "calls_in_range": bam["calls_in_range"]}
        if with_model and not args.no_parity:
            line["parity"] = parity_block(dev, args.precision)
            pt = parity_block(dev, args.precision, trained=True)
            if pt is not None:
                line["parity_trained_checkpoint"] = pt
            if x3 is not None:
                line["parity_" + x3[0]] = parity_block(dev, x3[0])
        if not args.no_cpu_baseline and world == 1:            # rank 0 at N=1 only (torchrun pins OMP_NUM_THREADS=1)
            sh0 = shards[args.warmup]
            val, threads, sample, legs = cpu_path_rate(sh0, args.stage)
            line["cpu_baseline"] = {"value": val * args.stride, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": sample, "legs": legs}
            # the sample the C port just encoded, against what the CUDA encoder wrote for the same windows of that shard
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import pileup_oracle as po
            ncheck = min(64, sh0.n_windows)
            want = po.encode_windows(sh0.reads, sh0.win_qs[:ncheck], sh0.win_lo[:ncheck], sh0.win_hi[:ncheck], sh0.ref,
                                     sh0.ref_base)[0]
            o = outs[args.warmup & 1]
            bufs[args.warmup].upload()
            enc.encode(bufs[args.warmup], o["img"], o["rowpos"], o["rowins"], o["status"])
            torch.cuda.synchronize(dev)
            same = bool(np.array_equal(o["img"][:ncheck].cpu().numpy(), want))
            assert same, "pileup tensors of the first timed shard differ from the CPU port"
            line["output_checked"] = "%d windows of the first timed shard byte-identical to the CPU port" % ncheck
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
