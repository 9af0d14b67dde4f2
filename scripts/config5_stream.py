#!/usr/bin/env python
"""BASELINE.json configs[4] (SURVEY 8d "config 5"), bounded: a whole genome streamed as 9 Mb contig chunks of the configs[1]
generator through the file -> calls pipeline (deepore_b200/pipeline.py), chunks sharded by interval across the GPUs of
one box -- rank r takes chunks r, r + N, ... -- with no data-path collective; the calls are gathered on the host.

Each rank pushes `--chunks` chunks (default 20 = 180 Mb of genome per GPU; the full 3.1 Gb genome is 345 chunks) through
ContigCaller.run: BGZF bytes in host memory -> pinned staging by a producer thread -> H2D + inflate + record split + window
index on a side stream, overlapped with encode -> model -> vote of the previous chunk.  Generating 30x reads for 9 Mb costs
~30 s of host time per chunk, far more than the GPU needs for it (~1.1 s), so each rank builds `--distinct` chunk images
with distinct seeds BEFORE the timed region and cycles through them; every pass re-does all device work.

    python scripts/config5_stream.py [--chunks 20] [--distinct 2] [--mb 9]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/config5_stream.py
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepore_b200 import model as dmodel, modelio, pipeline, sharding, synth          # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chunks", type=int, default=20)
    ap.add_argument("--distinct", type=int, default=2)
    ap.add_argument("--mb", type=float, default=9.0)
    ap.add_argument("--stride", type=int, default=100)
    ap.add_argument("--precision", default="fp16")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        host_group = dist.new_group(backend="gloo")
    L = int(args.mb * 1e6)
    t0 = time.perf_counter()
    images = []
    for k in range(args.distinct):
        seed = 1000 + 17 * (rank * args.distinct + k)                     # distinct genome interval = distinct seeds
        ref = synth.random_reference(L, seed=seed)
        rb = synth.simulate_reads(ref, coverage=30.0, seed=seed + 1)
        stream = synth.bam_stream(rb, "chr3", L)
        img = synth.bgzf_image(stream, level=6, threads=max(1, (os.cpu_count() or 8) // max(1, world)))
        wq = synth.window_starts(0, L, stride=args.stride)
        images.append((img, ref, int(wq[0]), int(wq.shape[0]), len(stream), rb.n_reads))
    prep_s = time.perf_counter() - t0
    m = dmodel.from_state(modelio.seeded_state(7), dev, args.precision)
    caller = pipeline.ContigCaller(m, dev, windows_per_call=4096, model_chunk=512, precision=args.precision)

    def chunk(i):                                                         # global chunk index of this rank's i-th chunk
        img, ref, q0, nw, _, _ = images[i % args.distinct]
        return pipeline.ContigChunk(img, "chr3", ref, 0, q0, nw, args.stride, tag=rank + i * world)

    list(caller.run([chunk(0)]))                                          # warm-up
    caller.timing = {"ingest_ms": 0.0, "compute_ms": 0.0, "chunks": 0}
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    n_windows, n_records, checks = 0, 0, []
    for out in caller.run(chunk(i) for i in range(args.chunks)):
        n_windows += int(out.calls.shape[0])
        scored = (out.calls >= 0) & (out.rowpos >= 0)
        n_records += int(scored.sum())
        checks.append(int(out.calls[scored].astype(np.int64).sum()))      # checksum of the call set of the chunk
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    gather_ms = 0.0
    per_rank = np.asarray([[n_windows, n_records, sum(checks)]], dtype=np.int64)
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        g0 = time.perf_counter()
        per_rank = sharding.gather_on_host(per_rank, group=host_group, dst=0)
        gather_ms = (time.perf_counter() - g0) * 1e3
    if rank == 0:
        tot_w = int(per_rank[:, 0].sum())
        comp = sum(int(images[i % args.distinct][0].size) for i in range(args.chunks))
        infl = sum(int(images[i % args.distinct][4]) for i in range(args.chunks))
        print(json.dumps({
            "workload": "BASELINE configs[4], bounded: %d chunks of %.0f Mb (30x nanopore-length reads) per GPU, %d distinct chunk "
                        "images per rank cycled" % (args.chunks, args.mb, args.distinct),
            "n_gpus": world, "precision": args.precision, "positions_per_s": tot_w * args.stride / dt,
            "windows": tot_w, "genome_mb_per_gpu": args.chunks * args.mb, "seconds": dt,
            "call_records": int(per_rank[:, 1].sum()), "call_checksum": int(per_rank[:, 2].sum()),
            "compressed_bytes_per_gpu": comp, "inflated_bytes_per_gpu": infl,
            "ingest_ms_per_chunk_side_stream": caller.timing["ingest_ms"] / max(1, caller.timing["chunks"]),
            "compute_ms_per_chunk": caller.timing["compute_ms"] / max(1, caller.timing["chunks"]),
            "host_gather_ms": gather_ms, "host_prep_s_untimed": prep_s,
            "timer": "host clock around ContigCaller.run over all chunks, device idle on both sides, max over ranks"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
