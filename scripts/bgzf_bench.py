"""Measurement of the device-side BAM ingest: BGZF inflate, record scan, SoA extract, on a synthetic
configs[1]-like BAM (30x nanopore-length reads), CUDA events on the launching stream; zlib on one host core beside it.
    python scripts/bgzf_bench.py [--mb 64] > gpurun_out/bgzf_bench.json"""
import argparse
import ctypes as C
import json
import os
import sys
import time
import zlib

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import golden_util as gu  # noqa: E402
from deepore_b200 import _lib, bamdev, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--kb", type=int, default=1500, help="reference length in kb (30x coverage of it is simulated)")
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()

ref = synth.random_reference(args.kb * 1000, seed=1)
rb = synth.simulate_reads(ref, coverage=30, seed=2)
data = gu.bam_bytes(rb)
image = np.frombuffer(gu.bgzf_compress(data, level=6), dtype=np.uint8)
idx = bamdev.index_blocks(image)
n = idx["n_blocks"]
dev = torch.device("cuda:0")
L = _lib.lib()
d_file = torch.from_numpy(image.copy()).to(dev)
dd = {k: torch.from_numpy(idx[k]).to(dev) for k in ("in_off", "in_len", "out_off", "out_len")}
out = torch.empty(len(data), dtype=torch.uint8, device=dev)
status = torch.zeros(n, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream(dev)


def inflate():
    _lib.check(L.dp_bgzf_inflate(d_file.data_ptr(), n, dd["in_off"].data_ptr(), dd["in_len"].data_ptr(), dd["out_off"].data_ptr(),
                                 dd["out_len"].data_ptr(), out.data_ptr(), status.data_ptr(), C.c_void_p(st.cuda_stream)), "inflate")


def timed(fn):
    fn(); torch.cuda.synchronize()
    ms = []
    for _ in range(args.reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms))


ms_inf = timed(inflate)
assert not status.cpu().numpy().any() and out.cpu().numpy().tobytes() == data
ms_rec = timed(lambda: bamdev.records_on_device(out, "chr3"))
t0 = time.perf_counter()
for b in range(n):
    zlib.decompress(image[idx["in_off"][b]:idx["in_off"][b] + idx["in_len"][b]].tobytes(), -15)
cpu_s = time.perf_counter() - t0
print(json.dumps({"workload": "synthetic 30x nanopore-length reads over %d kb, BGZF level 6" % args.kb,
                  "compressed_bytes": int(image.size), "inflated_bytes": len(data), "blocks": int(n), "reads": rb.n_reads,
                  "inflate_ms": ms_inf, "inflate_out_GBps": len(data) / ms_inf / 1e6,
                  "inflate_algorithmic_GBps": (len(data) + image.size) / ms_inf / 1e6,
                  "scan_plus_extract_ms": ms_rec, "records_GBps": len(data) / ms_rec / 1e6,
                  "zlib_one_core_s": cpu_s, "zlib_one_core_GBps": len(data) / cpu_s / 1e9,
                  "note": "warp-per-block inflate: %d blocks = %d warps on 148 SMs" % (n, n)}))
