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
ap.add_argument("--lib", default="", help="load this build of the library instead (scripts/inflate_variants.py)")
ap.add_argument("--sweep", action="store_true", help="only time the inflate kernel over growing prefixes of the block list")
args = ap.parse_args()
if args.lib:
    _lib.LIB_PATH = os.path.abspath(args.lib)

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


if args.sweep:
    rows = []
    for nb in sorted({int(n * f) for f in (0.02, 0.05, 0.1, 0.2, 0.35, 0.5, 0.65, 0.8, 0.9, 1.0)} | {148 * 4 * k for k in (1, 4, 8, 9, 10) if 148 * 4 * k <= n}):
        def part(nb=nb):
            _lib.check(L.dp_bgzf_inflate(d_file.data_ptr(), nb, dd["in_off"].data_ptr(), dd["in_len"].data_ptr(), dd["out_off"].data_ptr(),
                                         dd["out_len"].data_ptr(), out.data_ptr(), status.data_ptr(), C.c_void_p(st.cuda_stream)), "inflate")
        ms = timed(part)
        rows.append({"blocks": nb, "ms": round(ms, 3), "out_GBps": round(float(idx["out_len"][:nb].sum()) / ms / 1e6, 2)})
    print(json.dumps({"sweep": rows}))
    raise SystemExit(0)
ms_inf = timed(inflate)
assert not status.cpu().numpy().any() and out.cpu().numpy().tobytes() == data
ms_rec = timed(lambda: bamdev.records_on_device(out, "chr3"))      # the Python wrapper: allocations, the count read-back, slices
# the two kernels groups by themselves (C-ABI calls, CUDA events on the launching stream)
n_bytes = int(out.numel())
cap = n_bytes // 36 + 1
name = torch.tensor(list(b"chr3"), dtype=torch.uint8, device=dev)
rec_off = torch.empty(cap + 1, dtype=torch.int64, device=dev)
offs = [torch.empty(cap + 1, dtype=torch.int64, device=dev) for _ in range(3)]
counts = torch.zeros(8, dtype=torch.int64, device=dev)


def scan():
    _lib.check(L.dp_bam_scan(out.data_ptr(), n_bytes, name.data_ptr(), 4, cap, rec_off.data_ptr(), offs[0].data_ptr(),
                             offs[1].data_ptr(), offs[2].data_ptr(), counts.data_ptr(), C.c_void_p(st.cuda_stream)), "scan")


ms_scan = timed(scan)
c = counts.cpu().numpy()
assert c[6] == 0 and c[0] == rb.n_reads
nrec, ncig, nseq, nqual = (int(x) for x in c[:4])
dd2 = {"pos": torch.empty(nrec, dtype=torch.int32, device=dev), "flag": torch.empty(nrec, dtype=torch.int16, device=dev),
       "mapq": torch.empty(nrec, dtype=torch.uint8, device=dev), "l_seq": torch.empty(nrec, dtype=torch.int32, device=dev),
       "cigar": torch.empty(max(ncig, 1), dtype=torch.int32, device=dev), "seq4": torch.empty(max(nseq, 1), dtype=torch.uint8, device=dev),
       "qual": torch.empty(max(nqual, 1), dtype=torch.uint8, device=dev)}


def extract():
    _lib.check(L.dp_bam_extract(out.data_ptr(), nrec, rec_off.data_ptr(), offs[0].data_ptr(), offs[1].data_ptr(), offs[2].data_ptr(),
                                dd2["pos"].data_ptr(), dd2["flag"].data_ptr(), dd2["mapq"].data_ptr(), dd2["l_seq"].data_ptr(),
                                dd2["cigar"].data_ptr(), dd2["seq4"].data_ptr(), dd2["qual"].data_ptr(), C.c_void_p(st.cuda_stream)), "extract")


ms_extract = timed(extract)
os.environ["DP_BAM_SCAN"] = "serial"
ms_scan_serial = timed(scan)
del os.environ["DP_BAM_SCAN"]
t0 = time.perf_counter()
for b in range(n):
    zlib.decompress(image[idx["in_off"][b]:idx["in_off"][b] + idx["in_len"][b]].tobytes(), -15)
cpu_s = time.perf_counter() - t0
print(json.dumps({"workload": "synthetic 30x nanopore-length reads over %d kb, BGZF level 6" % args.kb,
                  "compressed_bytes": int(image.size), "inflated_bytes": len(data), "blocks": int(n), "reads": rb.n_reads,
                  "inflate_ms": ms_inf, "inflate_out_GBps": len(data) / ms_inf / 1e6,
                  "inflate_algorithmic_GBps": (len(data) + image.size) / ms_inf / 1e6,
                  "scan_plus_extract_ms": ms_rec, "records_GBps": len(data) / ms_rec / 1e6,
                  "scan_ms": ms_scan, "scan_serial_chain_walk_ms": ms_scan_serial, "extract_ms": ms_extract,
                  "extract_GBps": 2 * len(data) / ms_extract / 1e6,
                  "zlib_one_core_s": cpu_s, "zlib_one_core_GBps": len(data) / cpu_s / 1e9,
                  "note": "warp-per-block inflate: %d blocks = %d warps on 148 SMs" % (n, n)}))
