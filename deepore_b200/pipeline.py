"""File -> calls on the device: BGZF/BAM bytes in host memory in, per-row calls out, nothing in between touches the host.

The reference walks this path one window at a time through pysam (Pileup.py:36,413-415 -> htslib `sam_itr_queryi` /
`hts_itr_next`, dep/htslib/sam.c:669-677, hts.c:2211-2250; block inflate bgzf.c:456-499), writes a PNG per window
(Pileup.py:336-361), reads it back through a DataLoader (test.py:28-38) and runs the model row by row (test.py:59-89).
Here one CONTIG CHUNK (the records of one genomic interval, as a BGZF file image) is the unit:

  side stream     pinned file image --cudaMemcpyAsync--> HBM -> dp_bgzf_inflate -> dp_bam_scan / dp_bam_extract (read arrays in
                  HBM) -> dp_encode_prepare (per-read digests, per-base codes) -> dp_window_ranges (candidate reads per window)
  compute stream  per batch of windows: dp_encode_prepared -> dp_model_forward_windows -> dp_vote; calls + row -> position map of
                  the whole chunk leave the device in ONE copy per array

The ingest of chunk i+1 runs on the side stream while chunk i is on the compute stream (two buffer sets); a host thread
stages the next file image into pinned memory meanwhile.  Region access: a chunk is the whole BGZF image of its interval --
an interval shard of a whole-genome run reads its own slice of the file (by the .bai's linear index, `bai_chunk_range`) and
inflates only that; the per-window `fetch` inside the slice is dp_window_ranges.  There is no CPU fallback anywhere on
this path.
"""
from __future__ import annotations

import ctypes as C
import struct
import threading
from dataclasses import dataclass
from typing import Iterable, Iterator, Optional

import numpy as np

from . import _lib, bamdev


@dataclass
class ContigChunk:
    """One unit of work: a BGZF file image holding the BAM records of `contig` that overlap the windows, the reference
    bytes under them, and the windows as (first queryStart, count, stride)."""
    image: np.ndarray            # uint8, the BGZF bytes (host)
    contig: str
    ref: np.ndarray              # uint8 upper-case reference bases for [ref_base, ref_base + len)
    ref_base: int
    first_qs: int                # queryStart (= centre - flank) of the first window
    n_windows: int
    stride: int = 100
    tag: object = None           # carried through to the result


@dataclass
class ChunkCalls:
    tag: object
    contig: str
    win_qs: np.ndarray           # int32 [n_windows]
    calls: np.ndarray            # int8  [n_windows, H]   (-1: row not scored)
    rowpos: np.ndarray           # int32 [n_windows, H]   (-1: blank row)
    rowins: np.ndarray           # int32 [n_windows, H]
    n_reads: int
    n_bases: int
    fallback_windows: int        # windows the fast encoder kernel handed to the exact kernel (last batch)

    def records(self, window0: int = 0):
        from . import calls as dc
        return dc.call_records(self.calls, self.rowpos, self.rowins, window0)


class _Ingested:
    def release(self):
        """Drop every device buffer of the chunk (its results live on in the pinned host arrays ChunkCalls views)."""
        self.__dict__.clear()


class ContigCaller:
    """Drives contig chunks through ingest -> encode -> model -> vote on one device.  `run(chunks)` yields ChunkCalls in order."""

    def __init__(self, model, device="cuda:0", windows_per_call: int = 4096, model_chunk: int = 512, flank: int = 100,
                 coverageCutoff: int = 200, windowCutoff: int = 150, mapQualityCutoff: int = 20, precision: Optional[str] = None):
        import torch
        _lib.require_cuda()
        self.torch = torch
        self.device = torch.device(device)
        self.model = model
        self.wpc = int(windows_per_call)
        self.mchunk = min(int(model_chunk), self.wpc)
        self.flank, self.W, self.H = flank, coverageCutoff, windowCutoff
        self.params = _lib.EncodeParams(flank, coverageCutoff, windowCutoff, mapQualityCutoff)
        self.eng = model.engine(self.mchunk * self.H, self.device, precision)
        # HIGH priority: the GRU step kernels fill every SM (one 640-thread CTA each, nearly all registers), so the ingest
        # kernels of the next chunk only run in the gaps between two launches -- their blocks must be first in line there, or
        # the compute stream's next launch takes the SMs again and a 20 ms ingest stretches over hundreds of ms
        self.side = torch.cuda.Stream(self.device, priority=-1)
        self.img = torch.empty((self.wpc, self.H, self.W, 4), dtype=torch.uint8, device=self.device)
        self.pred = torch.empty((self.wpc * self.H, 3), dtype=torch.uint8, device=self.device)
        self.timing = {"ingest_ms": 0.0, "compute_ms": 0.0, "chunks": 0}

    # ------------------------------------------------------------------ ingest (side stream)
    def _ingest(self, ch: ContigChunk, pinned=None) -> _Ingested:
        torch = self.torch
        dev = self.device
        L = _lib.lib()
        g = _Ingested()
        g.chunk = ch
        image = np.ascontiguousarray(ch.image, dtype=np.uint8)
        idx = bamdev.index_blocks(image)                          # host walk over the block headers (a few thousand)
        g.ev0 = torch.cuda.Event(enable_timing=True)
        g.ev1 = torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(self.side):
            g.ev0.record(self.side)
            host = pinned if pinned is not None else torch.from_numpy(image).pin_memory()
            bam, _ = bamdev.inflate_on_device(image, dev, idx=idx, stream=self.side, pinned=host)
            d, info = bamdev.records_on_device(bam, ch.contig, stream=self.side)     # one host sync on the side stream (counts)
            g.bam, g.d, g.info = bam, d, info
            n, ncig, nb = info["n_reads"], info["n_cigar"], info["n_bases"]
            g.rstruct = _lib.Reads(n, ncig, d["pos"].data_ptr(), d["flag"].data_ptr(), d["mapq"].data_ptr(),
                                   d["l_seq"].data_ptr(), d["cigar_off"].data_ptr(), d["cigar"].data_ptr(),
                                   d["seq_off"].data_ptr(), d["seq4"].data_ptr(), d["qual_off"].data_ptr(),
                                   d["qual"].data_ptr(), nb)
            need = int(L.dp_encode_workspace_bytes(n, ncig, nb, self.wpc, C.byref(self.params)))
            g.ws = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
            st = C.c_void_p(self.side.cuda_stream)
            _lib.check(L.dp_encode_prepare(C.byref(g.rstruct), C.byref(self.params), self.wpc, g.ws.data_ptr(), g.ws.numel(), st),
                       "dp_encode_prepare")
            nw = int(ch.n_windows)
            g.win_qs = (torch.arange(nw, dtype=torch.int32, device=dev) * int(ch.stride) + int(ch.first_qs)).contiguous()
            g.win_lo = torch.empty(max(nw, 1), dtype=torch.int32, device=dev)
            g.win_hi = torch.empty(max(nw, 1), dtype=torch.int32, device=dev)
            g.unsorted = torch.zeros(1, dtype=torch.int32, device=dev)
            _lib.check(L.dp_window_ranges(C.byref(g.rstruct), g.ws.data_ptr(), g.ws.numel(), self.wpc, nw, g.win_qs.data_ptr(),
                                          self.flank, g.win_lo.data_ptr(), g.win_hi.data_ptr(), g.unsorted.data_ptr(), st),
                       "dp_window_ranges")
            ref = np.ascontiguousarray(ch.ref, dtype=np.uint8)
            g.ref = torch.from_numpy(ref if ref.size else np.zeros(1, np.uint8)).pin_memory().to(dev, non_blocking=True)
            g.ref_len = int(ref.shape[0])
            # read bases overlapping an average window: picks the shared-memory build of the encoder kernel
            span = max(1, nw * int(ch.stride))
            g.bases_per_window = int(nb * (2 * self.flank + 1) / span) if nw else 0
            g.ev1.record(self.side)
            g.ready = torch.cuda.Event()
            g.ready.record(self.side)
        return g

    # ------------------------------------------------------------------ encode + model + vote (compute stream)
    def _compute(self, g: _Ingested):
        torch = self.torch
        dev = self.device
        L = _lib.lib()
        ch = g.chunk
        nw, H = int(ch.n_windows), self.H
        cur = torch.cuda.current_stream(dev)
        cur.wait_event(g.ready)
        g.calls = torch.empty((max(nw, 1), H), dtype=torch.int8, device=dev)
        g.rowpos = torch.empty((max(nw, 1), H), dtype=torch.int32, device=dev)
        g.rowins = torch.empty((max(nw, 1), H), dtype=torch.int32, device=dev)
        g.status = torch.zeros(max(nw, 1), dtype=torch.int32, device=dev)
        g.c0 = torch.cuda.Event(enable_timing=True)
        g.c1 = torch.cuda.Event(enable_timing=True)
        g.c0.record(cur)
        st = C.c_void_p(cur.cuda_stream)
        rows = self.img.view(self.wpc * H, self.W, 4)
        for w0 in range(0, nw, self.wpc):
            n = min(self.wpc, nw - w0)
            _lib.check(L.dp_encode_prepared(
                C.byref(g.rstruct), n, g.win_qs[w0:].data_ptr(), g.win_lo[w0:].data_ptr(), g.win_hi[w0:].data_ptr(),
                g.ref.data_ptr(), int(ch.ref_base), g.ref_len, C.byref(self.params), self.img.data_ptr(),
                g.rowpos[w0:].data_ptr(), g.rowins[w0:].data_ptr(), g.status[w0:].data_ptr(), g.ws.data_ptr(), g.ws.numel(),
                self.wpc, g.bases_per_window, st), "dp_encode_prepared")
            for c0 in range(0, n, self.mchunk):
                c1 = min(n, c0 + self.mchunk)
                self.eng.forward_rows(rows[c0 * H:c1 * H], want_logits=False, want_pred=True, pred=self.pred[c0 * H:c1 * H],
                                      window_rows=H)
            _lib.check(L.dp_vote(self.pred.data_ptr(), n, H, 3, g.calls[w0:].data_ptr(), None, None, st), "dp_vote")
        g.c1.record(cur)
        # results of the whole chunk: one copy per array into pinned host memory
        g.h_calls = torch.empty((max(nw, 1), H), dtype=torch.int8).pin_memory()
        g.h_rowpos = torch.empty((max(nw, 1), H), dtype=torch.int32).pin_memory()
        g.h_rowins = torch.empty((max(nw, 1), H), dtype=torch.int32).pin_memory()
        g.h_status = torch.empty(max(nw, 1), dtype=torch.int32).pin_memory()
        g.h_unsorted = torch.empty(1, dtype=torch.int32).pin_memory()
        g.h_fb = torch.empty(1, dtype=torch.int32).pin_memory()
        g.h_calls.copy_(g.calls, non_blocking=True)
        g.h_rowpos.copy_(g.rowpos, non_blocking=True)
        g.h_rowins.copy_(g.rowins, non_blocking=True)
        g.h_status.copy_(g.status, non_blocking=True)
        g.h_unsorted.copy_(g.unsorted, non_blocking=True)
        g.h_fb.copy_(g.ws[:4].view(torch.int32), non_blocking=True)          # workspace word 0: fallback count of the last batch
        g.done = torch.cuda.Event()
        g.done.record(cur)

    def _collect(self, g: _Ingested) -> ChunkCalls:
        from .pileup import _STATUS_EXC
        g.done.synchronize()
        ch = g.chunk
        nw = int(ch.n_windows)
        if int(g.h_unsorted[0]) != 0:
            raise ValueError("records of %r are not coordinate-sorted (pysam.fetch needs an indexed, sorted BAM)" % ch.contig)
        status = g.h_status.numpy()[:nw]
        bad = np.nonzero(status != 0)[0]
        if bad.size:
            w = int(bad[0])
            raise _STATUS_EXC.get(int(status[w]), RuntimeError)(
                "window at %s:%d raised status %d" % (ch.contig, int(ch.first_qs) + w * int(ch.stride) + self.flank, int(status[w])))
        self.timing["ingest_ms"] += g.ev0.elapsed_time(g.ev1)
        self.timing["compute_ms"] += g.c0.elapsed_time(g.c1)
        self.timing["chunks"] += 1
        win_qs = (np.arange(nw, dtype=np.int64) * int(ch.stride) + int(ch.first_qs)).astype(np.int32)
        return ChunkCalls(tag=ch.tag, contig=ch.contig, win_qs=win_qs, calls=g.h_calls.numpy()[:nw], rowpos=g.h_rowpos.numpy()[:nw],
                          rowins=g.h_rowins.numpy()[:nw], n_reads=g.info["n_reads"], n_bases=g.info["n_bases"],
                          fallback_windows=int(g.h_fb[0]))

    # ------------------------------------------------------------------ driver
    def run(self, chunks: Iterable[ContigChunk]) -> Iterator[ChunkCalls]:
        """Chunks in, ChunkCalls out (same order).  While chunk i is on the compute stream, chunk i+1 is copied, inflated and
        split on the side stream, and a host thread stages chunk i+2's file image into pinned memory."""
        torch = self.torch
        it = iter(chunks)
        staged = {}

        def stage(ch):                                             # host thread: pageable/file-backed bytes -> pinned memory
            if ch is None:
                return
            t = torch.empty(max(int(np.asarray(ch.image).size), 1), dtype=torch.uint8).pin_memory()
            t.numpy()[:ch.image.size] = np.asarray(ch.image, dtype=np.uint8).reshape(-1)
            staged[id(ch)] = t

        def start_stage(ch):
            th = threading.Thread(target=stage, args=(ch,), daemon=True)
            th.start()
            return th

        def ingest_next():
            """Pull the next chunk, ingest it on the side stream, and start staging the one behind it."""
            nonlocal pending, th
            ch = pending
            if ch is None:
                return None
            th.join()
            g = self._ingest(ch, staged.pop(id(ch), None))
            pending = next(it, None)
            th = start_stage(pending)
            return g

        pending = next(it, None)
        if pending is None:
            return
        with torch.cuda.device(self.device):
            th = start_stage(pending)
            cur = ingest_next()
            self._compute(cur)                                     # asynchronous: the compute stream is now busy with `cur`
            nxt = ingest_next()                                    # side stream, overlaps `cur`
            while cur is not None:
                # the compute stream's queue never runs dry: the next chunk's work is enqueued behind `cur` BEFORE the host
                # waits for `cur` and hands it to the consumer, whose own work then overlaps the next chunk's kernels
                if nxt is not None:
                    self._compute(nxt)
                out = self._collect(cur)
                cur.release()
                yield out
                cur = nxt
                nxt = ingest_next() if cur is not None else None   # overlaps `cur`, which is running or queued


def bai_chunk_range(bai: bytes, tid: int, beg: int, end: int):
    """File range [virtual offset of the first BGZF block to read, virtual offset bound) for records of reference `tid`
    overlapping [beg, end), from a .bai index (SAMv1 5.2; the reference gets the same through
    dep/htslib/hts.c:2211-2250 hts_itr_query: bins of reg2bins + the 16 kb linear index as the lower bound).  Returns
    (min_voffset, max_voffset) or None when the index holds nothing for the region.  A shard then reads and inflates only
    compressed bytes [min_voffset >> 16, max_voffset >> 16 plus one block)."""
    if bai[:4] != b"BAI\1":
        raise ValueError("not a .bai index")
    n_ref, = struct.unpack_from("<i", bai, 4)
    off = 8
    for r in range(n_ref):
        n_bin, = struct.unpack_from("<i", bai, off)
        off += 4
        bins = {}
        for _ in range(n_bin):
            b, n_chunk = struct.unpack_from("<Ii", bai, off)
            off += 8
            if r == tid:
                bins[b] = struct.unpack_from("<%dQ" % (2 * n_chunk), bai, off)
            off += 16 * n_chunk
        n_intv, = struct.unpack_from("<i", bai, off)
        off += 4
        if r == tid:
            ioff = struct.unpack_from("<%dQ" % n_intv, bai, off) if n_intv else ()
            lo_bound = ioff[min(beg >> 14, n_intv - 1)] if n_intv else 0
            want = _reg2bins(beg, end)
            lo, hi = None, None
            for b in want:
                ch = bins.get(b)
                if not ch:
                    continue
                for k in range(0, len(ch), 2):
                    cb, ce = ch[k], ch[k + 1]
                    if ce <= lo_bound:
                        continue
                    lo = cb if lo is None else min(lo, cb)
                    hi = ce if hi is None else max(hi, ce)
            if lo is None:
                return None
            return max(lo, lo_bound), hi
        off += 8 * n_intv
    return None


def _reg2bins(beg: int, end: int):
    """Bins that may hold records overlapping [beg, end) (SAMv1 5.3, the UCSC binning scheme htslib uses)."""
    end -= 1
    out = [0]
    for shift, base in ((26, 1), (23, 9), (20, 73), (17, 585), (14, 4681)):
        out.extend(range(base + (beg >> shift), base + (end >> shift) + 1))
    return out
