"""Device-side BAM ingest: the compressed BGZF file image crosses PCIe once; inflate, record split and
the structure-of-arrays read batch (`dp_reads`) are produced in HBM by csrc/bgzf.cu.

Replaces, for whole-contig reads, the host stage `reads.BamFile` (and, in the reference, pysam -> htslib:
dep/htslib/bgzf.c:456-499 inflate_block, dep/htslib/sam.c bam_read1 behind Pileup.py:36,414).  Same result as
`reads.BamFile(path).read_contig(contig)`, array for array.  No CPU fallback: raises without the CUDA library / a GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np

from . import _lib
from .reads import ReadBatch

INFLATE_ERRORS = {-1: "input truncated", -2: "more than ISIZE bytes", -3: "reserved block type", -4: "stored LEN/NLEN mismatch",
                  -5: "bad Huffman table", -6: "bad symbol", -7: "distance before start of block", -8: "size differs from ISIZE"}


def index_blocks(image: np.ndarray) -> Dict[str, np.ndarray]:
    """Host walk over the BGZF block headers of a file image (uint8 array): payload offsets/lengths, inflated
    offsets/lengths, trailer CRC32s (dp_bgzf_index)."""
    L = _lib.lib()
    image = np.ascontiguousarray(image, dtype=np.uint8)
    n = int(L.dp_bgzf_index(image.ctypes.data, image.size, 0, None, None, None, None, None))
    if n < 0:
        _lib.check(int(n), "dp_bgzf_index")
    idx = {"in_off": np.zeros(max(n, 1), np.int64), "in_len": np.zeros(max(n, 1), np.int32),
           "out_off": np.zeros(n + 1, np.int64), "out_len": np.zeros(max(n, 1), np.int32),
           "crc32": np.zeros(max(n, 1), np.uint32)}
    if n > 0:
        m = int(L.dp_bgzf_index(image.ctypes.data, image.size, n, idx["in_off"].ctypes.data, idx["in_len"].ctypes.data,
                                idx["out_off"].ctypes.data, idx["out_len"].ctypes.data, idx["crc32"].ctypes.data))
        if m < 0:
            _lib.check(m, "dp_bgzf_index")
    idx["n_blocks"] = n
    return idx


def inflate_on_device(image: np.ndarray, device="cuda:0", idx: Optional[dict] = None, stream=None, pinned=None):
    """BGZF file image (host uint8) -> (inflated bytes as a device uint8 tensor, per-block status on the host).
    `pinned`: the same bytes already staged in a pinned torch tensor (the pipeline's producer thread does that)."""
    import torch
    _lib.require_cuda()
    image = np.ascontiguousarray(image, dtype=np.uint8)
    idx = index_blocks(image) if idx is None else idx
    n = int(idx["n_blocks"])
    total = int(idx["out_off"][n])
    dev = torch.device(device)
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    with torch.cuda.stream(st):
        if pinned is None:
            pinned = torch.from_numpy(image if image.size else np.zeros(1, np.uint8)).pin_memory()
        d_file = pinned.to(dev, non_blocking=True)
        d = {k: torch.from_numpy(idx[k]).to(dev, non_blocking=True) for k in ("in_off", "in_len", "out_off", "out_len")}
        out = torch.empty(max(total, 1), dtype=torch.uint8, device=dev)
        status = torch.full((max(n, 1),), 1, dtype=torch.int32, device=dev)
        _lib.check(_lib.lib().dp_bgzf_inflate(d_file.data_ptr(), n, d["in_off"].data_ptr(), d["in_len"].data_ptr(),
                                              d["out_off"].data_ptr(), d["out_len"].data_ptr(), out.data_ptr(),
                                              status.data_ptr(), C.c_void_p(st.cuda_stream)), "dp_bgzf_inflate")
        hstatus = status[:n].cpu().numpy()
    bad = np.nonzero(hstatus != 0)[0]
    if bad.size:
        b = int(bad[0])
        raise ValueError("BGZF block %d (of %d) does not inflate: %s" % (b, n, INFLATE_ERRORS.get(int(hstatus[b]), hstatus[b])))
    return out[:total], hstatus


def records_on_device(bam, contig: str = "", stream=None) -> Tuple[dict, dict]:
    """Inflated BAM stream (device uint8 tensor) -> (dict of device tensors named like ReadBatch's fields, counts)."""
    import torch
    _lib.require_cuda()
    dev = bam.device
    L = _lib.lib()
    st = stream if stream is not None else torch.cuda.current_stream(dev)
    n_bytes = int(bam.numel())
    cap = n_bytes // 36 + 1                                        # a record is at least 36 bytes
    with torch.cuda.stream(st):
        name = torch.tensor(list(contig.encode("ascii")) or [0], dtype=torch.uint8).to(dev)
        rec_off = torch.empty(cap + 1, dtype=torch.int64, device=dev)
        offs = {k: torch.empty(cap + 1, dtype=torch.int64, device=dev) for k in ("cigar_off", "seq_off", "qual_off")}
        counts = torch.zeros(8, dtype=torch.int64, device=dev)
        _lib.check(L.dp_bam_scan(bam.data_ptr(), n_bytes, name.data_ptr(), len(contig), cap, rec_off.data_ptr(),
                                 offs["cigar_off"].data_ptr(), offs["seq_off"].data_ptr(), offs["qual_off"].data_ptr(),
                                 counts.data_ptr(), C.c_void_p(st.cuda_stream)), "dp_bam_scan")
        c = counts.cpu().numpy()
        if c[6] == -3:
            raise ValueError("contig %r is not in the BAM header" % contig)      # list.index() in reads.BamFile
        if c[6] != 0:
            raise ValueError("BAM stream is malformed (%s)" % {-1: "bad magic", -2: "truncated", -4: "record or header field "
                             "inconsistent with its length (bam_read1 returns -4, dep/htslib/sam.c:403-406)"}.get(int(c[6]), int(c[6])))
        n, ncig, nseq, nqual = (int(x) for x in c[:4])
        d = {"pos": torch.empty(max(n, 1), dtype=torch.int32, device=dev),
             "flag": torch.empty(max(n, 1), dtype=torch.int16, device=dev),
             "mapq": torch.empty(max(n, 1), dtype=torch.uint8, device=dev),
             "l_seq": torch.empty(max(n, 1), dtype=torch.int32, device=dev),
             "cigar": torch.empty(max(ncig, 1), dtype=torch.int32, device=dev),
             "seq4": torch.empty(max(nseq, 1), dtype=torch.uint8, device=dev),
             "qual": torch.empty(max(nqual, 1), dtype=torch.uint8, device=dev)}
        _lib.check(L.dp_bam_extract(bam.data_ptr(), n, rec_off.data_ptr(), offs["cigar_off"].data_ptr(),
                                    offs["seq_off"].data_ptr(), offs["qual_off"].data_ptr(), d["pos"].data_ptr(),
                                    d["flag"].data_ptr(), d["mapq"].data_ptr(), d["l_seq"].data_ptr(), d["cigar"].data_ptr(),
                                    d["seq4"].data_ptr(), d["qual"].data_ptr(), C.c_void_p(st.cuda_stream)), "dp_bam_extract")
        for k in ("pos", "flag", "mapq", "l_seq"):
            d[k] = d[k][:n]
        d["cigar"], d["seq4"], d["qual"] = d["cigar"][:ncig], d["seq4"][:nseq], d["qual"][:nqual]
        for k, v in offs.items():
            d[k] = v[:n + 1].clone()
    info = {"n_reads": n, "n_cigar": ncig, "n_bases": nqual, "tid": int(c[4]), "records_seen": int(c[5])}
    return d, info


def read_contig_device(path: str, contig: str, device="cuda:0"):
    """BAM file -> device-resident read arrays of one contig (file order), via inflate + scan + extract on the GPU."""
    image = np.fromfile(path, dtype=np.uint8)
    bam, _ = inflate_on_device(image, device)
    return records_on_device(bam, contig)


def to_read_batch(d: dict, contig: str) -> ReadBatch:
    """Device arrays -> host ReadBatch (for checks and for callers that want the host-side window index)."""
    h = {k: v.cpu().numpy() for k, v in d.items()}
    return ReadBatch(contig=contig, pos=h["pos"], flag=h["flag"].view(np.uint16), mapq=h["mapq"], l_seq=h["l_seq"],
                     cigar_off=h["cigar_off"], cigar=h["cigar"].view(np.uint32), seq_off=h["seq_off"], seq4=h["seq4"],
                     qual_off=h["qual_off"], qual=h["qual"])
