"""File -> calls pipeline (deepore_b200/pipeline.py): the .bai region lookup on the CPU (against the bundled index of the
reference's data), and on the GPU the whole chain BGZF bytes -> dp_bgzf_inflate -> dp_bam_scan / dp_bam_extract ->
dp_encode_prepare -> dp_window_ranges -> dp_encode_prepared -> model -> dp_vote against the host-reader path
(reads.ReadBatch + numpy window index + dp_encode_windows), array for array."""
import os

import numpy as np
import pytest

import golden_util as gu

REF_BAM = "/root/reference/data/aln_NA12878.pg.chr3_100k-200k.bam"


@pytest.mark.skipif(not os.path.isfile(REF_BAM + ".bai"), reason="needs /root/reference (build container)")
def test_bai_chunk_range_covers_every_overlapping_record():
    """The range the bundled .bai gives for a region contains the virtual offset of every record overlapping it
    (dep/htslib/hts.c:2211-2250 hts_itr_query); virtual offsets recomputed here from the BGZF block table and the record chain."""
    import gzip
    import struct
    from deepore_b200 import pipeline
    from deepore_b200.reads import BamFile
    image = np.fromfile(REF_BAM, dtype=np.uint8)
    bai = open(REF_BAM + ".bai", "rb").read()
    # block table on the host: compressed offset and inflated offset of every BGZF block
    coff, uoff, o, u = [], [], 0, 0
    raw = image.tobytes()
    while o < len(raw):
        bsize = struct.unpack_from("<H", raw, o + 16)[0] + 1
        coff.append(o); uoff.append(u)
        u += struct.unpack_from("<I", raw, o + bsize - 4)[0]
        o += bsize
    coff, uoff = np.array(coff), np.array(uoff)
    stream = gzip.decompress(raw)
    l_text, = struct.unpack_from("<i", stream, 4)
    p = 8 + l_text
    n_ref, = struct.unpack_from("<i", stream, p)
    p += 4
    names = []
    for _ in range(n_ref):
        ln, = struct.unpack_from("<i", stream, p)
        names.append(stream[p + 4:p + 4 + ln - 1].decode())
        p += 8 + ln
    tid = names.index("chr3")
    rb = BamFile(REF_BAM).read_contig("chr3")
    voffs = []
    while p < len(stream):
        bs, = struct.unpack_from("<i", stream, p)
        b = int(np.searchsorted(uoff, p, side="right") - 1)
        voffs.append((int(coff[b]) << 16) | (p - int(uoff[b])))
        p += 4 + bs
    voffs = np.array(voffs)
    assert voffs.shape[0] == rb.n_reads
    for beg, end in [(100_000, 100_201), (150_000, 152_000), (199_800, 200_050), (99_000, 99_950)]:
        hit = (rb.pos < end) & (rb.endpos > beg)
        r = pipeline.bai_chunk_range(bai, tid, beg, end)
        if not hit.any():
            continue
        assert r is not None
        lo, hi = r
        assert voffs[hit].min() >= lo and voffs[hit].max() < hi, (beg, end)
        assert (lo >> 16) in set(coff.tolist())
    assert pipeline.bai_chunk_range(bai, tid + 1, 0, 1000) is None


@pytest.mark.gpu
@pytest.mark.parametrize("fixture", ["bundled_na12878", "nanopore_small", "quirks"])
def test_window_ranges_on_device_equal_the_host_index(fixture):
    import ctypes as C
    import torch
    from deepore_b200 import _lib
    from deepore_b200.pileup import ShardBuffers
    from deepore_b200.workload import IntervalShard
    g = gu.load(fixture)
    rb = g.reads
    L = _lib.lib()
    lo_pos, hi_pos = int(rb.pos.min()) - 300, int(rb.endpos.max()) + 300
    win_qs = np.arange(lo_pos, hi_pos, 37, dtype=np.int32)
    want_lo, want_hi = rb.window_ranges(win_qs, 100)
    sh = IntervalShard(rb, g.ref, g.ref_base, win_qs, want_lo, want_hi, 100)
    sb = ShardBuffers(sh, "cuda").upload()
    p = _lib.EncodeParams(100, 200, 150, 20)
    need = int(L.dp_encode_workspace_bytes(rb.n_reads, rb.n_cigar, int(rb.qual.shape[0]), 64, C.byref(p)))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(L.dp_encode_prepare(C.byref(sb.rstruct), C.byref(p), 64, ws.data_ptr(), ws.numel(), st), "prepare")
    lo = torch.empty(win_qs.shape[0], dtype=torch.int32, device="cuda")
    hi = torch.empty_like(lo)
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    _lib.check(L.dp_window_ranges(C.byref(sb.rstruct), ws.data_ptr(), ws.numel(), 64, win_qs.shape[0], sb.ptr["win_qs"], 100,
                                  lo.data_ptr(), hi.data_ptr(), flag.data_ptr(), st), "ranges")
    torch.cuda.synchronize()
    assert int(flag.item()) == 0
    assert np.array_equal(lo.cpu().numpy(), want_lo) and np.array_equal(hi.cpu().numpy(), want_hi)


def _image(rb, **kw):
    return np.frombuffer(gu.bgzf_compress(gu.bam_bytes(rb, **kw), level=6, chunk=0xff00), dtype=np.uint8)


@pytest.mark.gpu
def test_bam_bytes_to_calls_equal_the_host_reader_path():
    """Two chunks of tiled windows over the bundled NA12878 reads through ContigCaller (double-buffered ingest, window batches
    of 96, model chunks of 32) == host ReadBatch -> dp_encode_windows -> model -> dp_vote for the same windows."""
    import torch
    from deepore_b200 import model as dmodel, modelio, pipeline
    from deepore_b200.pileup import encode_windows
    g = gu.load("bundled_na12878")
    image = _image(g.reads, other_contig_every=9, names=("r", "read/2"))
    m = dmodel.from_state(modelio.seeded_state(7), "cuda")
    caller = pipeline.ContigCaller(m, "cuda:0", windows_per_call=96, model_chunk=32)
    chunks = [pipeline.ContigChunk(image, "chr3", g.ref, g.ref_base, 100_050 + 20_000 * k, 200, 100, tag=k) for k in range(2)]
    out = list(caller.run(chunks))
    assert [o.tag for o in out] == [0, 1] and caller.timing["chunks"] == 2
    for k, o in enumerate(out):
        assert np.array_equal(o.win_qs, 100_050 + 20_000 * k + 100 * np.arange(200))
        img, rowpos, rowins, status = encode_windows(g.reads, o.win_qs, g.ref, g.ref_base)
        assert (status.cpu().numpy() == 0).all()
        _, pred = m.forward_windows(img, want_logits=False)
        votes, _ = dmodel.vote(pred.view(-1, 3), 200, 150, 3)
        torch.cuda.synchronize()
        assert np.array_equal(o.rowpos, rowpos.cpu().numpy()) and np.array_equal(o.rowins, rowins.cpu().numpy())
        assert np.array_equal(o.calls, votes.cpu().numpy())
        assert o.n_reads == g.reads.n_reads
        rec = o.records(window0=200 * k)
        assert rec.shape[0] == int(((o.calls >= 0) & (o.rowpos >= 0)).sum()) > 0
    assert list(caller.run([])) == []


@pytest.mark.gpu
def test_unsorted_records_are_refused():
    from deepore_b200 import model as dmodel, modelio, pipeline
    from deepore_b200.reads import ReadBatch
    g = gu.load("quirks")
    rb = g.reads
    order = np.arange(rb.n_reads)[::-1]
    recs = [rb.slice(int(i), int(i) + 1) for i in order]
    rev = ReadBatch(contig=rb.contig, pos=np.concatenate([r.pos for r in recs]), flag=np.concatenate([r.flag for r in recs]),
                    mapq=np.concatenate([r.mapq for r in recs]), l_seq=np.concatenate([r.l_seq for r in recs]),
                    cigar_off=np.concatenate([[0], np.cumsum([r.n_cigar for r in recs])]).astype(np.int64),
                    cigar=np.concatenate([r.cigar for r in recs]),
                    seq_off=np.concatenate([[0], np.cumsum([r.seq4.shape[0] for r in recs])]).astype(np.int64),
                    seq4=np.concatenate([r.seq4 for r in recs]),
                    qual_off=np.concatenate([[0], np.cumsum([r.qual.shape[0] for r in recs])]).astype(np.int64),
                    qual=np.concatenate([r.qual for r in recs]))
    if not np.any(np.diff(rev.pos) < 0):
        pytest.skip("fixture has a single position")
    m = dmodel.from_state(modelio.seeded_state(7), "cuda")
    caller = pipeline.ContigCaller(m, "cuda:0", windows_per_call=64, model_chunk=32)
    ch = pipeline.ContigChunk(_image(rev), "chr3", g.ref, g.ref_base, int(g.win_qs.min()), 4, 100)
    with pytest.raises(ValueError):
        list(caller.run([ch]))
