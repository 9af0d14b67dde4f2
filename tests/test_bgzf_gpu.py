"""Device-side BAM ingest through the C-ABI: dp_bgzf_inflate against zlib (the library the reference's
htslib calls, dep/htslib/bgzf.c:456-499), dp_bam_scan / dp_bam_extract against the host reader reads.BamFile, and the
ingested batch through the encoder against the committed reference images."""
import zlib

import numpy as np
import pytest

import golden_util as gu

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _bam_stream(name="nanopore_small", **kw):
    return gu.bam_bytes(gu.load(name).reads, **kw)


@pytest.mark.parametrize("level,strategy", [(0, 0), (1, 0), (6, 0), (9, 0), (6, zlib.Z_FIXED), (6, zlib.Z_HUFFMAN_ONLY)])
def test_inflate_bit_exact_vs_zlib(level, strategy):
    from deepore_b200 import bamdev
    data = _bam_stream()
    img = np.frombuffer(gu.bgzf_compress(data, level=level, strategy=strategy), dtype=np.uint8)
    out, status = bamdev.inflate_on_device(img)
    assert status.shape[0] == (len(data) + 0xff00 - 1) // 0xff00 + 1 and not status.any()
    assert out.cpu().numpy().tobytes() == data


def test_inflate_ragged_blocks_and_noise():
    from deepore_b200 import bamdev
    rng = np.random.default_rng(17)
    noise = bytes(rng.integers(0, 256, 300_000, dtype=np.uint8))
    runs = b"\0" * 200_000 + b"ACGT" * 50_000
    for data, chunk in ((noise, 0xf000), (runs, 0xff00), (runs + noise, 777), (b"", 0xff00), (b"x", 1)):
        img = np.frombuffer(gu.bgzf_compress(data, level=6, chunk=chunk), dtype=np.uint8)
        out, status = bamdev.inflate_on_device(img)
        assert not status.any() and out.cpu().numpy().tobytes() == data


def test_inflate_reports_corrupt_blocks():
    from deepore_b200 import bamdev
    data = _bam_stream("quirks")
    img = np.frombuffer(gu.bgzf_compress(data, level=6), dtype=np.uint8).copy()
    idx = bamdev.index_blocks(img)
    rng = np.random.default_rng(2)
    for _ in range(20):                                   # flipped payload bits: an error or different bytes, no fault
        bad = img.copy()
        b = int(rng.integers(0, idx["n_blocks"] - 1))
        for _k in range(4):
            bad[int(idx["in_off"][b]) + int(rng.integers(0, idx["in_len"][b]))] ^= 1 << int(rng.integers(0, 8))
        try:
            out, status = bamdev.inflate_on_device(bad, idx=idx)
            assert out.numel() == len(data)
        except ValueError as e:
            assert "does not inflate" in str(e)
    torch.cuda.synchronize()
    out, status = bamdev.inflate_on_device(img, idx=idx)  # the device is still healthy
    assert out.cpu().numpy().tobytes() == data


@pytest.mark.parametrize("name", ["quirks", "nanopore_small", "bundled_na12878"])
def test_records_match_host_reader(name, tmp_path):
    from deepore_b200 import bamdev
    from deepore_b200.reads import BamFile
    g = gu.load(name)
    data = gu.bam_bytes(g.reads, other_contig_every=3, names=("r", "read/2", "a-much-longer-read-name-0123456789"))
    p = str(tmp_path / "x.bam")
    with open(p, "wb") as fh:
        fh.write(gu.bgzf_compress(data, level=6))
    want = BamFile(p).read_contig("chr3")
    d, info = bamdev.read_contig_device(p, "chr3")
    got = bamdev.to_read_batch(d, "chr3")
    assert info["n_reads"] == want.n_reads == g.reads.n_reads and info["tid"] == 0
    assert info["records_seen"] == g.reads.n_reads + (g.reads.n_reads + 2) // 3
    for k in ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4", "qual_off", "qual"):
        assert np.array_equal(getattr(got, k), getattr(want, k)), k
        assert np.array_equal(getattr(got, k), getattr(g.reads, k)), k
    other, oinfo = bamdev.read_contig_device(p, "chrOther")
    assert oinfo["tid"] == 1 and oinfo["n_reads"] == (g.reads.n_reads + 2) // 3
    every, einfo = bamdev.read_contig_device(p, "")
    assert einfo["n_reads"] == einfo["records_seen"] == info["records_seen"]
    with pytest.raises(ValueError):
        bamdev.read_contig_device(p, "chr4")


def test_ingested_reads_encode_to_the_reference_images(tmp_path):
    from deepore_b200 import bamdev
    from deepore_b200.pileup import encode_windows
    g = gu.load("bundled_na12878")
    p = str(tmp_path / "b.bam")
    with open(p, "wb") as fh:
        fh.write(gu.bgzf_compress(gu.bam_bytes(g.reads), level=6))
    d, _ = bamdev.read_contig_device(p, "chr3")
    rb = bamdev.to_read_batch(d, "chr3")
    img, _, _, status = encode_windows(rb, g.win_qs, g.ref, g.ref_base)
    assert np.array_equal(status.cpu().numpy(), g.status)
    ok = g.status == 0
    assert np.array_equal(img.cpu().numpy()[ok], g.img[ok])


@pytest.mark.parametrize("what", ["negative_l_seq", "huge_n_cigar", "zero_l_name", "negative_l_text", "negative_n_ref"])
def test_corrupt_record_fields_are_refused(what):
    """The consistency rule of bam_read1 (dep/htslib/sam.c:397-406, which returns -4): l_seq < 0, l_read_name < 1 or
    variable-length parts larger than block_size must end the scan with an error -- never with offsets that walk
    backwards or reads past the record (the extract kernel trusts the scan)."""
    import struct
    from deepore_b200 import bamdev
    data = bytearray(_bam_stream("quirks"))
    l_text, = struct.unpack_from("<i", data, 4)
    p = 8 + l_text
    n_ref, = struct.unpack_from("<i", data, p)
    q = p + 4
    for _ in range(n_ref):
        ln, = struct.unpack_from("<i", data, q)
        q += 8 + ln
    bs, = struct.unpack_from("<i", data, q)
    q2 = q + 4 + bs                                               # second record: corrupt one in the middle of the chain
    if what == "negative_l_seq":
        struct.pack_into("<i", data, q2 + 20, -40)
    elif what == "huge_n_cigar":
        struct.pack_into("<H", data, q2 + 16, 65535)
    elif what == "zero_l_name":
        data[q2 + 12] = 0
    elif what == "negative_l_text":
        struct.pack_into("<i", data, 4, -8)
    else:
        struct.pack_into("<i", data, p, -1)
    img = np.frombuffer(gu.bgzf_compress(bytes(data), level=1), dtype=np.uint8)
    bam, _ = bamdev.inflate_on_device(img)
    with pytest.raises(ValueError):
        bamdev.records_on_device(bam, "chr3")
    torch.cuda.synchronize()                                      # no sticky CUDA error left behind
    ok, _ = bamdev.inflate_on_device(np.frombuffer(gu.bgzf_compress(_bam_stream("quirks"), level=1), dtype=np.uint8))
    d, info = bamdev.records_on_device(ok, "chr3")
    assert info["n_reads"] == gu.load("quirks").reads.n_reads
