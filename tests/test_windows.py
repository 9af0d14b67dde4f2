"""Window / label producer (deepore_b200/windows.py) against the unmodified pileupGenerator.py of the reference
(tests/golden/bundled_windows.npz, made by oracle/make_windows_golden.py): record dictionary, window list, label drafts and
insert lengths on the CPU; the batched CUDA driver (PNG + summary CSV) on the GPU."""
import json
import os

import numpy as np
import pytest

import golden_util as gu
from deepore_b200 import windows

VCF = os.path.join(gu.GOLD, "bundled_confident.vcf")
FULL_VCF = "/root/reference/data/vcf_NA12878_S1.small.vcf"


@pytest.fixture()
def gold():
    windows.allVariantRecord.clear()
    yield np.load(os.path.join(gu.GOLD, "bundled_windows.npz"))
    windows.allVariantRecord.clear()


@pytest.mark.parametrize("vcf", [VCF, FULL_VCF, FULL_VCF + ".gz"])
def test_record_dictionary_and_labels_match_reference(gold, vcf):
    if not os.path.isfile(vcf):
        pytest.skip("the full bundled VCF only exists in the build container")
    d = windows.populateRecordDictionary("3", vcf)
    keys = np.array(sorted(d), np.int64)
    assert np.array_equal(keys, gold["record_keys"])
    vals = np.array([[d[k][0], d[k][1] or 0, d[k][2] or 0] for k in keys], np.int32)
    assert np.array_equal(vals, gold["record_vals"])
    recs = windows.qualifying_records("3", vcf)
    assert [r.pos for r in recs] == gold["vcf_pos"].tolist()
    for i, r in enumerate(recs):
        lab, ins, dele = windows.getLabel(r.pos - 100 - 1, r.pos + 100)
        assert lab == str(gold["label_draft"][i])
        assert {str(k): v for k, v in ins.items()} == json.loads(str(gold["insert_lengths"][i]))
        assert {str(k): v for k, v in dele.items()} == json.loads(str(gold["delete_lengths"][i]))


def test_genotype_classes_and_record_fields():
    assert windows.getClassForGenotype(["0", "0"]) == 1
    assert windows.getClassForGenotype([".", "."]) == 1
    assert windows.getClassForGenotype(["0", "1"]) == 2
    assert windows.getClassForGenotype(["1", "2"]) == 2
    assert windows.getClassForGenotype(["1", "1"]) == 3
    r = windows.VcfRecord("chr3\t100\t.\tAT\tA,ATTT\t80.5\tPASS\tDP=3\tGT:DP\t1|2:7\n")
    assert (r.start, r.stop, r.alleles, r.qual) == (99, 101, ("AT", "A", "ATTT"), 80.5)
    assert windows.getGTField(r) == ["1", "2"]
    assert windows.getGTField(windows.VcfRecord("chr3\t5\t.\tA\tC\t.\t.\t.\tGT\t0\\1\n")) == ["0", "1"]
    b = windows.VcfRecord("chr3\t99914\t.\tA\t.\t.\tPASS\tEND=100026;BLOCKAVG_min30p3a\tGT:DP\t0/0:50\n")
    assert (b.qual, b.alleles, b.start, b.stop) == (None, ("A",), 99913, 100026)


def test_insert_and_delete_runs(tmp_path):
    """An insertion allele marks its anchor with the allele length; a deletion spans len(REF) positions and opens one run."""
    windows.allVariantRecord.clear()
    p = tmp_path / "t.vcf"
    p.write_text("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\tS\n"
                 "chr3\t1010\t.\tA\tACGT\t99\tPASS\t.\tGT\t0/1\n"
                 "chr3\t1020\t.\tGTTT\tG\t99\tPASS\t.\tGT\t1/1\n"
                 "chr3\t1030\t.\tC\tT\t12\tPASS\t.\tGT\t1/1\n")
    try:
        d = windows.populateRecordDictionary("3", str(p))
        assert d[1009] == (2, 4, None)
        assert [d[k] for k in range(1019, 1023)] == [(3, None, 1)] * 4          # longest CALLED allele is "G"
        assert 1029 not in d                                                      # QUAL 12
        lab, ins, dele = windows.getLabel(1000, 1030)
        assert lab == "1" * 9 + "2" + "1" * 9 + "3333" + "1" * 7
        assert ins == {9: 4} and dele == {19: 1, 20: 1, 21: 1, 22: 1}
        assert [r.pos for r in windows.qualifying_records("3", str(p))] == [1010, 1020]
        assert [r.pos for r in windows.read_vcf(str(p), "chr3:1020-1025")] == [1020]
    finally:
        windows.allVariantRecord.clear()


@pytest.mark.gpu
def test_batched_driver_writes_the_reference_files(gold, tmp_path):
    """generatePileupBasedonVCF on the bundled data (BAM and FASTA rebuilt from the committed arrays): the summary CSV lines, the
    file names and every PNG pixel equal what the unmodified reference driver wrote."""
    from PIL import Image
    g = gu.load("bundled_na12878")
    bam, fa = str(tmp_path / "b.bam"), str(tmp_path / "r.fa")
    gu.write_bam(bam, g.reads)
    gu.write_fasta(fa, g.ref, g.ref_base)
    out = str(tmp_path) + "/"
    n = windows.generatePileupBasedonVCF("3", bam, fa, VCF, out, 100)
    assert n == 260
    lines = [ln.rstrip("\n").split(",") for ln in open(out + "summary-3.csv")]
    assert [os.path.basename(p) for p, _ in lines] == [str(x) for x in gold["png_name"]]
    assert [lab for _, lab in lines] == [str(x) for x in gold["label_out"]]
    for i, (p, _) in enumerate(lines):
        assert os.path.isabs(p)
        assert np.array_equal(np.array(Image.open(p)), g.img[i]), "window %d" % i


@pytest.mark.gpu
def test_labelled_batch_any_order(gold):
    """Window centres in arbitrary order come back in the caller's order, labels included."""
    g = gu.load("bundled_na12878")
    pick = [200, 3, 77, 3, 259, 0]
    pos = [int(gold["vcf_pos"][k]) - 1 for k in pick]
    drafts = [str(gold["label_draft"][k]) for k in pick]
    ins = [{int(a): b for a, b in json.loads(str(gold["insert_lengths"][k])).items()} for k in pick]
    img, labels, rowpos, rowins = windows.encode_labelled_batch(g.reads, g.ref, g.ref_base, pos, drafts, ins)
    for a, k in enumerate(pick):
        assert np.array_equal(img[a], g.img[k])
        assert labels[a] == str(gold["label_out"][k])


@pytest.mark.gpu
def test_device_png_files_decode_to_the_same_pixels(tmp_path):
    """dp_png_encode (csrc/png.cu): the byte streams written on the device are valid PNG files -- PIL (what
    modules.dataset.PileupDataset uses, dataset.py:47-52) reads back exactly the encoder's pixels, for the default image
    shape and for one whose scanline stream does not fill a deflate block; sizes match dp_png_size."""
    import io
    import zlib
    import torch
    from PIL import Image
    from deepore_b200 import _lib, png
    g = gu.load("bundled_na12878")
    pick = [0, 1, 57, 131, 259]
    files = png.encode_batch(torch.from_numpy(g.img[pick]).cuda())
    assert tuple(files.shape) == (len(pick), int(_lib.lib().dp_png_size(150, 200)))
    for a, k in enumerate(pick):
        raw = files[a].numpy().tobytes()
        im = Image.open(io.BytesIO(raw))
        im.load()
        assert im.mode == "RGBA" and im.size == (200, 150)
        assert np.array_equal(np.asarray(im), g.img[k])
        # the IDAT payload is a well-formed zlib stream (Adler-32 checked by zlib) of filter-0 scanlines
        n = int.from_bytes(raw[33:37], "big")
        assert raw[37:41] == b"IDAT" and zlib.crc32(raw[37:41 + n]) == int.from_bytes(raw[41 + n:45 + n], "big")
        lines = zlib.decompress(raw[41:41 + n])
        assert len(lines) == 150 * 801 and all(lines[r * 801] == 0 for r in range(150))
    rng = np.random.default_rng(4)
    small = rng.integers(0, 256, size=(3, 7, 12, 4), dtype=np.uint8)
    files = png.encode_batch(torch.from_numpy(small).cuda())
    for a in range(3):
        assert np.array_equal(np.asarray(Image.open(io.BytesIO(files[a].numpy().tobytes()))), small[a])
    paths = [str(tmp_path / ("w%d.png" % i)) for i in range(3)]
    assert png.write_files(paths, torch.from_numpy(small).cuda()) == 3 * files.shape[1]
    assert np.array_equal(np.asarray(Image.open(paths[2])), small[2])


@pytest.mark.gpu
def test_labels_of_windows_with_more_than_256_rows_match_the_reference():
    """Long-read windows with 3 % insertions have 270 .. 340 image rows in total; the image keeps 150 of them but the label
    is widened by EVERY insert row of the window (Pileup.labelInsertRegion, Pileup.py:309,320-333).  The uncut row layout
    comes from dp_window_layout; labels must equal what the unmodified reference returned (tests/golden/nanopore_small.npz:
    label draft '1' * 201, no variant lengths -- oracle/ref_harness.run_reference_window)."""
    g = gu.load("nanopore_small")
    n = g.win_qs.shape[0]
    assert max(len(str(x)) for x in g.labels) > 256
    pos = [int(q) + 100 for q in g.win_qs]
    img, labels, rowpos, rowins = windows.encode_labelled_batch(g.reads, g.ref, g.ref_base, pos, ["1" * 201] * n, [{}] * n)
    assert np.array_equal(img, g.img)
    assert labels == [str(x) for x in g.labels]
    # the Pileup class walks the same layout
    from deepore_b200.pileup import rowins_of_layout, window_layout
    colrows, status = window_layout(g.reads, g.win_qs[:3], g.ref, g.ref_base)
    assert (status == 0).all() and colrows.shape == (3, 201)
    for k in range(3):
        full = rowins_of_layout(colrows[k])
        assert len(full) - int((colrows[k] > 0).sum()) == len(str(g.labels[k])) - 201      # one label character per insert row
        assert np.array_equal(full[:150], rowins[k][:150])
