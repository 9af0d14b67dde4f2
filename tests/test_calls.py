"""Call-set records (deepore_b200/calls.py): votes + row map -> (position, insert index, class) records, merge rule for
overlapping windows, TSV round trip.  CPU only; the GPU side of the same artefact is in test_model_gpu.py."""
import numpy as np

from deepore_b200 import calls as dc


def _toy():
    # two windows of H = 8 rows, window centres 100 bp apart (flank 100): the second one starts 100 bp further right
    H = 8
    calls = np.full((2, H), -1, np.int8)
    rowpos = np.full((2, H), -1, np.int32)
    rowins = np.zeros((2, H), np.int32)
    # window 0 (qs = 1000): rows 0..6 populated, an insertion of two bases after position 1002; row 7 blank
    rowpos[0, :7] = [1000, 1001, 1002, 1002, 1002, 1003, 1004]
    rowins[0, :7] = [0, 0, 0, 1, 2, 0, 0]
    calls[0, 2:7] = [1, 2, 2, 0, 1]
    # window 1 (qs = 1100) sees positions 1100.. -- plus, for the merge rule, pretend it also scored 1003 and 1004
    rowpos[1, :6] = [1003, 1004, 1100, 1101, 1102, 1103]
    calls[1, 2:6] = [0, 1, 1, 2]
    calls[1, 0:2] = [2, 2]
    return calls, rowpos, rowins, np.array([1000, 1100], np.int32)


def test_records_skip_unscored_and_blank_rows():
    calls, rowpos, rowins, _ = _toy()
    rec = dc.call_records(calls, rowpos, rowins)
    assert rec.shape[0] == 5 + 6
    assert rec["window"].tolist() == [0] * 5 + [1] * 6
    assert rec["row"].tolist() == [2, 3, 4, 5, 6, 0, 1, 2, 3, 4, 5]
    assert rec["pos"][:5].tolist() == [1002, 1002, 1002, 1003, 1004]
    assert rec["ins"][:5].tolist() == [0, 1, 2, 0, 0]
    assert rec["call"][:5].tolist() == [1, 2, 2, 0, 1]
    calls[0, 7] = 1                                     # scored by the vote kernel but the row is blank: no record
    assert dc.call_records(calls, rowpos, rowins).shape[0] == 11
    off = dc.call_records(calls, rowpos, rowins, window0=40)
    assert off["window"].min() == 40


def test_merge_keeps_the_nearest_window():
    calls, rowpos, rowins, win_qs = _toy()
    rec = dc.call_records(calls, rowpos, rowins)
    m = dc.merge_records(rec, win_qs, flank=100)
    key = list(zip(m["pos"].tolist(), m["ins"].tolist()))
    assert key == sorted(set(key))                      # one record per (pos, ins), sorted
    # 1003 / 1004: centre of window 0 is 1100 (97 / 96 away), centre of window 1 is 1200 (197 / 196 away) -> window 0
    d = {(int(x["pos"]), int(x["ins"])): (int(x["window"]), int(x["call"])) for x in m}
    assert d[(1003, 0)] == (0, 0) and d[(1004, 0)] == (0, 1)
    assert d[(1002, 1)] == (0, 2) and d[(1100, 0)] == (1, 0)
    # equal distance: the earlier window wins
    r2 = np.zeros(2, dc.RECORD)
    r2["window"] = [1, 0]; r2["pos"] = [1150, 1150]; r2["call"] = [2, 1]
    assert int(dc.merge_records(r2, win_qs, flank=100)["call"][0]) == 1
    assert dc.merge_records(rec[:0], win_qs).shape[0] == 0


def test_tsv_round_trip(tmp_path):
    calls, rowpos, rowins, win_qs = _toy()
    rec = dc.merge_records(dc.call_records(calls, rowpos, rowins), win_qs)
    p = str(tmp_path / "calls.tsv")
    n = dc.write_tsv(p, "chr3", rec)
    assert n == rec.shape[0]
    lines = open(p).read().splitlines()
    assert lines[0] == dc.HEADER.strip("\n") and len(lines) == n + 1
    assert lines[1].split("\t")[:4] == ["chr3", "1003", "0", "1"]       # 1-based position of record (1002, 0)
    chroms, back = dc.read_tsv(p)
    assert set(chroms) == {"chr3"} and np.array_equal(back, rec)
    assert dc.write_tsv(p, "chr3", rec[:2], append=True) == 2
    assert dc.read_tsv(p)[1].shape[0] == n + 2
    k = dc.write_tsv(p, "chr3", rec, only_class=2)
    assert k == int((rec["call"] == 2).sum())


def test_vcf_writer_sites_and_genotypes(tmp_path):
    calls, rowpos, rowins, win_qs = _toy()
    rec = dc.merge_records(dc.call_records(calls, rowpos, rowins), win_qs)
    ref = (b"ACGT" * 60)                                 # positions 1000..1239
    p = str(tmp_path / "calls.vcf")
    n = dc.write_vcf(p, "chr3", rec, ref, ref_base=1000, sample="NA12878")
    body = [ln.rstrip("\n").split("\t") for ln in open(p) if not ln.startswith("#")]
    want = [x for x in rec if int(x["call"]) in (2, 3)]  # hom-ref (1) and unlabelled (0) rows are not variant sites
    assert n == len(body) == len(want) > 0
    for ln, x in zip(body, want):
        assert ln[0] == "chr3" and int(ln[1]) == int(x["pos"]) + 1
        assert ln[3] == chr(ref[int(x["pos"]) - 1000])
        assert ln[4] == ("<INS>" if x["ins"] > 0 else "<NON_REF>")
        assert ("INSROW=%d" % x["ins"] in ln[7]) == bool(x["ins"] > 0)
        assert ln[9] == dc.CLASS_GENOTYPES[int(x["call"])]
    head = [ln for ln in open(p) if ln.startswith("#")]
    assert head[0].startswith("##fileformat=VCFv4.2") and head[-1].rstrip("\n").endswith("FORMAT\tNA12878")
    # hom-ref lines on request; a custom class map; unsorted records and positions outside the slice are refused
    assert dc.write_vcf(p, "chr3", rec, ref, 1000, emit_hom_ref=True) == int(np.isin(rec["call"], (1, 2, 3)).sum())
    assert dc.write_vcf(p, "chr3", rec, ref, 1000, class_genotypes={0: "1/1"}) == int((rec["call"] == 0).sum())
    import pytest
    with pytest.raises(ValueError):
        dc.write_vcf(p, "chr3", rec[::-1], ref, 1000)
    with pytest.raises(IndexError):
        dc.write_vcf(p, "chr3", rec, ref, 1100)
