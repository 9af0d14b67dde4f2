"""TEST INFRASTRUCTURE ONLY -- golden fixture for the window / label producer.

Imports the UNMODIFIED /root/reference/src/utils/pileupGenerator.py (through the pysam / pyfaidx stand-ins of
oracle/standins) and runs its own `generatePileupBasedonVCF` on the bundled NA12878 data: 260 windows, each with the label
string and insert lengths `getLabel` derives from the VCF (pileupGenerator.py:65-95), the PNG and the summary CSV line
(:119).  Stored in tests/golden/bundled_windows.npz: record positions, the label drafts, the insert-length dictionaries,
the output labels of the summary CSV and the genotype-class dictionary `allVariantRecord` (+ bundled_confident.vcf: the QUAL > 60
lines of the bundled VCF, the only ones that reach the dictionary or the window list) -- the images themselves are
those of bundled_na12878.npz (the label does not touch the image; checked here).

    python oracle/make_windows_golden.py
"""
import json
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness as rh                                    # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")


def main():
    assert rh.reference_available(), "needs /root/reference"
    rh.import_reference_pileup()                            # stand-ins + src/utils on sys.path
    import pileupGenerator as pg                            # the unmodified reference driver
    from PIL import Image
    fasta = rh.ensure_fixture_fasta()
    # what getLabel returns per record, before the driver runs (same dictionary the driver fills)
    pg.allVariantRecord.clear()
    pg.populateRecordDictionary("3", rh.REF_VCF)
    keys = np.array(sorted(pg.allVariantRecord.keys()), np.int64)
    vals = np.array([[pg.allVariantRecord[k][0], pg.allVariantRecord[k][1] or 0, pg.allVariantRecord[k][2] or 0]
                     for k in keys], np.int32)
    pos, drafts, inserts, deletes = [], [], [], []
    for rec in pg.VariantFile(rh.REF_VCF).fetch(region="chr3"):
        if rec.qual is not None and rec.qual > 60 and pg.getClassForGenotype(pg.getGTField(rec)) != 1:
            lab, il, dl = pg.getLabel(rec.pos - 100 - 1, rec.pos + 100)
            pos.append(rec.pos); drafts.append(lab); inserts.append(json.dumps(il)); deletes.append(json.dumps(dl))
    out = tempfile.mkdtemp(prefix="dpwin_") + "/"
    pg.allVariantRecord.clear()
    pg.generatePileupBasedonVCF("3", rh.REF_BAM, fasta, rh.REF_VCF, out, 100)
    lines = [ln.rstrip("\n").split(",") for ln in open(out + "summary-3.csv")]
    assert len(lines) == len(pos)
    names = [os.path.basename(p) for p, _ in lines]
    labels = [l for _, l in lines]
    gold = np.load(os.path.join(GOLD, "bundled_na12878.npz"))
    assert gold["vcf_pos"].tolist() == pos
    for i, (p, _) in enumerate(lines):
        assert np.array_equal(np.array(Image.open(p)), gold["img"][i]), "image %d differs from bundled_na12878.npz" % i
    for p in sorted(set(p for p, _ in lines)):              # two records may share a position (one file name)
        os.remove(p)
    os.remove(out + "summary-3.csv")
    os.rmdir(out)
    np.savez_compressed(os.path.join(GOLD, "bundled_windows.npz"), vcf_pos=np.asarray(pos, np.int32),
                        label_draft=np.array(drafts), insert_lengths=np.array(inserts), delete_lengths=np.array(deletes),
                        label_out=np.array(labels), png_name=np.array(names), record_keys=keys, record_vals=vals)
    # the confident records (QUAL > 60) are the only lines that reach the dictionary or the window list: a cut-down VCF of
    # exactly those lines is the input fixture of the producer tests (the full file does not travel to the GPU box)
    with open(rh.REF_VCF) as fh, open(os.path.join(GOLD, "bundled_confident.vcf"), "w") as out_vcf:
        for ln in fh:
            f = ln.split("\t")
            if ln.startswith("#CHROM") or ln.startswith("##fileformat"):
                out_vcf.write(ln)
            elif not ln.startswith("#") and f[5] != "." and float(f[5]) > 60:
                out_vcf.write(ln)
    n_var = sum(1 for l in labels if set(l) - {"1", "0"})
    print("windows:", len(pos), "labels with a variant digit:", n_var, "dictionary entries:", len(keys))


if __name__ == "__main__":
    main()
