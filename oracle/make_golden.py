"""TEST INFRASTRUCTURE ONLY -- generates the committed fixtures under tests/golden/ by running the
UNMODIFIED reference encoder (/root/reference/src/utils/Pileup.py) through the stand-ins.

    python oracle/make_golden.py            # needs /root/reference (this build container only)

Outputs (np.savez_compressed):
  tests/golden/bundled_na12878.npz   all chr3 records of data/aln_NA12878.pg.chr3_100k-200k.bam as SoA, the
                                     MD-reconstructed reference segment, the 260 qualifying VCF windows
                                     (pileupGenerator.py:104-112) and the reference's 260 images + labels
  tests/golden/quirks.npz            hand-built reads for every quirk of SURVEY.md A.1 (+ ops the BAM lacks)
  tests/golden/fuzz.npz              seeded random CIGAR soup, all ten op codes
  tests/golden/nanopore_small.npz    deepore_b200.synth reads (long, indel-rich) over a 30 kb reference
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness as rh  # noqa: E402
from deepore_b200.reads import ReadBatch, FastaFile, NT16  # noqa: E402
from deepore_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
E_KEY, E_INDEX = -3, -4


def batch_arrays(rb, prefix=""):
    return {prefix + k: getattr(rb, k) for k in
            ("pos", "flag", "mapq", "l_seq", "cigar_off", "cigar", "seq_off", "seq4", "qual_off", "qual")}


def records_for_standin(rb):
    import pysam
    recs = []
    for i in range(rb.n_reads):
        c0, c1 = int(rb.cigar_off[i]), int(rb.cigar_off[i + 1])
        cig = [(int(c) & 0xF, int(c) >> 4) for c in rb.cigar[c0:c1]]
        l = int(rb.l_seq[i])
        pk = rb.seq4[int(rb.seq_off[i]):int(rb.seq_off[i + 1])]
        s = []
        for b in pk.tolist():
            s.append(NT16[b >> 4]); s.append(NT16[b & 0xF])
        seq = "".join(s[:l])
        q = rb.qual[int(rb.qual_off[i]):int(rb.qual_off[i + 1])].tobytes()
        recs.append(pysam.make_record(int(rb.pos[i]), int(rb.mapq[i]), cig, seq, q, flag=int(rb.flag[i])))
    return recs


def run_reference_on_batch(P, rb, ref_bytes, ref_base, win_qs, flank=100, tag="mem", label=None,
                           variantLengths=None, **kw):
    """Run the unmodified reference on in-memory records.  ref_bytes covers genomic [ref_base, ...)."""
    import pysam
    import pyfaidx
    bam = "mem:%s.bam" % tag
    fa = "mem:%s.fa" % tag
    pysam.register_memory_file(bam, ["chr3"], [ref_base + len(ref_bytes)], records_for_standin(rb))
    pyfaidx.register_memory_fasta(fa, {"chr3": "N" * ref_base + bytes(ref_bytes).decode("ascii")})
    H = kw.get("windowCutoff", 150)
    W = kw.get("coverageCutoff", 200)
    imgs = np.zeros((len(win_qs), H, W, 4), np.uint8)
    status = np.zeros(len(win_qs), np.int32)
    labels = []
    tmp = tempfile.mkdtemp(prefix="dpgold_")
    for i, qs in enumerate(win_qs):
        try:
            img, lab = rh.run_reference_window(P, bam, fa, "3", int(qs) + flank, flank=flank, tmpdir=tmp,
                                               label=label, variantLengths=variantLengths, **kw)
            imgs[i] = img
            labels.append(lab)
        except KeyError:
            status[i] = E_KEY; labels.append("")
        except IndexError:
            status[i] = E_INDEX; labels.append("")
    os.rmdir(tmp)
    return imgs, status, np.array(labels)


# ------------------------------------------------------------------------------------------------
def rand_seq(rng, n, alphabet="ACGT"):
    return "".join(alphabet[i] for i in rng.integers(0, len(alphabet), size=n))


def qlen_of(cig):
    return sum(n for op, n in cig if op in (0, 1, 4, 7, 8))


def seq_from_ref(rng, ref, pos, cig, err=0.1):
    """A read that mostly matches the reference along its CIGAR (so 'M' entries dominate)."""
    out = []
    p = pos
    for op, n in cig:
        if op in (0, 7, 8):
            for i in range(n):
                c = chr(ref[p + i]) if 0 <= p + i < len(ref) else "A"
                if c not in "ACGT" or rng.random() < err:
                    c = "ACGT"[rng.integers(0, 4)]
                out.append(c)
            p += n
        elif op in (2, 3):
            p += n
        elif op in (1, 4):
            out.append(rand_seq(rng, n))
    return "".join(out)


def make_quirks(P):
    rng = np.random.default_rng(1234)
    G = 3000
    ref = synth.random_reference(G, seed=77).copy()
    ref[1190:1193] = ord("N")
    ref[1300] = ord("N")
    qs0 = 1100                                   # main window [1100, 1301)
    recs = []

    def add(pos, cig, mapq=60, flag=0, seq=None, qual=None, err=0.1):
        if isinstance(cig, str):
            from deepore_b200.reads import parse_cigar_string
            cig = parse_cigar_string(cig)
        if seq is None:
            seq = seq_from_ref(rng, ref, pos, cig, err)
        if qual is None:
            qual = rng.integers(2, 42, size=len(seq)).astype(np.uint8).tobytes()
        recs.append(dict(pos=pos, cigar=cig, mapq=mapq, flag=flag, seq=seq, qual=qual))

    # (a) plain matches at every alignment of refStart vs qs  (SURVEY A.1 "rel" shift when refStart == qs)
    for d in (-150, -101, -100, -99, -50, -2, -1, 0, 1, 2, 50, 99, 100, 150, 199, 200, 201):
        add(qs0 + d, "101M", err=0.0)
    add(qs0 - 101, "101M")      # endpos == qs  -> not fetched
    add(qs0 - 100, "101M")      # endpos == qs+1 -> fetched
    add(qs0 + 200, "50M")       # pos == qe-1
    add(qs0 + 201, "50M")       # pos == qe -> not fetched
    # (b) soft clips
    add(qs0 + 10, "4S74M23S"); add(qs0 - 30, "20S60M"); add(qs0 + 150, "30M70S"); add(qs0 + 190, "5S20M")
    add(qs0 - 60, "10S50M10S"); add(qs0 - 60, "60M8S"); add(qs0, "3S40M"); add(qs0 + 1, "3S40M")
    add(qs0 + 200, "7S30M"); add(qs0 + 40, "81S20M")
    # (c) hard clips
    add(qs0 + 20, "5H50M"); add(qs0 + 25, "50M7H"); add(qs0 + 30, "3H4S40M2S6H"); add(qs0 - 20, "4H30M")
    # (d) N, =, X, P, B
    add(qs0 + 5, "20M30N20M"); add(qs0 - 40, "30M50N30M"); add(qs0 + 60, "10=2X10=5M"); add(qs0 + 70, "5M3P5M")
    add(qs0 + 80, "10M2B10M"); add(qs0 + 100, "4X4=4X"); add(qs0 - 3, "6=20M")
    # (e) inserts
    add(qs0 - 10, "10M3I40M")        # anchor == qs      (needs A > qs -> dropped)
    add(qs0 - 10, "11M2I40M")        # anchor == qs+1
    add(qs0 + 100, "20M1I20M"); add(qs0 + 100, "20M4I20M"); add(qs0 + 100, "20M2I20M", mapq=25)
    add(qs0 + 150, "50M6I10M")       # anchor == qe-1
    add(qs0 + 150, "51M3I10M")       # anchor == qe
    add(qs0 + 150, "60M3I10M")       # anchor > qe
    add(qs0 + 30, "30M5I")           # insert as last op: anchor column not populated by this read
    add(qs0 + 31, "29M2I4S")         # insert followed by trailing S
    add(qs0 + 130, "5M2I1M3I1M1I20M")
    # (f) deletions and n=1 ops
    add(qs0 - 20, "15M10D60M"); add(qs0 + 180, "15M10D60M"); add(qs0 + 90, "1M1D1M1D1M1I1M2D30M")
    add(qs0 + 95, "10M1D10M"); add(qs0 - 5, "5M5D50M"); add(qs0 + 196, "5M5D50M")
    # (h) filters
    add(qs0 + 50, "60M", mapq=19); add(qs0 + 50, "60M", mapq=20); add(qs0 + 50, "60M", mapq=0)
    add(qs0 + 50, "60M", flag=4)     # FUNMAP with a cigar: endpos = pos+1, still fetched and parsed
    add(qs0 + 55, [], seq="ACGTACGTAC")          # no cigar
    add(qs0 + 56, "30S"); add(qs0 + 57, "5S10I5S")
    # (i) N bases in reads
    s = seq_from_ref(rng, ref, qs0 + 85, [(0, 120)], 0.05)
    s = s[:10] + "N" + s[11:50] + "NN" + s[52:]
    add(qs0 + 85, "120M", seq=s)
    s = seq_from_ref(rng, ref, qs0 + 60, [(0, 30), (1, 3), (0, 30)], 0.05)
    s = s[:31] + "N" + s[32:]
    add(qs0 + 60, "30M3I30M", seq=s)
    # (j) CIGAR that starts with D / N (refStart != pos)
    add(qs0 + 12, "3D40M"); add(qs0 - 12, "5N40M"); add(qs0 + 14, "2S3D40M")
    # long read starting far left of the window (block-skip rule) and one spanning it entirely
    cig = []
    for i in range(60):
        cig += [(0, int(rng.integers(5, 30))), (int(rng.choice([1, 2])), int(rng.integers(1, 4)))]
    cig += [(0, 50)]
    add(100, cig)
    add(300, cig[10:])
    recs.sort(key=lambda r: r["pos"])
    rb = ReadBatch.from_records(recs)
    wins = [qs0, qs0 - 100, qs0 + 100, qs0 + 57, 1199, 900, 1500, 2500, 150, 700]
    imgs, status, labels = run_reference_on_batch(P, rb, ref.tobytes(), 0, wins, tag="quirks")
    # a second pass with a label/variantLengths to pin labelInsertRegion (Pileup.py:320-333)
    lab_in = "".join(str(1 + (i % 3 == 0)) for i in range(201))
    vl = {21: 3, 33: 2, 101: 5, 140: 2, 151: 4}
    _, _, labels2 = run_reference_on_batch(P, rb, ref.tobytes(), 0, wins, tag="quirks", label=lab_in, variantLengths=vl)
    # narrow / shallow image parameters
    imgs_small, st_small, _ = run_reference_on_batch(P, rb, ref.tobytes(), 0, wins[:4], tag="quirks", flank=40,
                                                     coverageCutoff=32, windowCutoff=60, mapQualityCutoff=30)
    out = dict(batch_arrays(rb), ref=ref, ref_base=np.int64(0), win_qs=np.asarray(wins, np.int32), img=imgs,
               status=status, labels=labels, label_in=np.array(lab_in), vl_keys=np.array(list(vl.keys())),
               vl_vals=np.array(list(vl.values())), labels2=labels2,
               small_win_qs=np.asarray([w + 60 for w in wins[:4]], np.int32),
               small_img=imgs_small, small_status=st_small)
    # NB run_reference_on_batch derives position = qs + flank; with flank=40 the qs used is wins[i] (position
    # = wins[i]+40 -> queryStart = wins[i]); record what was actually used:
    out["small_win_qs"] = np.asarray(wins[:4], np.int32)
    np.savez_compressed(os.path.join(GOLD, "quirks.npz"), **out)
    print("quirks:", rb.n_reads, "reads", len(wins), "windows; status", status.tolist())


def make_deep(P):
    """Depth > coverageCutoff (truncation after the category sort) and > windowCutoff rows (insert rows)."""
    rng = np.random.default_rng(99)
    G = 1200
    ref = synth.random_reference(G, seed=5)
    recs = []
    for i in range(420):
        pos = int(rng.integers(200, 520))
        cig = []
        left = 180
        while left > 0:
            m = int(min(left, rng.integers(1, 25)))
            cig.append((0, m)); left -= m
            if left > 0:
                op = int(rng.choice([1, 2], p=[0.7, 0.3]))
                cig.append((op, int(rng.integers(1, 5))))
        seq = seq_from_ref(rng, ref, pos, cig, 0.15)
        q = rng.integers(1, 45, size=len(seq)).astype(np.uint8).tobytes()
        recs.append(dict(pos=pos, cigar=cig, mapq=int(rng.choice([60, 30, 22, 10])), seq=seq, qual=q))
    recs.sort(key=lambda r: r["pos"])
    rb = ReadBatch.from_records(recs)
    wins = [400, 450, 520, 300]
    imgs, status, labels = run_reference_on_batch(P, rb, ref.tobytes(), 0, wins, tag="deep")
    np.savez_compressed(os.path.join(GOLD, "deep.npz"), **batch_arrays(rb), ref=ref, ref_base=np.int64(0),
                        win_qs=np.asarray(wins, np.int32), img=imgs, status=status, labels=labels)
    print("deep:", rb.n_reads, "reads; status", status.tolist())


def make_fuzz(P):
    rng = np.random.default_rng(4321)
    G = 2500
    ref = synth.random_reference(G, seed=11, n_frac=0.01)
    recs = []
    for i in range(260):
        pos = int(rng.integers(0, G - 400))
        nops = int(rng.integers(1, 12))
        cig = []
        for j in range(nops):
            op = int(rng.choice([0, 1, 2, 3, 4, 5, 6, 7, 8, 9], p=[0.42, 0.12, 0.12, 0.05, 0.06, 0.04, 0.03, 0.07, 0.06, 0.03]))
            cig.append((op, int(rng.integers(1, 40))))
        ql = qlen_of(cig)
        if ql == 0:
            continue
        seq = seq_from_ref(rng, ref, pos, cig, 0.2)
        if rng.random() < 0.1:
            k = int(rng.integers(0, len(seq)))
            seq = seq[:k] + "N" + seq[k + 1:]
        q = rng.integers(0, 60, size=len(seq)).astype(np.uint8).tobytes()
        recs.append(dict(pos=pos, cigar=cig, mapq=int(rng.choice([60, 40, 21, 20, 19, 3])), seq=seq, qual=q,
                         flag=int(rng.choice([0, 16, 1024, 4], p=[0.6, 0.3, 0.07, 0.03]))))
    recs.sort(key=lambda r: r["pos"])
    rb = ReadBatch.from_records(recs)
    wins = [int(x) for x in rng.integers(0, G - 260, size=48)]
    imgs, status, labels = run_reference_on_batch(P, rb, ref.tobytes(), 0, wins, tag="fuzz")
    np.savez_compressed(os.path.join(GOLD, "fuzz.npz"), **batch_arrays(rb), ref=ref, ref_base=np.int64(0),
                        win_qs=np.asarray(wins, np.int32), img=imgs, status=status, labels=labels)
    print("fuzz:", rb.n_reads, "reads", len(wins), "windows; status", status.tolist())


def make_nanopore(P):
    G = 30_000
    base = 5000
    ref = synth.random_reference(G, seed=1)
    rb = synth.simulate_reads(ref, ref_base=base, coverage=30, seed=2, mean_len=4000, min_len=300, max_len=20000)
    wins = synth.window_starts(base + 2000, base + 8000, stride=150)
    imgs, status, labels = run_reference_on_batch(P, rb, ref.tobytes(), base, wins.tolist(), tag="nano")
    np.savez_compressed(os.path.join(GOLD, "nanopore_small.npz"), **batch_arrays(rb), ref=ref, ref_base=np.int64(base),
                        win_qs=wins, img=imgs, status=status, labels=labels)
    print("nanopore:", rb.n_reads, "reads", len(wins), "windows; status", status.tolist())


def make_bundled(P):
    rb = ReadBatch.from_bam(rh.REF_BAM, "chr3")
    fa = FastaFile(rh.ensure_fixture_fasta())
    lo_g, hi_g = 99_700, 200_400
    ref = fa.contig("chr3")[lo_g:hi_g].copy()
    vpos = rh.qualifying_vcf_positions()
    wins = np.asarray([p - 1 - 100 for p in vpos], np.int32)      # position = rec.pos - 1 (pileupGenerator.py:112)
    imgs = np.zeros((len(wins), 150, 200, 4), np.uint8)
    labels = []
    for i, p in enumerate(vpos):
        imgs[i], lab = rh.run_reference_window(P, rh.REF_BAM, rh.FIXTURE_FASTA, "3", p - 1)
        labels.append(lab)
    np.savez_compressed(os.path.join(GOLD, "bundled_na12878.npz"), **batch_arrays(rb), ref=ref,
                        ref_base=np.int64(lo_g), win_qs=wins, vcf_pos=np.asarray(vpos, np.int32), img=imgs,
                        status=np.zeros(len(wins), np.int32), labels=np.array(labels))
    print("bundled:", rb.n_reads, "reads", len(wins), "windows")


if __name__ == "__main__":
    assert rh.reference_available(), "needs /root/reference"
    os.makedirs(GOLD, exist_ok=True)
    P = rh.import_reference_pileup()
    which = sys.argv[1:] or ["quirks", "deep", "fuzz", "nanopore", "bundled"]
    for w in which:
        globals()["make_" + w](P)
