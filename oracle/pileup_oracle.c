/*
 * TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C) of the reference pileup encoder.
 *
 * Follows /root/reference/src/utils/Pileup.py statement by statement; each block cites the lines it
 * restates.  It is the checker for the CUDA encoder (tests/, __graft_entry__.smoke(), bench.py's
 * cpu_baseline / --impl reference legs) and is never linked into or called by the product library.
 *
 * Parity status: PINNED against the reference itself -- tests/test_oracle_pins.py runs the unmodified
 * Pileup.py (through the pysam/pyfaidx stand-ins in oracle/standins) on the bundled NA12878 BAM and on
 * synthetic quirk reads and requires byte equality with this file's output; the outputs of that run are
 * committed under tests/golden/.  (The reference ships no golden vectors of its own: SURVEY.md 8c.)
 *
 * Inputs use the same SoA read batch as the product C-ABI (include/deepore_b200.h), with HOST pointers.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/deepore_b200.h"

#define NCAT 7
/* sortingKey (Pileup.py:66-73): A0 C1 G2 T3 D4 M5 (I6) N7 -> dense order A,C,G,T,D,M,N */
enum { CAT_A = 0, CAT_C = 1, CAT_G = 2, CAT_T = 3, CAT_D = 4, CAT_M = 5, CAT_N = 6 };

static const char NT16[] = "=ACMGRSVTWYHKDBN";              /* dep/htslib/hts.c:82 */
/* deltaRef / deltaRead / deltaRefRelative  (Pileup.py:47-51), ops M I D N S H P = X B */
static const int D_REF[16]  = {1, 0, 1, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
static const int D_READ[16] = {1, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
static const int D_REL[16]  = {1, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};

/* SNPtoRGB (Pileup.py:56-64) */
static void colour_of_char(int ch, uint8_t *rgb, int *ok) {
    *ok = 1;
    switch (ch) {
    case 'M': rgb[0] = 255; rgb[1] = 255; rgb[2] = 255; break;
    case 'A': rgb[0] = 255; rgb[1] = 0;   rgb[2] = 0;   break;
    case 'C': rgb[0] = 255; rgb[1] = 255; rgb[2] = 0;   break;
    case 'G': rgb[0] = 0;   rgb[1] = 255; rgb[2] = 0;   break;
    case 'T': rgb[0] = 0;   rgb[1] = 0;   rgb[2] = 255; break;
    case 'I': rgb[0] = 255; rgb[1] = 0;   rgb[2] = 255; break;
    case 'D': rgb[0] = 0;   rgb[1] = 255; rgb[2] = 255; break;
    case 'N': rgb[0] = 0;   rgb[1] = 0;   rgb[2] = 0;   break;
    default:  rgb[0] = rgb[1] = rgb[2] = 0; *ok = 0;    break;     /* KeyError in the reference */
    }
}
static const char CAT_CHAR[NCAT] = {'A', 'C', 'G', 'T', 'D', 'M', 'N'};

static int cat_of_char(int ch) {
    switch (ch) {
    case 'A': return CAT_A; case 'C': return CAT_C; case 'G': return CAT_G; case 'T': return CAT_T;
    case 'D': return CAT_D; case 'M': return CAT_M; case 'N': return CAT_N; default: return -1;
    }
}

typedef struct { uint8_t *v; int n, cap; } bytevec;
static int bv_push(bytevec *b, uint8_t x) {
    if (b->n == b->cap) {
        int nc = b->cap ? b->cap * 2 : 8;
        uint8_t *nv = (uint8_t *)realloc(b->v, (size_t)nc);
        if (!nv) return -1;
        b->v = nv; b->cap = nc;
    }
    b->v[b->n++] = x;
    return 0;
}

/* insertColumns[anchor] = list over k of {A,C,G,T,N: list}  (Pileup.py:103-115) */
typedef struct { bytevec *cells; int nk, capk; } insert_col;     /* cells[k*5 + base] */

typedef struct {
    int L;                    /* 2*flank+1 */
    bytevec *cols;            /* [L][NCAT]  pileupColumns[qs+c][cat] quality bytes in append order */
    uint8_t *populated;       /* [L]  `i in self.pileupColumns` */
    insert_col *ins;          /* [L]  insertColumns[qs+c] (anchors outside [qs,qe) can never be rendered) */
} window_state;

static int python_index(long idx, long len, long *out) {       /* list/str indexing incl. negatives */
    if (idx < 0) idx += len;
    if (idx < 0 || idx >= len) return -1;
    *out = idx;
    return 0;
}

typedef struct {
    const dp_reads *R;
    int r;
    long aln_start, aln_len;  /* query_alignment_start / len(query_alignment_sequence) */
} read_view;

static int seq_char(const read_view *rv, long i) {             /* readSequence[i] on the clipped sequence */
    long k;
    if (python_index(i, rv->aln_len, &k)) return -1;
    k += rv->aln_start;
    uint8_t b = rv->R->seq4[rv->R->seq_off[rv->r] + (k >> 1)];
    return NT16[(k & 1) ? (b & 0xF) : (b >> 4)];
}
static int qual_at(const read_view *rv, long i) {              /* readQualities[i] on the FULL query */
    long k;
    if (python_index(i, rv->R->l_seq[rv->r], &k)) return -1;
    return rv->R->qual[rv->R->qual_off[rv->r] + k];
}

/* Pileup.parseRead (Pileup.py:174-264) */
static int parse_read(window_state *W, const dp_reads *R, int r, long qs, long qe, const uint8_t *refseq,
                      long reflen, const uint8_t *lut) {
    const uint32_t *cig = R->cigar + R->cigar_off[r];
    long ncig = (long)(R->cigar_off[r + 1] - R->cigar_off[r]);
    long mapq = R->mapq[r];
    /* get_reference_positions(): first M/=/X position  (:183,:191) */
    long refStart = -1;
    {
        long p = R->pos[r];
        for (long c = 0; c < ncig; c++) {
            int op = cig[c] & 0xF; long n = cig[c] >> 4;
            if (op == 0 || op == 7 || op == 8) { if (n > 0) { refStart = p; break; } }
            else if (op == 2 || op == 3) p += n;
        }
        if (refStart < 0) return 0;                           /* len(refPositions) == 0  (:187-189) */
    }
    read_view rv; rv.R = R; rv.r = r;
    {   /* query_alignment_sequence (:197): strip leading/trailing S (H skipped) */
        long l = R->l_seq[r], s = 0, e = l;
        for (long c = 0; c < ncig; c++) { int op = cig[c] & 0xF; if (op == 5) continue; if (op == 4) s += cig[c] >> 4; else break; }
        for (long c = ncig - 1; c >= 0; c--) { int op = cig[c] & 0xF; if (op == 5) continue; if (op == 4) e -= cig[c] >> 4; else break; }
        rv.aln_start = s; rv.aln_len = e - s; if (rv.aln_len < 0) rv.aln_len = 0;
    }
    long refPosition = 0, readPosition = -1, rel = 0;         /* :199-201 */
    if (refStart > qs) rel += refStart - qs - 1;              /* :203-204 */

    for (long c = 0; c < ncig; c++) {
        int snp = cig[c] & 0xF; long n = cig[c] >> 4;
        long A = refPosition + refStart;                      /* :212 */
        if (A < qs - n) {                                     /* :214-216 */
            refPosition += D_REF[snp] * n; readPosition += D_READ[snp] * n;
            continue;
        }
        if (snp == 1) {                                       /* :220-230 */
            if (A > qs) {
                /* slices readSequence[rp+1:rp+n+1], readQualities[rp+1:rp+n+1]; addInsertEntry indexes [k]
                   for k < n, so a short slice raises IndexError */
                if (A < qe) {                                 /* anchors >= qe are stored but never rendered */
                    insert_col *ic = &W->ins[A - qs];
                    for (long k = 0; k < n; k++) {
                        long si = readPosition + 1 + k;
                        if (si < 0 || si >= rv.aln_len) return DP_EINDEX;
                        int ch = seq_char(&rv, si);
                        if (si >= R->l_seq[r]) return DP_EINDEX;
                        int bq = qual_at(&rv, si);
                        int base;
                        switch (ch) { case 'A': base = 0; break; case 'C': base = 1; break; case 'G': base = 2; break;
                                      case 'T': base = 3; break; case 'N': base = 4; break; default: return DP_EKEY; }
                        if (k >= ic->nk) {
                            if (k >= ic->capk) {
                                int nc = ic->capk ? ic->capk * 2 : 4; while (nc <= k) nc *= 2;
                                bytevec *nv = (bytevec *)realloc(ic->cells, sizeof(bytevec) * 5 * (size_t)nc);
                                if (!nv) return DP_ENOMEM;
                                memset(nv + 5 * ic->capk, 0, sizeof(bytevec) * 5 * (size_t)(nc - ic->capk));
                                ic->cells = nv; ic->capk = nc;
                            }
                            ic->nk = (int)k + 1;
                        }
                        long q = bq < mapq ? bq : mapq;       /* calculateQuality (:166-171) */
                        if (bv_push(&ic->cells[k * 5 + base], lut[q])) return DP_ENOMEM;
                    }
                } else {
                    /* still evaluates the slices' characters -> KeyError possible; we only validate bounds */
                }
            }
            refPosition += D_REF[snp] * n; readPosition += D_READ[snp] * n;
            continue;
        }
        for (long i = 0; i < n; i++) {                        /* :235-262 */
            if (A >= qs && A < qe) {
                int rc = seq_char(&rv, readPosition);         /* :240 */
                if (rc < 0) return DP_EINDEX;
                if (rel < 0 || rel >= reflen) return DP_EINDEX;
                int fc = refseq[rel];                         /* :241 */
                int bq = qual_at(&rv, readPosition);          /* :243 */
                if (bq < 0) return DP_EINDEX;
                long q = bq < mapq ? bq : mapq;
                int enc = (snp == 2) ? 'D' : (rc == fc ? 'M' : rc);   /* getPileupEncoding (:128-134) */
                int cat = cat_of_char(enc);
                if (cat < 0) return DP_EKEY;                  /* pileupColumns[i][enc] KeyError (:100) */
                long col = A - qs;
                W->populated[col] = 1;
                if (bv_push(&W->cols[col * NCAT + cat], lut[q])) return DP_ENOMEM;
                rel += D_REL[snp];                            /* :252 */
                A = refPosition + refStart;                   /* :254  (assigned BEFORE the bump) */
                refPosition += D_REF[snp]; readPosition += D_READ[snp];
                if (snp == 4) break;                          /* :257-258 */
            } else {
                A = refPosition + refStart;
                refPosition += D_REF[snp]; readPosition += D_READ[snp];
            }
        }
    }
    return 0;
}

static void free_window(window_state *W) {
    for (int i = 0; i < W->L * NCAT; i++) free(W->cols[i].v);
    for (int i = 0; i < W->L; i++) {
        for (int k = 0; k < W->ins[i].capk * 5; k++) free(W->ins[i].cells[k].v);
        free(W->ins[i].cells);
    }
    free(W->cols); free(W->populated); free(W->ins);
}

/* one window: Pileup.__init__ .. savePileupRGB */
static int encode_one(const dp_reads *R, long qs, int lo, int hi, const uint8_t *ref, long ref_base,
                      long ref_len, const dp_encode_params *P, const uint8_t *lut, uint8_t *img,
                      int32_t *rowpos, int32_t *rowins) {
    const int L = 2 * P->flank + 1, Wd = P->coverage_cutoff, H = P->window_cutoff;
    const long qe = qs + L;                                   /* :28-29 */
    /* refSequence = fasta[qs .. qe] inclusive = L+1 chars (:41) */
    if (qs < ref_base || qe + 1 > ref_base + ref_len) return DP_EINVAL;
    const uint8_t *refseq = ref + (qs - ref_base);
    const long reflen = L + 1;

    window_state W;
    W.L = L;
    W.cols = (bytevec *)calloc((size_t)L * NCAT, sizeof(bytevec));
    W.populated = (uint8_t *)calloc((size_t)L, 1);
    W.ins = (insert_col *)calloc((size_t)L, sizeof(insert_col));
    if (!W.cols || !W.populated || !W.ins) return DP_ENOMEM;

    int rc = 0;
    for (int r = lo; r < hi && rc == 0; r++) {                /* iterateReads (:152-163), fetch rule hts.c:2242 */
        long pos = R->pos[r];
        long ncig = (long)(R->cigar_off[r + 1] - R->cigar_off[r]);
        long endpos;
        if (!(R->flag[r] & 4) && ncig > 0) {                  /* bam_endpos (sam.c:347-353) */
            long rl = 0;
            const uint32_t *cig = R->cigar + R->cigar_off[r];
            for (long c = 0; c < ncig; c++) { int op = cig[c] & 0xF; if (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) rl += cig[c] >> 4; }
            endpos = pos + rl;
        } else endpos = pos + 1;
        if (!(pos < qe && endpos > qs)) continue;
        if (R->mapq[r] < P->map_quality_cutoff) continue;     /* :160 */
        rc = parse_read(&W, R, r, qs, qe, refseq, reflen, lut);
    }
    if (rc) { free_window(&W); return rc; }

    /* generatePileupImage (:295-317) + encodeColumn (:268-292) + savePileupRGB (:336-361), emitting only the
       first H rows x first Wd pixels.  refMut mirrors the in-place 'I' splices of :315. */
    int maxrows = L;
    for (int c = 0; c < L; c++) if (W.populated[c]) maxrows += W.ins[c].nk;
    uint8_t *refMut = (uint8_t *)malloc((size_t)(reflen + maxrows + 1));
    if (!refMut) { free_window(&W); return DP_ENOMEM; }
    memcpy(refMut, refseq, (size_t)reflen);
    long refMutLen = reflen;

    for (int y = 0; y < H; y++) {
        for (int x = 0; x < Wd; x++) { uint8_t *px = img + ((size_t)y * Wd + x) * 4; px[0] = px[1] = px[2] = 0; px[3] = 255; }
        if (rowpos) rowpos[y] = -1;
        if (rowins) rowins[y] = 0;
    }
    int cidx = 0;
    for (int col = 0; col < L && rc == 0; col++) {
        if (!W.populated[col]) continue;
        /* encodeColumn(c, segs): ntRef = refSequence[c] */
        if (cidx < H) {
            uint8_t *row = img + (size_t)cidx * Wd * 4;
            int ok;
            if (cidx >= refMutLen) { rc = DP_EINDEX; break; }
            colour_of_char(refMut[cidx], row, &ok); row[3] = 255;
            if (!ok) { rc = DP_EKEY; break; }
            int x = 1;
            for (int cat = 0; cat < NCAT && x < Wd; cat++) {
                bytevec *b = &W.cols[col * NCAT + cat];
                uint8_t rgb[3]; colour_of_char(CAT_CHAR[cat], rgb, &ok);
                for (int e = 0; e < b->n && x < Wd; e++, x++) {
                    uint8_t *px = row + x * 4; px[0] = rgb[0]; px[1] = rgb[1]; px[2] = rgb[2]; px[3] = b->v[e];
                }
            }
            /* filler (0,0,0,255) already there */
            if (rowpos) rowpos[cidx] = (int32_t)(qs + col);
        } else {
            /* rows past the cut are still built by the reference (and would raise on a bad ref char) */
            int ok; uint8_t tmp[3];
            if (cidx < refMutLen) { colour_of_char(refMut[cidx], tmp, &ok); if (!ok) { rc = DP_EKEY; break; } }
            else { rc = DP_EINDEX; break; }
        }
        int nk = W.ins[col].nk;
        for (int k = 0; k < nk; k++) {
            cidx++;
            if (cidx < H) {
                uint8_t *row = img + (size_t)cidx * Wd * 4;
                for (int x = 0; x < Wd; x++) { uint8_t *px = row + x * 4; px[0] = 255; px[1] = 0; px[2] = 255; px[3] = 255; }
                int x = 1;
                static const char INS_CHAR[5] = {'A', 'C', 'G', 'T', 'N'};
                for (int base = 0; base < 5 && x < Wd; base++) {
                    bytevec *b = &W.ins[col].cells[k * 5 + base];
                    uint8_t rgb[3]; int ok; colour_of_char(INS_CHAR[base], rgb, &ok);
                    for (int e = 0; e < b->n && x < Wd; e++, x++) {
                        uint8_t *px = row + x * 4; px[0] = rgb[0]; px[1] = rgb[1]; px[2] = rgb[2]; px[3] = b->v[e];
                    }
                }
                if (rowpos) rowpos[cidx] = (int32_t)(qs + col);
                if (rowins) rowins[cidx] = k + 1;
            }
            /* self.refSequence = refSequence[:c] + 'I' + refSequence[c:]  (:315) */
            long at = cidx < refMutLen ? cidx : refMutLen;
            memmove(refMut + at + 1, refMut + at, (size_t)(refMutLen - at));
            refMut[at] = 'I'; refMutLen++;
        }
        cidx++;
    }
    free(refMut);
    free_window(&W);
    return rc;
}

/* Host twin of dp_encode_windows (include/deepore_b200.h).  lut = 256-entry calculateQuality table. */
int dpo_encode_windows(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs,
                       const int32_t *win_lo, const int32_t *win_hi,
                       const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                       const dp_encode_params *p, const uint8_t *lut,
                       uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins, int32_t *out_status) {
    if (!reads || !win_qs || !win_lo || !win_hi || !ref || !p || !lut || !out_img) return DP_EINVAL;
    const size_t img_bytes = (size_t)p->window_cutoff * p->coverage_cutoff * 4;
    int first = 0;
    for (int w = 0; w < n_windows; w++) {
        int rc = encode_one(reads, win_qs[w], win_lo[w], win_hi[w], ref, ref_base, ref_len, p, lut,
                            out_img + img_bytes * w,
                            out_rowpos ? out_rowpos + (size_t)w * p->window_cutoff : 0,
                            out_rowins ? out_rowins + (size_t)w * p->window_cutoff : 0);
        if (out_status) out_status[w] = rc;
        if (rc && !first) first = rc;
    }
    return first;
}
