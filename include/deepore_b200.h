/*
 * deepore_b200 -- C-ABI of the B200-native deePore hot path.
 *
 * The reference (kishwarshafin/deePore) has no FFI of its own: its boundary is the Python class API
 * of src/utils/Pileup.py and src/modules/model.py.  The Python shims in deepore_b200/ keep that API and
 * call the entry points below through ctypes / torch custom ops.  Every function takes plain pointers,
 * extents and a CUDA stream handle (void* == cudaStream_t); no torch types, no C++ exceptions.
 * Return value: 0 on success, a negative DP_E* code on failure; dp_last_error() gives the message of
 * the last failure on the calling thread.
 *
 * Each entry point cites the reference code it replaces (paths relative to the reference root).
 */
#ifndef DEEPORE_B200_H
#define DEEPORE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DP_OK            0
#define DP_EINVAL       -1   /* bad argument (null pointer, negative extent, window outside ref) */
#define DP_ECUDA        -2   /* CUDA runtime error, see dp_last_error() */
#define DP_EKEY         -3   /* a base outside {A,C,G,T,N} reached the colour table: the reference raises
                                KeyError at Pileup.py:100 / :281 */
#define DP_EINDEX       -4   /* an index the reference would raise IndexError on (Pileup.py:240-241) */
#define DP_ENOMEM       -5
#define DP_EUNSUPPORTED -6

/* ---- read batch: BAM records of ONE contig in file (coordinate-sorted) order, structure of arrays.
 * Field meaning follows the BAM record (dep/htslib/htslib/sam.h:154-195); nothing is pre-digested, so a
 * host reader can memcpy straight out of inflated BGZF blocks.  All pointers are DEVICE pointers for the
 * dp_* functions (HOST pointers for the dpo_* oracle twin in oracle/pileup_oracle.h). */
typedef struct dp_reads {
    int32_t        n_reads;
    int64_t        n_cigar;    /* total CIGAR ops == cigar_off[n_reads] (known to the host reader)      */
    const int32_t *pos;        /* [n_reads]   0-based leftmost coordinate (core.pos)                  */
    const uint16_t*flag;       /* [n_reads]   only BAM_FUNMAP (0x4) is consulted (bam_endpos)          */
    const uint8_t *mapq;       /* [n_reads]                                                            */
    const int32_t *l_seq;      /* [n_reads]   query length incl. soft clips                            */
    const int64_t *cigar_off;  /* [n_reads+1] offsets into cigar[] (in ops)                            */
    const uint32_t*cigar;      /* [cigar_off[n_reads]]  len<<4|op, ops "MIDNSHP=XB" (sam.h:65-109)     */
    const int64_t *seq_off;    /* [n_reads+1] BYTE offsets into seq4[]; read i has (l_seq+1)/2 bytes   */
    const uint8_t *seq4;       /* 4-bit packed bases, high nibble first, alphabet "=ACMGRSVTWYHKDBN"   */
    const int64_t *qual_off;   /* [n_reads+1] byte offsets into qual[]; read i has l_seq bytes         */
    const uint8_t *qual;       /* phred bytes of the FULL query (soft clips included)                  */
} dp_reads;

/* ---- encoder parameters == keyword arguments of PileUpGenerator.generatePileup (Pileup.py:418) */
typedef struct dp_encode_params {
    int32_t flank;             /* flankLength; window length = 2*flank+1 (Pileup.py:22)    default 100 */
    int32_t coverage_cutoff;   /* coverageCutoff = image width W  (Pileup.py:346-349)      default 200 */
    int32_t window_cutoff;     /* windowCutoff   = image height H                          default 150 */
    int32_t map_quality_cutoff;/* mapQualityCutoff (Pileup.py:160)                         default 20  */
} dp_encode_params;

const char *dp_last_error(void);
int  dp_version(void);
int  dp_device_count(void);

/*
 * dp_encode_windows -- replaces Pileup.__init__ + iterateReads + parseRead + generatePileupImage +
 * savePileupRGB (Pileup.py:21-93,152-264,268-317,336-361) for a BATCH of windows.
 *
 *   reads        device SoA (above)
 *   n_windows    number of windows
 *   win_qs       [n_windows] queryStart = position - flank  (Pileup.py:426), device
 *   win_lo/hi    [n_windows] candidate read range [lo,hi) in file order that contains every read with
 *                pos < queryEnd && endpos > queryStart (dep/htslib/hts.c:2242); the kernel re-applies the
 *                exact rule, so the range may be loose.  device
 *   ref          reference bases (upper-case ASCII) for genomic [ref_base, ref_base+ref_len), device
 *   out_img      [n_windows][H][W][4] uint8 RGBA, device -- byte-identical to the PNG pixels the reference
 *                writes (row = pileup column index, x = depth, pixel 0 = reference base)
 *   out_rowpos   optional [n_windows][H] int32: genomic position of the anchor of each image row, -1 for
 *                blank rows
 *   out_rowins   optional [n_windows][H] int32: 0 for an anchor row, k+1 for the k-th insert row
 *   out_status   optional [n_windows] int32: 0 or a DP_E* code raised inside that window
 *   workspace    device scratch of at least dp_encode_workspace_bytes(...) bytes (may be NULL if 0)
 */
int64_t dp_encode_workspace_bytes(int32_t n_reads, int64_t n_cigar, int32_t n_windows,
                                  const dp_encode_params *p);
int dp_encode_windows(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs,
                      const int32_t *win_lo, const int32_t *win_hi,
                      const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                      const dp_encode_params *p,
                      uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins, int32_t *out_status,
                      void *workspace, int64_t workspace_bytes, void *stream);

/* The 256-entry quality table int(round((1-10**(q/-10))*255)) of Pileup.calculateQuality
 * (Pileup.py:166-171); evaluated once on the host with the reference's own expression and uploaded. */
int dp_set_quality_lut(const uint8_t lut[256]);

#ifdef __cplusplus
}
#endif
#endif /* DEEPORE_B200_H */
