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
    int64_t        n_bases;    /* total query bases == qual_off[n_reads] (known to the host reader)     */
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
int64_t dp_encode_workspace_bytes(int32_t n_reads, int64_t n_cigar, int64_t n_bases, int32_t n_windows,
                                  const dp_encode_params *p);
int dp_encode_windows(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs,
                      const int32_t *win_lo, const int32_t *win_hi,
                      const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                      const dp_encode_params *p,
                      uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins, int32_t *out_status,
                      void *workspace, int64_t workspace_bytes, void *stream);

/*
 * Contig-resident form of the same path (the file -> calls pipeline, deepore_b200/pipeline.py): the read arrays of a whole
 * contig stay in HBM, their per-read digests and per-base codes are built ONCE, and the windows are encoded in batches.
 *
 *   dp_encode_prepare    the read-dependent half of dp_encode_windows (per-read CIGAR prefix sums, bam_endpos, checkpoints,
 *                        per-base category/quality codes) into `workspace`, which must hold
 *                        dp_encode_workspace_bytes(n_reads, n_cigar, n_bases, max_windows, p) bytes
 *   dp_window_ranges     replaces the host index of reads.ReadBatch.window_ranges (the `fetch` of Pileup.py:36 /
 *                        dep/htslib/hts.c:2242 as a candidate range): win_lo[w] = number of reads whose running maximum of
 *                        bam_endpos is <= win_qs[w], win_hi[w] = number of reads with pos < win_qs[w] + 2*flank + 1.
 *                        *unsorted is set to 1 when pos[] is not ascending (pysam.fetch needs a coordinate-sorted file).
 *                        Needs a prepared workspace (it reads the bam_endpos values dp_encode_prepare left there).
 *   dp_encode_prepared   the window-dependent half: same arguments and results as dp_encode_windows; n_windows <= max_windows;
 *                        bases_per_window = read bases overlapping an average window (picks the shared-memory build of the
 *                        kernel; <= 0: n_bases / n_windows, right only when `reads` holds just this batch's reads)
 * All pointers are device pointers; nothing synchronises.
 */
int dp_encode_prepare(const dp_reads *reads, const dp_encode_params *p, int32_t max_windows, void *workspace,
                      int64_t workspace_bytes, void *stream);
int dp_window_ranges(const dp_reads *reads, const void *workspace, int64_t workspace_bytes, int32_t max_windows,
                     int32_t n_windows, const int32_t *win_qs, int32_t flank, int32_t *win_lo, int32_t *win_hi,
                     int32_t *unsorted, void *stream);
int dp_encode_prepared(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs, const int32_t *win_lo,
                       const int32_t *win_hi, const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                       const dp_encode_params *p, uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins,
                       int32_t *out_status, void *workspace, int64_t workspace_bytes, int32_t max_windows,
                       int32_t bases_per_window, void *stream);

/*
 * dp_window_layout -- the row layout of whole windows, NOT cut at windowCutoff: out_colrows[w][c] = number of image rows
 * column c of window w produces (0: no entry lands there, else 1 + the longest insertion anchored there,
 * Pileup.py:295-317).  The label of a window is widened by every insert row of the window, also those below the image
 * (Pileup.labelInsertRegion, Pileup.py:309,320-333), so the label code needs this when a window's rows were cut.  Same
 * arguments as dp_encode_windows; out_colrows is a device int32 [n_windows][2*flank+1]; runs the read preparation and the
 * first pass of the exact kernel.
 */
int dp_window_layout(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs, const int32_t *win_lo,
                     const int32_t *win_hi, const uint8_t *ref, int64_t ref_base, int64_t ref_len, const dp_encode_params *p,
                     int32_t *out_colrows, int32_t *out_status, void *workspace, int64_t workspace_bytes, void *stream);

/* Diagnostics: how many windows of the LAST dp_encode_windows call on this workspace left the fast kernel's
 * envelope and were recomputed by the exact general kernel (results are identical either way).  Synchronises
 * the stream. */
int dp_encode_fallback_count(const void *workspace, int32_t n_reads, int64_t n_cigar, int32_t *out_count,
                             void *stream);

/* The 256-entry quality table int(round((1-10**(q/-10))*255)) of Pileup.calculateQuality
 * (Pileup.py:166-171); evaluated once on the host with the reference's own expression and uploaded. */
int dp_set_quality_lut(const uint8_t lut[256]);

/* ================================================================================================
 * Model forward: replaces modules.model.Model.forward (src/modules/model.py:137-156) and the sliding
 * 3-row decode loop of src/test.py:59-89.
 *
 * dp_model_weights carries HOST pointers to fp32 tensors in torch layout:
 *   fold_w / fold_b   the CNN + reinterpreting view + cnn_fc_rnn part of the forward (model.py:138-151) composed
 *                     into ONE affine operator per time step -- legal because the five residual blocks hold only
 *                     convolutions and eval-mode BatchNorm (no activation, model.py:93-105):
 *                       y[t][h] = fold_b[t][h] + sum_{s,w,ch} fold_w[t][h][s][w][ch] * pixel_u8[row+s][w][ch]
 *                     (s = row slot 0..seq_len-1; the 1/255 of ToTensor, dataset.py:52, is folded in).  The host helper
 *                     deepore_b200/fold.py builds it from a Model's state in float64.
 *   gru_*             rnns[k][1].{weight_ih,weight_hh,bias_ih,bias_hh}_l0[_reverse]  (model.py:83-87; gate order r,z,n)
 *   fc1_*             fc1.{weight,bias}  (model.py:90)
 */
#define DP_MAX_GRU_LAYERS 3
typedef struct dp_model_weights {
    int32_t in_ch;        /* 4   */
    int32_t width;        /* coverage_depth = 200 */
    int32_t hidden;       /* 500 */
    int32_t n_class;      /* 3   */
    int32_t seq_len;      /* 3 (test.py:48) */
    int32_t n_layers;     /* 3 bidirectional GRU layers */
    const float *fold_w;  /* [seq_len][hidden][seq_len][width][in_ch] */
    const float *fold_b;  /* [seq_len][hidden] */
    const float *gru_w_ih[DP_MAX_GRU_LAYERS][2];   /* [3*hidden][hidden]; [.][0] forward, [.][1] reverse */
    const float *gru_w_hh[DP_MAX_GRU_LAYERS][2];   /* [3*hidden][hidden] */
    const float *gru_b_ih[DP_MAX_GRU_LAYERS][2];   /* [3*hidden] */
    const float *gru_b_hh[DP_MAX_GRU_LAYERS][2];   /* [3*hidden] */
    const float *fc1_w;   /* [n_class][hidden] */
    const float *fc1_b;   /* [n_class] */
} dp_model_weights;

typedef struct dp_model dp_model;   /* opaque; immutable after creation, shared by all streams of its device */

/* precision: 0 = bf16 operands / fp32 accumulate (tensor cores, one pass); 1 = split-bf16 (hi+lo operands, three
 * passes, fp32-class accuracy).  max_rows = capacity in image rows per forward call (n_windows * windowCutoff). */
int dp_model_create(const dp_model_weights *w, int64_t max_rows, int32_t precision, dp_model **out);
int dp_model_destroy(dp_model *m);

/*
 * dp_model_forward_rows -- every 3-row window of a stack of pileup image rows in one pass.
 *   img      device uint8 [n_rows][width][in_ch]: the encoder's out_img viewed as rows (n_windows*H rows)
 *   logits   optional device fp32 [n_rows][seq_len][n_class]: logits[r] = Model.forward(rows r..r+seq_len-1)
 *            (rows whose window runs past n_rows see zero rows there)
 *   pred     optional device uint8 [n_rows][seq_len]: argmax over classes (torch.topk(1), first maximum)
 */
int dp_model_forward_rows(dp_model *m, const uint8_t *img, int64_t n_rows, float *logits, uint8_t *pred, void *stream);

/*
 * dp_model_forward_windows -- dp_model_forward_rows for a stack of n_windows WHOLE pileups of H rows each: the calls of
 * src/test.py:59-66 (`for row in range(0, H - seq_len + 1)`: rows 0 .. H-seq_len of every pileup) and nothing else.  The
 * calls at the last seq_len-1 rows of a pileup, whose 3-row window would run into the next pileup and which the reference
 * never makes, are not computed (1.35 % of every GRU launch at H = 150): their entries of logits / pred are left
 * untouched; all other entries equal dp_model_forward_rows' bit for bit.  dp_vote reads none of the skipped entries.
 *   img      device uint8 [n_windows][H][width][in_ch]
 *   logits   optional device fp32 [n_windows * H][seq_len][n_class];  pred optional device uint8 [n_windows * H][seq_len]
 */
int dp_model_forward_windows(dp_model *m, const uint8_t *img, int64_t n_windows, int32_t H, float *logits, uint8_t *pred,
                             void *stream);

/*
 * dp_model_forward_calls -- Model.forward(x, hidden) as src/test.py:63-66 calls it: n_calls INDEPENDENT windows of seq_len
 * rows each (one per sample of the batch), no sliding.
 *   img      device uint8 [n_calls][seq_len][width][in_ch]   (n_calls * seq_len <= max_rows)
 *   logits   optional device fp32 [n_calls][seq_len][n_class] = Model.forward's return value (model.py:154-156)
 *   pred     optional device uint8 [n_calls][seq_len]
 */
int dp_model_forward_calls(dp_model *m, const uint8_t *img, int64_t n_calls, float *logits, uint8_t *pred, void *stream);

/*
 * dp_vote -- test.py:71-89: calls[w][row] = mode(pred[w][row-2][2], pred[w][row-1][1], pred[w][row][0]) for
 * 2 <= row <= H-3 (smallest class on a three-way tie), -1 elsewhere; optional confusion[label][call] += 1.
 *   pred [n_windows][H][3] uint8 (dp_model_forward_rows with n_rows = n_windows*H), calls [n_windows][H] int8,
 *   labels optional [n_windows][H] uint8, confusion optional [n_class*n_class] uint64 (accumulated, not cleared)
 */
int dp_vote(const uint8_t *pred, int32_t n_windows, int32_t H, int32_t n_class, int8_t *calls, const uint8_t *labels,
            unsigned long long *confusion, void *stream);

/*
 * dp_cross_entropy -- the loss of the reference's validation pass (src/train.py:21-78: forward, then
 * nn.CrossEntropyLoss on outputs.view(-1, 3) against the labels of the call's rows, :36,75), element-wise:
 * loss[i] = logsumexp(logits[i][:]) - logits[i][labels[i]] for n (call, t) elements; the mean over a batch is the caller's.
 *   logits device fp32 [n][n_class]; labels device uint8 [n]; loss optional device fp32 [n]; sum optional device double
 *   (accumulated, not cleared); bad_labels optional device int32 counter of labels >= n_class (their loss is +inf and is
 *   left out of the sum; torch raises on them).  Training itself (backward, SGD: train.py:101-203) is not part of this library.
 */
int dp_cross_entropy(const float *logits, const uint8_t *labels, int64_t n, int32_t n_class, float *loss, double *sum,
                     int32_t *bad_labels, void *stream);

/* Debug/inspection: copy an internal activation of the last forward to the host as fp32 [n_rows][512].
 * what: 0..2 = y[t] (folded CNN+Linear output); 100*k+10*d+t = GRU layer k hidden state, d=0 forward / 1 reverse. */
int dp_model_peek(dp_model *m, int32_t what, int64_t n_rows, float *host_out);

/* Per-kernel device timing for the roofline report (CUDA events on the launching stream around each launch of
 * dp_model_forward_rows while enabled).  dp_model_profile_read synchronises the device and returns accumulated
 * milliseconds and launch counts for: [0] u8->bf16 convert, [1] folded CNN+Linear GEMM, [2] GRU steps, [3] fc1+argmax,
 * then clears the accumulators. */
int dp_model_profile(dp_model *m, int32_t enable);
int dp_model_profile_read(dp_model *m, float ms[4], int32_t launches[4]);

/* Plain dense GEMM on the same tcgen05 kernel (parity tests): out[M][N] = A[M][K] . B[N][K]^T + bias, device bf16
 * row-major operands (lda, ldb in elements, multiples of 8), out bf16 or fp32 (ldc), N multiple of 16. */
int dp_gemm_bf16(const void *A, int64_t lda, const void *B, int64_t ldb, const float *bias, void *out, int64_t ldc,
                 int32_t out_fp32, int64_t M, int64_t N, int64_t K, int32_t block_n, void *stream);

/* ---- device-side BAM ingest: BGZF inflate + BAM record -> dp_reads arrays in HBM.
 * Replaces the host stage behind pysam: dep/htslib/bgzf.c:456-499 (inflate_block, zlib raw inflate of one block),
 * bgzf.c block header / trailer (18-byte header with the "BC" BSIZE sub-field; CRC32 + ISIZE), and
 * dep/htslib/sam.c bam_read1 (32-byte core, name, CIGAR, 4-bit bases, qualities). */

/* dp_bgzf_index -- HOST: walks the block headers of a BGZF file image.  Returns the number of blocks (the EOF block
 * included) or a negative DP_E* code.  With max_blocks == 0 it only counts; otherwise fills, per block, the byte
 * offset / length of the raw DEFLATE payload, the offset / length (ISIZE) of its inflated bytes (out_off has
 * n_blocks + 1 entries, the last one = total inflated size) and the trailer CRC32.  Any out pointer may be NULL. */
int64_t dp_bgzf_index(const uint8_t *file, int64_t n_bytes, int64_t max_blocks, int64_t *in_off, int32_t *in_len,
                      int64_t *out_off, int32_t *out_len, uint32_t *crc32);

/* dp_bgzf_inflate -- DEVICE pointers throughout: inflates every block of the (device-resident) file image into
 * out[out_off[b] .. +out_len[b]).  status[b] = 0 or a negative DPI_E_* code (csrc/inflate_core.h); malformed blocks
 * never read or write out of bounds. */
int dp_bgzf_inflate(const uint8_t *file, int64_t n_blocks, const int64_t *in_off, const int32_t *in_len,
                    const int64_t *out_off, const int32_t *out_len, uint8_t *out, int32_t *status, void *stream);

/* dp_bam_scan -- DEVICE pointers: parses the BAM header of the inflated stream, resolves `contig` (contig_len bytes,
 * device memory; contig_len == 0 keeps every record) to its tid and walks the record chain.  For every kept record i
 * (< max_records) writes rec_off[i] (byte offset of its length prefix) and the running cigar_off / seq_off / qual_off
 * of dp_reads (arrays of max_records + 1 entries; entry [kept] closes them).  counts (7 x int64): records kept, CIGAR
 * ops, packed sequence bytes, quality bytes, tid, records seen, status (0 ok, -1 bad magic, -2 truncated, -3 contig
 * not in the header, -4 a record whose l_seq / l_read_name / n_cigar_op do not fit its block_size or a negative header
 * length -- the cases bam_read1 rejects with -4, dep/htslib/sam.c:397-406; nothing is extracted from such a stream). */
int dp_bam_scan(const uint8_t *bam, int64_t n_bytes, const char *contig, int32_t contig_len, int64_t max_records,
                int64_t *rec_off, int64_t *cigar_off, int64_t *seq_off, int64_t *qual_off, int64_t *counts, void *stream);

/* dp_bam_scan_static -- dp_bam_scan (same arguments, same results) for callers that must not allocate: code under CUDA
 * graph capture, serving loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
 * The segment scratch is a file-scope device array (98 384 bytes per GPU that loads the library,
 * held until exit) instead of stream-ordered pool memory; the call allocates and frees nothing.  n_bytes above
 * DP_BAM_SCAN_STATIC_MAX_BYTES is rejected with DP_EINVAL (use dp_bam_scan).  All calls on one device share the array:
 * calls that overlap on the GPU (different streams without ordering) race -- order them on one stream or with events. */
#define DP_BAM_SCAN_STATIC_MAX_BYTES ((int64_t)1 << 30)
int dp_bam_scan_static(const uint8_t *bam, int64_t n_bytes, const char *contig, int32_t contig_len, int64_t max_records,
                       int64_t *rec_off, int64_t *cigar_off, int64_t *seq_off, int64_t *qual_off, int64_t *counts, void *stream);

/* dp_bam_extract -- DEVICE pointers: fills the dp_reads arrays of the n_records records found by dp_bam_scan. */
int dp_bam_extract(const uint8_t *bam, int64_t n_records, const int64_t *rec_off, const int64_t *cigar_off,
                   const int64_t *seq_off, const int64_t *qual_off, int32_t *pos, uint16_t *flag, uint8_t *mapq,
                   int32_t *l_seq, uint32_t *cigar, uint8_t *seq4, uint8_t *qual, void *stream);

/* ---- PNG files of pileup tensors, written on the device.  Replaces the per-window host work of
 * Pileup.savePileupRGB (src/utils/Pileup.py:336-361: PIL image, 30 000 pixel stores, PNG save): out + i * out_stride
 * receives the complete PNG byte stream (8-bit RGBA, W x H, filter 0, one IDAT of stored deflate blocks, Adler-32, chunk
 * CRC-32s) of image i; every file is dp_png_size(H, W) bytes long.  DEVICE pointers; img = uint8 [n_images][H][W][4] as
 * dp_encode_windows writes it. */
int64_t dp_png_size(int32_t H, int32_t W);
int dp_png_encode(const uint8_t *img, int32_t n_images, int32_t H, int32_t W, uint8_t *out, int64_t out_stride, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* DEEPORE_B200_H */
