// Device-side BAM ingest: BGZF blocks -> inflated BAM stream -> structure-of-arrays read batch.
//
// Replaces the host stage of the reference's data access (pysam -> htslib): dep/htslib/bgzf.c:456-499 (inflate_block),
// bgzf.c:93-110 / :610-655 (block header: 18 bytes, BSIZE in the "BC" extra field; trailer CRC32 + ISIZE), and
// dep/htslib/sam.c:380-440 (bam_read1: 32-byte core, name, CIGAR, 4-bit sequence, qualities).  The compressed file is
// what crosses PCIe; the inflated stream and the SoA arrays only ever exist in HBM.
//
// Kernels (first correct version; byte work, HBM/latency-bound, no tensor cores):
//   bgzf_inflate_kernel   one warp per BGZF block (blocks are independent DEFLATE streams of <= 64 KB): the lanes decode
//                         in lock step and share the match / stored-block copies;
//                         Huffman tables + a 9-bit literal/length lookup table per warp in shared memory (inflate_core.h)
//   bam_scan_kernel       one thread: header (contig name -> tid), then the record chain (each record's length prefix
//                         gives the next one); writes record offsets and the running CIGAR / sequence / quality offsets
//   bam_extract_kernel    one warp per kept record: core fields + coalesced copies of CIGAR, packed bases, qualities
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/deepore_b200.h"
#include "common.cuh"
#include "inflate_core.h"

namespace {

constexpr int INFLATE_WARPS = 4;                                     // BGZF blocks per CTA: one per warp
constexpr int INFLATE_STRIDE_U16 = DPI_SCRATCH_U16 + 1;

// One BGZF block per WARP.  A thread per block inside one warp (the first version) diverges on every symbol -- 32
// different streams serialised on one warp scheduler: 0.21 GB/s on the bench file, the speed of ONE host core.  Here the 32
// lanes of a warp run the same decoder in lock step on the same stream (no divergence) and split only the output bytes of
// matches and stored blocks between them (inflate_core.h).
__global__ void __launch_bounds__(INFLATE_WARPS * 32)
bgzf_inflate_kernel(const uint8_t *__restrict__ file, int64_t n_blocks, const int64_t *__restrict__ in_off,
                    const int32_t *__restrict__ in_len, const int64_t *__restrict__ out_off,
                    const int32_t *__restrict__ out_len, uint8_t *out, int32_t *__restrict__ status)
{
    __shared__ uint16_t scratch[INFLATE_WARPS][INFLATE_STRIDE_U16];
    const int warp = threadIdx.x >> 5;
    const int64_t b = (int64_t)blockIdx.x * INFLATE_WARPS + warp;
    if (b >= n_blocks) return;                                      // warp-uniform
    uint32_t produced = 0;
    int rc = dpi_inflate(file + in_off[b], (uint32_t)in_len[b], out + out_off[b], (uint32_t)out_len[b], scratch[warp],
                         &produced);                                // all 32 lanes, in lock step (inflate_core.h)
    if (rc == DPI_OK && produced != (uint32_t)out_len[b]) rc = DPI_E_LENGTH;
    if ((threadIdx.x & 31) == 0) status[b] = rc;
}

__device__ __forceinline__ int32_t ld_i32(const uint8_t *p)
{
    return (int32_t)((uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24));
}

// counts: [0] records kept, [1] CIGAR ops, [2] packed sequence bytes, [3] quality bytes, [4] tid (-1: all / not found),
//         [5] records seen, [6] status (0 ok, -1 bad magic, -2 truncated, -3 contig not found, -4 inconsistent record /
//         header field: what bam_read1 answers with -4, dep/htslib/sam.c:403-406)
__global__ void bam_scan_kernel(const uint8_t *__restrict__ bam, int64_t n_bytes, const char *__restrict__ contig,
                                int32_t contig_len, int64_t max_records, int64_t *__restrict__ rec_off,
                                int64_t *__restrict__ cigar_off, int64_t *__restrict__ seq_off,
                                int64_t *__restrict__ qual_off, int64_t *__restrict__ counts)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    for (int i = 0; i < 7; ++i) counts[i] = 0;
    counts[4] = -1;
    if (n_bytes < 12 || bam[0] != 'B' || bam[1] != 'A' || bam[2] != 'M' || bam[3] != 1) { counts[6] = -1; return; }
    const int32_t l_text = ld_i32(bam + 4);
    if (l_text < 0) { counts[6] = -4; return; }                  // a negative length would walk backwards
    int64_t off = 8 + (int64_t)l_text;
    if (off + 4 > n_bytes) { counts[6] = -2; return; }
    int32_t n_ref = ld_i32(bam + off);
    if (n_ref < 0) { counts[6] = -4; return; }
    off += 4;
    int32_t tid = -1;
    for (int32_t r = 0; r < n_ref; ++r) {
        if (off + 4 > n_bytes) { counts[6] = -2; return; }
        int32_t l_name = ld_i32(bam + off);
        if (l_name < 1 || off + 8 + l_name > n_bytes) { counts[6] = -2; return; }
        if (tid < 0 && contig_len > 0 && l_name - 1 == contig_len) {
            bool same = true;
            for (int32_t k = 0; k < contig_len && same; ++k) same = bam[off + 4 + k] == (uint8_t)contig[k];
            if (same) tid = r;
        }
        off += 8 + l_name;
    }
    if (contig_len > 0 && tid < 0) { counts[6] = -3; return; }
    counts[4] = tid;
    int64_t kept = 0, seen = 0, ncig = 0, nseq = 0, nqual = 0;
    while (off + 36 <= n_bytes) {
        int32_t block_size = ld_i32(bam + off);
        if (block_size < 32 || off + 4 + block_size > n_bytes) { counts[6] = -2; break; }
        ++seen;
        {   // the consistency rule of bam_read1 (dep/htslib/sam.c:397-406, which returns -4): the variable-length parts must
            // fit the record, or dp_bam_extract would read past it and the running offsets would go wrong
            const int32_t l_name = bam[off + 12];
            const int32_t n_cig0 = (int32_t)bam[off + 16] | ((int32_t)bam[off + 17] << 8);
            const int32_t l_seq0 = ld_i32(bam + off + 20);
            if (l_seq0 < 0 || l_name < 1 ||
                32ll + l_name + 4ll * n_cig0 + (((int64_t)l_seq0 + 1) >> 1) + l_seq0 > (int64_t)block_size) { counts[6] = -4; break; }
        }
        if (tid < 0 || ld_i32(bam + off + 4) == tid) {
            int32_t n_cig = (int32_t)bam[off + 16] | ((int32_t)bam[off + 17] << 8);
            int32_t l_seq = ld_i32(bam + off + 20);
            if (kept < max_records) {
                rec_off[kept] = off;
                cigar_off[kept] = ncig;
                seq_off[kept] = nseq;
                qual_off[kept] = nqual;
            }
            ++kept;
            ncig += n_cig;
            nseq += (l_seq + 1) >> 1;
            nqual += l_seq;
        }
        off += 4 + (int64_t)block_size;
    }
    if (counts[6] == 0 && off != n_bytes) counts[6] = -2;
    if (kept <= max_records) {                                   // closing offsets (arrays hold max_records + 1 entries)
        cigar_off[kept] = ncig;
        seq_off[kept] = nseq;
        qual_off[kept] = nqual;
    }
    counts[0] = kept; counts[1] = ncig; counts[2] = nseq; counts[3] = nqual; counts[5] = seen;
}

__global__ void __launch_bounds__(256)
bam_extract_kernel(const uint8_t *__restrict__ bam, int64_t n_records, const int64_t *__restrict__ rec_off,
                   const int64_t *__restrict__ cigar_off, const int64_t *__restrict__ seq_off,
                   const int64_t *__restrict__ qual_off, int32_t *__restrict__ pos, uint16_t *__restrict__ flag,
                   uint8_t *__restrict__ mapq, int32_t *__restrict__ l_seq_out, uint32_t *__restrict__ cigar,
                   uint8_t *__restrict__ seq4, uint8_t *__restrict__ qual)
{
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = (((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); r < n_records; r += warps) {
        const uint8_t *rec = bam + rec_off[r];
        const int32_t l_name = rec[12];
        const int32_t n_cig = (int32_t)rec[16] | ((int32_t)rec[17] << 8);
        const int32_t l_seq = ld_i32(rec + 20);
        if (lane == 0) {
            pos[r] = ld_i32(rec + 8);
            mapq[r] = rec[13];
            flag[r] = (uint16_t)((uint32_t)rec[18] | ((uint32_t)rec[19] << 8));
            l_seq_out[r] = l_seq;
        }
        const uint8_t *p = rec + 36 + l_name;
        uint32_t *co = cigar + cigar_off[r];
        for (int32_t i = lane; i < n_cig; i += 32) co[i] = (uint32_t)ld_i32(p + 4 * i);
        p += 4 * (int64_t)n_cig;
        const int32_t nb = (l_seq + 1) >> 1;
        uint8_t *so = seq4 + seq_off[r];
        for (int32_t i = lane; i < nb; i += 32) so[i] = p[i];
        p += nb;
        uint8_t *qo = qual + qual_off[r];
        for (int32_t i = lane; i < l_seq; i += 32) qo[i] = p[i];
    }
}

}  // namespace

extern "C" {

int64_t dp_bgzf_index(const uint8_t *file, int64_t n_bytes, int64_t max_blocks, int64_t *in_off, int32_t *in_len,
                      int64_t *out_off, int32_t *out_len, uint32_t *crc32)
{
    if (!file || n_bytes < 0) { dp::set_error("dp_bgzf_index: null file"); return DP_EINVAL; }
    int64_t off = 0, n = 0, total = 0;
    while (off < n_bytes) {
        if (off + 18 > n_bytes) { dp::set_error("dp_bgzf_index: truncated block header at byte %lld", (long long)off); return DP_EINVAL; }
        const uint8_t *h = file + off;
        if (h[0] != 31 || h[1] != 139 || h[2] != 8 || !(h[3] & 4)) { dp::set_error("dp_bgzf_index: not a BGZF block at byte %lld", (long long)off); return DP_EINVAL; }
        const int xlen = h[10] | (h[11] << 8);
        if (off + 12 + xlen > n_bytes) { dp::set_error("dp_bgzf_index: truncated extra field at byte %lld", (long long)off); return DP_EINVAL; }
        int bsize = -1;
        for (int x = 0; x + 4 <= xlen;) {                          // sub-fields: SI1 SI2 SLEN(2) data
            const uint8_t *e = h + 12 + x;
            const int slen = e[2] | (e[3] << 8);
            if (e[0] == 'B' && e[1] == 'C' && slen == 2 && x + 6 <= xlen) bsize = (e[4] | (e[5] << 8)) + 1;
            x += 4 + slen;
        }
        if (bsize < 12 + xlen + 8 || off + bsize > n_bytes) { dp::set_error("dp_bgzf_index: bad BSIZE at byte %lld", (long long)off); return DP_EINVAL; }
        const uint8_t *t = file + off + bsize - 8;
        const uint32_t crc = (uint32_t)t[0] | ((uint32_t)t[1] << 8) | ((uint32_t)t[2] << 16) | ((uint32_t)t[3] << 24);
        const uint32_t isize = (uint32_t)t[4] | ((uint32_t)t[5] << 8) | ((uint32_t)t[6] << 16) | ((uint32_t)t[7] << 24);
        if (isize > 65536u) { dp::set_error("dp_bgzf_index: ISIZE %u > 64 KB at byte %lld", isize, (long long)off); return DP_EINVAL; }
        if (max_blocks > 0) {
            if (n >= max_blocks) { dp::set_error("dp_bgzf_index: more than %lld blocks", (long long)max_blocks); return DP_EINVAL; }
            if (in_off) in_off[n] = off + 12 + xlen;
            if (in_len) in_len[n] = bsize - 12 - xlen - 8;
            if (out_off) out_off[n] = total;
            if (out_len) out_len[n] = (int32_t)isize;
            if (crc32) crc32[n] = crc;
        }
        total += isize;
        ++n;
        off += bsize;
    }
    if (max_blocks > 0 && out_off) out_off[n] = total;               // arrays of max_blocks + 1 offsets
    return n;
}

int dp_bgzf_inflate(const uint8_t *file, int64_t n_blocks, const int64_t *in_off, const int32_t *in_len,
                    const int64_t *out_off, const int32_t *out_len, uint8_t *out, int32_t *status, void *stream)
{
    if (n_blocks < 0 || (n_blocks > 0 && (!file || !in_off || !in_len || !out_off || !out_len || !out || !status))) {
        dp::set_error("dp_bgzf_inflate: null argument");
        return DP_EINVAL;
    }
    if (n_blocks == 0) return DP_OK;
    const int64_t grid = (n_blocks + INFLATE_WARPS - 1) / INFLATE_WARPS;
    bgzf_inflate_kernel<<<(unsigned)grid, INFLATE_WARPS * 32, 0, (cudaStream_t)stream>>>(file, n_blocks, in_off, in_len, out_off,
                                                                                         out_len, out, status);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int dp_bam_scan(const uint8_t *bam, int64_t n_bytes, const char *contig, int32_t contig_len, int64_t max_records,
                int64_t *rec_off, int64_t *cigar_off, int64_t *seq_off, int64_t *qual_off, int64_t *counts, void *stream)
{
    if (!bam || !rec_off || !cigar_off || !seq_off || !qual_off || !counts || max_records < 0 || (contig_len > 0 && !contig)) {
        dp::set_error("dp_bam_scan: null argument");
        return DP_EINVAL;
    }
    bam_scan_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(bam, n_bytes, contig, contig_len, max_records, rec_off, cigar_off,
                                                        seq_off, qual_off, counts);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int dp_bam_extract(const uint8_t *bam, int64_t n_records, const int64_t *rec_off, const int64_t *cigar_off,
                   const int64_t *seq_off, const int64_t *qual_off, int32_t *pos, uint16_t *flag, uint8_t *mapq,
                   int32_t *l_seq, uint32_t *cigar, uint8_t *seq4, uint8_t *qual, void *stream)
{
    if (n_records < 0) { dp::set_error("dp_bam_extract: negative record count"); return DP_EINVAL; }
    if (n_records == 0) return DP_OK;
    if (!bam || !rec_off || !cigar_off || !seq_off || !qual_off || !pos || !flag || !mapq || !l_seq) {
        dp::set_error("dp_bam_extract: null argument");
        return DP_EINVAL;
    }
    int64_t ctas = (n_records + 7) / 8;
    if (ctas > 148 * 8) ctas = 148 * 8;
    bam_extract_kernel<<<(unsigned)ctas, 256, 0, (cudaStream_t)stream>>>(bam, n_records, rec_off, cigar_off, seq_off, qual_off,
                                                                         pos, flag, mapq, l_seq, cigar, seq4, qual);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

}  // extern "C"
