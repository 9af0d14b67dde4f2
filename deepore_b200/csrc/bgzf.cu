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
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/deepore_b200.h"
#include "common.cuh"
#include "inflate_core.h"

namespace {

#ifndef INFLATE_WARPS_PER_CTA
#define INFLATE_WARPS_PER_CTA 4
#endif
constexpr int INFLATE_WARPS = INFLATE_WARPS_PER_CTA;                                     // BGZF blocks per CTA: one per warp
constexpr int INFLATE_STRIDE_U16 = (DPI_SCRATCH_U16 + 2 + 7) & ~7;   // per-warp scratch stays 16-byte aligned, banks staggered by the +2

// One BGZF block per WARP.  A thread per block inside one warp (the first version) diverges on every symbol -- 32
// different streams serialised on one warp scheduler: 0.21 GB/s on the bench file, the speed of ONE host core.  Here the 32
// lanes of a warp run the same decoder in lock step on the same stream (no divergence) and split only the output bytes of
// matches and stored blocks between them (inflate_core.h).
// no launch bound: ptxas settles on 48 registers by itself; a bound that only RESTATES 48 (128 threads, 10 CTAs) made it
// recompute the stream's word base at every window slide (+25 % time), and a looser one takes 62 registers
#ifdef INFLATE_MIN_CTAS
__global__ void __launch_bounds__(INFLATE_WARPS * 32, INFLATE_MIN_CTAS)
#else
__global__ void
#endif
bgzf_inflate_kernel(const uint8_t *__restrict__ file, int64_t n_blocks, const int64_t *__restrict__ in_off,
                    const int32_t *__restrict__ in_len, const int64_t *__restrict__ out_off,
                    const int32_t *__restrict__ out_len, uint8_t *out, int32_t *__restrict__ status)
{
    __shared__ __align__(16) uint16_t scratch[INFLATE_WARPS][INFLATE_STRIDE_U16];
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
                                int64_t *__restrict__ qual_off, int64_t *__restrict__ counts, const int *__restrict__ gate)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    if (gate && *gate == 0) return;                              // the parallel scan below already produced the result
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

// ---- parallel scan.  The record chain is sequential (each length prefix gives the next record): one thread chasing it
// pays a DRAM round trip per record (12.4 ms for the 17 650 records of a 355 MB contig).  Here the stream is cut into
// segments; a warp per segment GUESSES the first record start inside it (the first offset whose 36 leading bytes look like
// a record -- htslib's own consistency rule, dep/htslib/sam.c:397-406, plus a NUL-terminated name and reference ids inside
// the header's range -- and from which the chain stays consistent up to the segment's end) and walks the segment from
// there.  The guesses are then PROVEN: starting from the true first record behind the header, the exit offset of every
// segment on the chain must be exactly the guessed entry of the segment it lands in.  If any link fails (or the stream is
// malformed) the serial kernel above runs instead, so the result is exact either way; on htslib-written files every link
// holds.
struct ScanSeg {
    int64_t entry;                 // first record start inside the segment (-1: none found)
    int64_t exit;                  // first record start at or behind the segment's end
    int64_t kept, ncig, nseq, nqual;
};
constexpr int SCAN_MAXSEG = 2048;
constexpr int SCAN_TRIES = 48;     // candidates walked per segment before giving up

__device__ __forceinline__ bool bam_record_plausible(const uint8_t *bam, int64_t n_bytes, int64_t o, int32_t n_ref)
{
    if (o + 36 > n_bytes) return false;
    // cheapest, most selective test first: the lanes of a warp test 32 offsets in lock step, so a round is only as short as
    // its slowest lane -- random bytes pass "reference id inside the header's range" with probability ~n_ref / 2^32, but
    // "32 <= block_size <= bytes left" with several per cent
    {
        const int32_t tid0 = ld_i32(bam + o + 4);
        if (tid0 < -1 || tid0 >= n_ref) return false;
    }
    const int32_t bs = ld_i32(bam + o);
    if (bs < 32 || o + 4 + (int64_t)bs > n_bytes) return false;
    const int32_t tid = ld_i32(bam + o + 4), pos = ld_i32(bam + o + 8);
    const int32_t l_name = bam[o + 12];
    const int32_t n_cig = (int32_t)bam[o + 16] | ((int32_t)bam[o + 17] << 8);
    const int32_t l_seq = ld_i32(bam + o + 20);
    const int32_t mtid = ld_i32(bam + o + 24), mpos = ld_i32(bam + o + 28);
    if (tid < -1 || tid >= n_ref || mtid < -1 || mtid >= n_ref || pos < -1 || mpos < -1) return false;
    if (l_seq < 0 || l_name < 1) return false;
    if (32ll + l_name + 4ll * n_cig + (((int64_t)l_seq + 1) >> 1) + l_seq > (int64_t)bs) return false;
    return bam[o + 36 + l_name - 1] == 0;                        // read names are NUL-terminated
}

// hdr: [0] offset of the first record, [1] tid of the wanted contig (-1: all), [2] n_ref, [3] status (0 ok)
__global__ void bam_header_kernel(const uint8_t *__restrict__ bam, int64_t n_bytes, const char *__restrict__ contig,
                                  int32_t contig_len, int64_t *__restrict__ hdr)
{
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    hdr[0] = 0; hdr[1] = -1; hdr[2] = 0; hdr[3] = 0;
    if (n_bytes < 12 || bam[0] != 'B' || bam[1] != 'A' || bam[2] != 'M' || bam[3] != 1) { hdr[3] = -1; return; }
    const int32_t l_text = ld_i32(bam + 4);
    if (l_text < 0) { hdr[3] = -4; return; }
    int64_t off = 8 + (int64_t)l_text;
    if (off + 4 > n_bytes) { hdr[3] = -2; return; }
    const int32_t n_ref = ld_i32(bam + off);
    if (n_ref < 0) { hdr[3] = -4; return; }
    off += 4;
    int32_t tid = -1;
    for (int32_t r = 0; r < n_ref; ++r) {
        if (off + 4 > n_bytes) { hdr[3] = -2; return; }
        const int32_t l_name = ld_i32(bam + off);
        if (l_name < 1 || off + 8 + l_name > n_bytes) { hdr[3] = -2; return; }
        if (tid < 0 && contig_len > 0 && l_name - 1 == contig_len) {
            bool same = true;
            for (int32_t k = 0; k < contig_len && same; ++k) same = bam[off + 4 + k] == (uint8_t)contig[k];
            if (same) tid = r;
        }
        off += 8 + l_name;
    }
    if (contig_len > 0 && tid < 0) { hdr[3] = -3; return; }
    hdr[0] = off; hdr[1] = tid; hdr[2] = n_ref;
}

// walks the chain from `o` while it starts inside [.., seg_end); false if a record fails the consistency rule
__device__ __forceinline__ bool bam_walk_segment(const uint8_t *bam, int64_t n_bytes, int64_t o, int64_t seg_end, int32_t n_ref,
                                                 int32_t tid, ScanSeg &g)
{
    g.kept = g.ncig = g.nseq = g.nqual = 0;
    while (o < seg_end) {
        if (o == n_bytes) break;
        if (!bam_record_plausible(bam, n_bytes, o, n_ref)) return false;
        if (tid < 0 || ld_i32(bam + o + 4) == tid) {
            const int32_t n_cig = (int32_t)bam[o + 16] | ((int32_t)bam[o + 17] << 8);
            const int32_t l_seq = ld_i32(bam + o + 20);
            g.kept++; g.ncig += n_cig; g.nseq += (l_seq + 1) >> 1; g.nqual += l_seq;
        }
        o += 4 + (int64_t)ld_i32(bam + o);
    }
    g.exit = o;
    return true;
}

__global__ void __launch_bounds__(128)
bam_seg_kernel(const uint8_t *__restrict__ bam, int64_t n_bytes, int64_t seg_bytes, int n_seg, const int64_t *__restrict__ hdr,
               ScanSeg *__restrict__ segs)
{
    const int lane = threadIdx.x & 31;
    const int s = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (s >= n_seg || hdr[3] != 0) return;
    const int64_t first = hdr[0];
    const int32_t tid = (int32_t)hdr[1], n_ref = (int32_t)hdr[2];
    const int64_t b = (int64_t)s * seg_bytes, e = min(b + seg_bytes, n_bytes);
    ScanSeg g;
    g.entry = -1; g.exit = -1; g.kept = g.ncig = g.nseq = g.nqual = 0;
    if (e > first) {
        if (b <= first) {                                        // the segment holding the true first record: no guess needed
            if (bam_walk_segment(bam, n_bytes, first, e, n_ref, tid, g)) g.entry = first;
        } else {
            int tries = 0;
            // 128 offsets per iteration (four independent tests per lane: the loop is bound by load latency, not by issue)
            for (int64_t base = b; base < e && g.entry < 0 && tries < SCAN_TRIES; base += 128) {
                unsigned m4[4];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const int64_t o = base + 32 * u + lane;
                    m4[u] = __ballot_sync(0xffffffffu, o < e && bam_record_plausible(bam, n_bytes, o, n_ref));
                }
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    unsigned m = m4[u];
                    while (m && g.entry < 0 && tries < SCAN_TRIES) {
                        const int k = __ffs(m) - 1;
                        m &= m - 1;
                        tries++;
                        if (bam_walk_segment(bam, n_bytes, base + 32 * u + k, e, n_ref, tid, g)) g.entry = base + 32 * u + k;   // all lanes, same loads
                    }
                }
            }
        }
    }
    if (lane == 0) segs[s] = g;
}

// proof + prefix sums.  link[0] = 1 when the serial kernel has to run.  counts as bam_scan_kernel writes them.
__global__ void __launch_bounds__(1024)
bam_link_kernel(int64_t n_bytes, int64_t seg_bytes, int n_seg, const int64_t *__restrict__ hdr, ScanSeg *__restrict__ segs,
                int64_t max_records, int64_t *__restrict__ counts, int *__restrict__ link, int64_t *__restrict__ cigar_off,
                int64_t *__restrict__ seq_off, int64_t *__restrict__ qual_off)
{
    __shared__ int64_t s_entry[SCAN_MAXSEG], s_exit[SCAN_MAXSEG];
    __shared__ uint8_t s_on[SCAN_MAXSEG];
    __shared__ int64_t s_w[32][4];
    __shared__ int64_t s_carry[4];
    __shared__ int s_ok;
    const int tid_ = threadIdx.x, lane = tid_ & 31, warp = tid_ >> 5;
    if (hdr[3] != 0) {                                           // header errors: reported here, nothing else runs
        if (tid_ == 0) { for (int i = 0; i < 7; ++i) counts[i] = 0; counts[4] = -1; counts[6] = hdr[3]; link[0] = 0; }
        return;
    }
    for (int i = tid_; i < n_seg; i += blockDim.x) { s_entry[i] = segs[i].entry; s_exit[i] = segs[i].exit; s_on[i] = 0; }
    __syncthreads();
    if (tid_ == 0) {
        int64_t cur = hdr[0];
        bool ok = true;
        while (cur < n_bytes) {
            const int s = (int)(cur / seg_bytes);
            if (s >= n_seg || s_entry[s] != cur) { ok = false; break; }
            s_on[s] = 1;
            const int64_t nx = s_exit[s];                       // shared memory: the chase is a chain of dependent reads
            if (nx <= cur) { ok = false; break; }
            cur = nx;
        }
        if (cur != n_bytes) ok = false;                          // a truncated tail is the serial kernel's to report
        s_ok = ok ? 1 : 0;
        for (int k = 0; k < 4; k++) s_carry[k] = 0;
    }
    __syncthreads();
    if (!s_ok) { if (tid_ == 0) link[0] = 1; return; }
    // exclusive prefix sums of (records kept, CIGAR ops, sequence bytes, quality bytes) over the segments on the chain
    for (int base = 0; base < n_seg; base += blockDim.x) {
        const int i = base + tid_;
        int64_t v[4] = {0, 0, 0, 0};
        if (i < n_seg && s_on[i]) { v[0] = segs[i].kept; v[1] = segs[i].ncig; v[2] = segs[i].nseq; v[3] = segs[i].nqual; }
        int64_t inc[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            inc[k] = v[k];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int64_t t = __shfl_up_sync(0xffffffffu, inc[k], d);
                if (lane >= d) inc[k] += t;
            }
            if (lane == 31) s_w[warp][k] = inc[k];
        }
        __syncthreads();
        int64_t before[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            before[k] = s_carry[k];
            for (int w = 0; w < warp; w++) before[k] += s_w[w][k];
        }
        if (i < n_seg) {
            ScanSeg g = segs[i];
            g.kept = before[0] + inc[0] - v[0]; g.ncig = before[1] + inc[1] - v[1];
            g.nseq = before[2] + inc[2] - v[2]; g.nqual = before[3] + inc[3] - v[3];
            if (!s_on[i]) g.entry = -1;                          // a guess inside a record the chain jumps over: not a record
            segs[i] = g;
        }
        __syncthreads();
        if (tid_ == blockDim.x - 1) {
#pragma unroll
            for (int k = 0; k < 4; k++) s_carry[k] = before[k] + inc[k];
        }
        __syncthreads();
    }
    if (tid_ == 0) {
        const int64_t kept = s_carry[0];
        counts[0] = kept; counts[1] = s_carry[1]; counts[2] = s_carry[2]; counts[3] = s_carry[3];
        counts[4] = hdr[1]; counts[5] = 0; counts[6] = 0;
        if (kept <= max_records) { cigar_off[kept] = s_carry[1]; seq_off[kept] = s_carry[2]; qual_off[kept] = s_carry[3]; }
        link[0] = 0;
    }
}

// second walk of every segment on the chain: record offsets and running array offsets at their final places
__global__ void __launch_bounds__(128)
bam_emit_kernel(const uint8_t *__restrict__ bam, int64_t n_bytes, int64_t seg_bytes, int n_seg, const int64_t *__restrict__ hdr,
                const ScanSeg *__restrict__ segs, const int *__restrict__ link, int64_t max_records, int64_t *__restrict__ rec_off,
                int64_t *__restrict__ cigar_off, int64_t *__restrict__ seq_off, int64_t *__restrict__ qual_off,
                int64_t *__restrict__ counts)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg || hdr[3] != 0 || link[0] != 0) return;
    const ScanSeg g = segs[s];
    if (g.entry < 0) return;
    const int32_t tid = (int32_t)hdr[1];
    const int64_t e = min((int64_t)(s + 1) * seg_bytes, n_bytes);
    int64_t o = g.entry, k = g.kept, ncig = g.ncig, nseq = g.nseq, nqual = g.nqual, seen = 0;
    while (o < e && o < n_bytes) {
        ++seen;
        if (tid < 0 || ld_i32(bam + o + 4) == tid) {
            const int32_t n_cig = (int32_t)bam[o + 16] | ((int32_t)bam[o + 17] << 8);
            const int32_t l_seq = ld_i32(bam + o + 20);
            if (k < max_records) { rec_off[k] = o; cigar_off[k] = ncig; seq_off[k] = nseq; qual_off[k] = nqual; }
            ++k; ncig += n_cig; nseq += (l_seq + 1) >> 1; nqual += l_seq;
        }
        o += 4 + (int64_t)ld_i32(bam + o);
    }
    atomicAdd(reinterpret_cast<unsigned long long *>(counts + 5), (unsigned long long)seen);
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
    cudaStream_t st = (cudaStream_t)stream;
    const char *e = getenv("DP_BAM_SCAN");                       // "serial": the one-thread chain walk only
    if (e && !strcmp(e, "serial")) {
        bam_scan_kernel<<<1, 32, 0, st>>>(bam, n_bytes, contig, contig_len, max_records, rec_off, cigar_off, seq_off, qual_off,
                                          counts, nullptr);
        DP_CUDA_CHECK(cudaGetLastError());
        return DP_OK;
    }
    // segments of >= 64 KB, at most SCAN_MAXSEG of them (the link kernel keeps their entries in shared memory)
    int64_t seg_bytes = 65536;
    while ((n_bytes + seg_bytes - 1) / seg_bytes > SCAN_MAXSEG) seg_bytes *= 2;
    const int n_seg = (int)((n_bytes + seg_bytes - 1) / seg_bytes) > 0 ? (int)((n_bytes + seg_bytes - 1) / seg_bytes) : 1;
    uint8_t *scratch = nullptr;
    const size_t bytes = 64 + 16 + sizeof(ScanSeg) * (size_t)n_seg;
    {   // stream-ordered scratch (~100 KB): keep freed blocks in the device's default pool instead of returning them to the
        // driver at every synchronisation (the default release threshold is 0, which turns each call into a cudaMalloc)
        static bool pool_set[64] = {false};
        int dev = 0;
        DP_CUDA_CHECK(cudaGetDevice(&dev));
        if (dev < 64 && !pool_set[dev]) {
            cudaMemPool_t pool;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
                uint64_t keep = 64ull << 20;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            cudaGetLastError();
            pool_set[dev] = true;
        }
    }
    DP_CUDA_CHECK(cudaMallocAsync((void **)&scratch, bytes, st));
    int64_t *hdr = (int64_t *)scratch;
    int *link = (int *)(scratch + 64);
    ScanSeg *segs = (ScanSeg *)(scratch + 80);
    bam_header_kernel<<<1, 32, 0, st>>>(bam, n_bytes, contig, contig_len, hdr);
    bam_seg_kernel<<<(n_seg + 3) / 4, 128, 0, st>>>(bam, n_bytes, seg_bytes, n_seg, hdr, segs);
    bam_link_kernel<<<1, 1024, 0, st>>>(n_bytes, seg_bytes, n_seg, hdr, segs, max_records, counts, link, cigar_off, seq_off, qual_off);
    bam_emit_kernel<<<(n_seg + 127) / 128, 128, 0, st>>>(bam, n_bytes, seg_bytes, n_seg, hdr, segs, link, max_records, rec_off,
                                                         cigar_off, seq_off, qual_off, counts);
    bam_scan_kernel<<<1, 32, 0, st>>>(bam, n_bytes, contig, contig_len, max_records, rec_off, cigar_off, seq_off, qual_off, counts,
                                      link);                     // runs only when a link of the proof failed
    cudaError_t le = cudaGetLastError();
    cudaFreeAsync(scratch, st);
    if (le != cudaSuccess) { dp::set_error("dp_bam_scan: %s", cudaGetErrorString(le)); return DP_ECUDA; }
    return DP_OK;
}

// dp_bam_scan for callers that must not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc):
// the same five launches, with the segment scratch in a file-scope device array instead of cudaMallocAsync memory.
// The array is sized for the scan's own ceiling of SCAN_MAXSEG segments (98 384 bytes; every GPU that loads the library
// holds it until exit), and the input is capped at DP_BAM_SCAN_STATIC_MAX_BYTES so that a segment stays <= 512 KB.
// ALL calls on a device share the array: two calls that overlap on the GPU (different streams, no ordering) race.
// Nothing in it has to start at zero -- the header kernel writes hdr, the segment kernel every segment it is asked for,
// the link kernel link[0] -- the header words are cleared anyway so that a call never sees the previous call's.
static __device__ __align__(256) uint8_t g_scan_static[80 + sizeof(ScanSeg) * SCAN_MAXSEG];

static int scan_static_address(uint8_t **out)
{
    // the array's name is a host-side stand-in here, not a device address: resolved once per device, that device current
    static std::mutex mu;
    static uint8_t *ptr[64] = {nullptr};
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { dp::set_error("dp_bam_scan_static: device ordinal %d out of range", dev); return DP_EINVAL; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        uint8_t *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_scan_static);
        if (e != cudaSuccess || !p) {
            dp::set_error("dp_bam_scan_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return DP_ECUDA;
        }
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return DP_OK;
}

int dp_bam_scan_static(const uint8_t *bam, int64_t n_bytes, const char *contig, int32_t contig_len, int64_t max_records,
                       int64_t *rec_off, int64_t *cigar_off, int64_t *seq_off, int64_t *qual_off, int64_t *counts, void *stream)
{
    if (!bam || !rec_off || !cigar_off || !seq_off || !qual_off || !counts || max_records < 0 || (contig_len > 0 && !contig)) {
        dp::set_error("dp_bam_scan_static: null argument");
        return DP_EINVAL;
    }
    if (n_bytes < 0 || n_bytes > DP_BAM_SCAN_STATIC_MAX_BYTES) {
        dp::set_error("dp_bam_scan_static: %lld bytes exceed the static scratch's limit of %lld (use dp_bam_scan)",
                      (long long)n_bytes, (long long)DP_BAM_SCAN_STATIC_MAX_BYTES);
        return DP_EINVAL;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int64_t seg_bytes = 65536;
    while ((n_bytes + seg_bytes - 1) / seg_bytes > SCAN_MAXSEG) seg_bytes *= 2;
    const int n_seg = (int)((n_bytes + seg_bytes - 1) / seg_bytes) > 0 ? (int)((n_bytes + seg_bytes - 1) / seg_bytes) : 1;
    uint8_t *scratch = nullptr;
    int rc = scan_static_address(&scratch);
    if (rc != DP_OK) return rc;
    DP_CUDA_CHECK(cudaMemsetAsync(scratch, 0, 80, st));
    int64_t *hdr = (int64_t *)scratch;
    int *link = (int *)(scratch + 64);
    ScanSeg *segs = (ScanSeg *)(scratch + 80);
    bam_header_kernel<<<1, 32, 0, st>>>(bam, n_bytes, contig, contig_len, hdr);
    bam_seg_kernel<<<(n_seg + 3) / 4, 128, 0, st>>>(bam, n_bytes, seg_bytes, n_seg, hdr, segs);
    bam_link_kernel<<<1, 1024, 0, st>>>(n_bytes, seg_bytes, n_seg, hdr, segs, max_records, counts, link, cigar_off, seq_off, qual_off);
    bam_emit_kernel<<<(n_seg + 127) / 128, 128, 0, st>>>(bam, n_bytes, seg_bytes, n_seg, hdr, segs, link, max_records, rec_off,
                                                         cigar_off, seq_off, qual_off, counts);
    bam_scan_kernel<<<1, 32, 0, st>>>(bam, n_bytes, contig, contig_len, max_records, rec_off, cigar_off, seq_off, qual_off, counts,
                                      link);                     // runs only when a link of the proof failed
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
