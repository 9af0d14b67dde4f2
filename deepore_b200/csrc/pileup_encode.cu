// deepore_b200 pileup encoder -- hand-written CUDA for sm_100a.
//
// Replaces /root/reference/src/utils/Pileup.py : Pileup.__init__ (:21-93), iterateReads (:152-163),
// parseRead (:174-264), generatePileupImage/encodeColumn (:268-317) and savePileupRGB (:336-361) for a
// batch of windows, bit-exactly (tests/test_encoder_gpu.py).
//
// Formulation (DESIGN.md "encoder"): parseRead's state machine is closed-form per CIGAR op -- op j that
// starts at reference position A_j with length n emits its step i at column A_j (i==0) or
// A_j + dRef*(i-1) (i>=1), reading query index qp_j + dRead*i (qp lags by one: it starts at -1) and
// reference character ref[rel], rel = rel0 + number of earlier in-window M/D steps of the read.  The
// block-skip test of :214 never changes what is emitted, so every op is handled by the same rule.
// That turns the scatter into an ORDERED GATHER: one CTA per window; reads are visited in file order;
// one thread owns one pileup column (pass 0) / one image row (passes 1-2) and pulls its own entries out
// of each read, so per-(row,category) order is (read order, step order) with no atomics and no sort:
//
//   prep kernel   per read: first aligned position, bam_endpos, soft-clip extent; per op: exclusive
//                 prefix sums of reference / query advance (lets a window enter a long read in O(log ops))
//   pass 0        per column: populated?  longest insert anchored here  -> row map (sparse column axis,
//                 insert rows spliced after their anchor, Pileup.py:295-317)
//   pass 1        per image row: entry count per category (A,C,G,T,D,M,N / insert rows A,C,G,T,N)
//   pass 2        per image row: place each quality byte at base[category] + running rank (first W-1 kept)
//   emit          RGBA rows written with 16-byte stores: colour from the category boundaries, alpha from the
//                 staged quality bytes, filler elsewhere.  120 000 B per default window; this is the
//                 HBM-bound part.
#include <limits.h>
#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

namespace dp {

static thread_local char g_err[512];
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// Pileup.calculateQuality (Pileup.py:166-171): int(round((1-10**(q/-10))*255)), q = min(baseQ, mapQ).
// Table generated with the reference's own Python expression (tests/test_oracle_pins.py checks it);
// every q >= 28 maps to 255.
__constant__ uint8_t c_qlut[256];
static const uint8_t h_qlut_head[28] = {0,   52,  94,  127, 153, 174, 191, 204, 215, 223, 230, 235, 239, 242,
                                        245, 247, 249, 250, 251, 252, 252, 253, 253, 254, 254, 254, 254, 254};
static bool g_lut_ready[64] = {false};
static bool g_lut_flat = true;      // table constant from index 28 on (true for the reference's formula): v3 codes need it

static int ensure_lut() {
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 64 && g_lut_ready[dev]) return DP_OK;
    uint8_t lut[256];
    for (int i = 0; i < 256; i++) lut[i] = i < 28 ? h_qlut_head[i] : 255;
    DP_CUDA_CHECK(cudaMemcpyToSymbol(c_qlut, lut, 256));
    if (dev < 64) g_lut_ready[dev] = true;
    return DP_OK;
}

// ------------------------------------------------------------------------------------------------
// categories in sortingKey order (Pileup.py:66-73): A C G T D M N ; inserts use A C G T N (0,1,2,3,6)
enum { CAT_A = 0, CAT_C = 1, CAT_G = 2, CAT_T = 3, CAT_D = 4, CAT_M = 5, CAT_N = 6, NCAT = 7 };

// RGBA little-endian words with alpha 0 (SNPtoRGB, Pileup.py:56-64)
__device__ __forceinline__ uint32_t cat_rgb(int cat) {
    // A (255,0,0) C (255,255,0) G (0,255,0) T (0,0,255) D (0,255,255) M (255,255,255) N (0,0,0)
    const uint32_t t[NCAT] = {0x000000FFu, 0x0000FFFFu, 0x0000FF00u, 0x00FF0000u, 0x00FFFF00u, 0x00FFFFFFu, 0x00000000u};
    return t[cat];
}
// 4-bit BAM code -> category of the read character, or -1 (KeyError in the reference)
// "=ACMGRSVTWYHKDBN": A=1 C=2 G=4 T=8 N=15
__device__ __forceinline__ int cat_of_code(int code) {
    switch (code) {
    case 1: return CAT_A; case 2: return CAT_C; case 4: return CAT_G; case 8: return CAT_T; case 15: return CAT_N;
    default: return -1;
    }
}
__device__ __forceinline__ int ascii_of_code(int code) { return "=ACMGRSVTWYHKDBN"[code]; }
__device__ __forceinline__ int cat_of_ascii(int ch) {
    switch (ch) {
    case 'A': return CAT_A; case 'C': return CAT_C; case 'G': return CAT_G; case 'T': return CAT_T; case 'N': return CAT_N;
    default: return -1;
    }
}

struct ReadAux {      // per read, produced by the prep kernel
    int ref_start;    // get_reference_positions()[0] (Pileup.py:191) or INT_MIN when the read has no M/=/X base
    int endpos;       // bam_endpos (dep/htslib/sam.c:347-353)
    int aln_start;    // query_alignment_start: leading soft clip (H skipped)
    int aln_len;      // len(query_alignment_sequence)
};

__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

// deltaRef / deltaRead of Pileup.py:47-48 : only M,D,N advance the reference; only M,I advance the query
__device__ __forceinline__ bool op_adv_ref(int op) { return op == 0 || op == 2 || op == 3; }
__device__ __forceinline__ bool op_adv_read(int op) { return op == 0 || op == 1; }

// ------------------------------------------------------------------------------------------------
// Per-read digest for the fast path (v3), written by the prep kernel.  `clean` reads are those on which the
// reference cannot raise and whose CIGAR stays inside the fast path's closed form: ops M/I/D/S only, no
// zero-length op, S only as first/last op, no I next to an I or an S, first non-S op is M, query advance equal
// to the aligned length, every base in {A,C,G,T,N}.  A window that meets any other read is recomputed by the
// exact (v1) kernel.
struct __align__(16) ReadInfo {
    int ref_start;     // first aligned reference position (INT_MIN: none)
    int endk;          // bam_endpos, or INT_MIN when the read can never contribute (mapq < cut, no aligned base)
    int aln_start;     // leading soft clip
    int aln_len;       // len(query_alignment_sequence)
    int l_seq;
    int mapq28;        // min(mapq, 28): the quality table is flat from 28 on
    int nops;
    int flags;         // bit 0: clean; bits 8..15: checkpoint shift
    long long c0;      // cigar offset
    long long seq_off;
    long long qual_off;
    int span;          // reference advance of the whole CIGAR (Pileup.py deltaRef rule)
    int pad;
};
constexpr int RI_CLEAN = 1;

// prep: GW warps per read (1: short CIGARs, one warp per read; 8: long CIGARs, one block per read).  Each thread takes PK
// consecutive ops per round (independent loads), scans run per warp and are stitched across the block's warps.
template <int GW>
__global__ void __launch_bounds__(256) prep_reads_kernel(dp_reads R, ReadAux *__restrict__ aux,
                                                         int *__restrict__ RP, int *__restrict__ QP,
                                                         ReadInfo *__restrict__ info, int *__restrict__ ckpt,
                                                         int *__restrict__ endk, int mapq_cut) {
    constexpr int PK = 4;
    constexpr int GT = GW * 32;                               // threads per read
    __shared__ int s_part[8][4];                              // per-warp totals of a round
    __shared__ int s_first[2];                                // first aligned op, D/N advance before it
    const int lane = threadIdx.x & 31, wig = GW == 1 ? 0 : (threadIdx.x >> 5);
    const int gt = wig * 32 + lane;
    const int r = GW == 1 ? (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5) : (int)blockIdx.x;
    if (r >= R.n_reads) return;
    const int64_t c0 = R.cigar_off[r];
    const int nops = (int)(R.cigar_off[r + 1] - c0);
    const int pos = R.pos[r];
    int carryR = 0, carryQ = 0, carryH = 0, carryDN = 0;
    int first_j = INT_MAX, first_dn = 0;                      // this thread's first aligned op
    bool dirty = false;
    const int op_first = nops > 0 ? (int)(R.cigar[c0] & 15) : -1;
    if (GW > 1) { if (threadIdx.x == 0) s_first[0] = INT_MAX; __syncthreads(); }
    for (int base = 0; base < nops; base += GT * PK) {
        const int jb = base + gt * PK;
        uint32_t w[PK + 2];                                  // w[0] = op before the thread's first, w[PK+1] = op after its last
#pragma unroll
        for (int t = 0; t < PK + 2; t++) {
            const int j = jb + t - 1;
            w[t] = (j >= 0 && j < nops) ? __ldg(R.cigar + c0 + j) : 0xFFFFFFFFu;
        }
        int sR = 0, sQ = 0, sH = 0, sDN = 0;
#pragma unroll
        for (int t = 1; t <= PK; t++) {
            if (jb + t - 1 >= nops) continue;
            const int op = w[t] & 15, n = (int)(w[t] >> 4);
            sR += op_adv_ref(op) ? n : 0;
            sQ += op_adv_read(op) ? n : 0;
            sH += (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) ? n : 0;   // bam_cigar2rlen
            sDN += (op == 2 || op == 3) ? n : 0;
        }
        const int iR = warp_incl_scan(sR, lane), iQ = warp_incl_scan(sQ, lane);
        const int iH = warp_incl_scan(sH, lane), iDN = warp_incl_scan(sDN, lane);
        int oR = 0, oQ = 0, oDN = 0;                          // totals of the warps before this one
        int tR = __shfl_sync(0xffffffffu, iR, 31), tQ = __shfl_sync(0xffffffffu, iQ, 31);
        int tH = __shfl_sync(0xffffffffu, iH, 31), tDN = __shfl_sync(0xffffffffu, iDN, 31);
        if (GW > 1) {
            if (lane == 31) { s_part[wig][0] = iR; s_part[wig][1] = iQ; s_part[wig][2] = iH; s_part[wig][3] = iDN; }
            __syncthreads();
            tR = tQ = tH = tDN = 0;
#pragma unroll
            for (int k = 0; k < GW; k++) {
                const int pR = s_part[k][0], pQ = s_part[k][1], pH = s_part[k][2], pDN = s_part[k][3];
                if (k < wig) { oR += pR; oQ += pQ; oDN += pDN; }
                tR += pR; tQ += pQ; tH += pH; tDN += pDN;
            }
            __syncthreads();
        }
        int eR = carryR + oR + iR - sR, eQ = carryQ + oQ + iQ - sQ, eDN = carryDN + oDN + iDN - sDN;
#pragma unroll
        for (int t = 1; t <= PK; t++) {
            const int j = jb + t - 1;
            if (j >= nops) continue;
            const int op = w[t] & 15, n = (int)(w[t] >> 4);
            RP[c0 + r + j] = eR;                             // nops+1 entries per read, base c0 + r
            QP[c0 + j] = eQ;
            // ---- fast-path envelope
            const int opp = j > 0 ? (int)(w[t - 1] & 15) : -1;
            const int opn = j + 1 < nops ? (int)(w[t + 1] & 15) : -1;
            if (!(op == 0 || op == 1 || op == 2 || op == 4) || n == 0) dirty = true;
            if (op == 4 && !(j == 0 || j == nops - 1)) dirty = true;
            if (op == 1 && (opp == 1 || opp == 4 || opn == 1 || opn == 4)) dirty = true;
            if (op != 0 && op != 4 && (j == 0 || (j == 1 && op_first == 4))) dirty = true;   // first non-S op must be M
            if (first_j == INT_MAX && (op == 0 || op == 7 || op == 8) && n > 0) { first_j = j; first_dn = eDN; }
            eR += op_adv_ref(op) ? n : 0;
            eQ += op_adv_read(op) ? n : 0;
            eDN += (op == 2 || op == 3) ? n : 0;
        }
        carryR += tR; carryQ += tQ; carryH += tH; carryDN += tDN;
    }
    // ---- first aligned reference position (get_reference_positions()[0], Pileup.py:191)
    int ref_start = INT_MIN;
    if (GW == 1) {
        const unsigned m = __ballot_sync(0xffffffffu, first_j != INT_MAX);     // lanes hold ascending op ranges per round
        int best = first_j;
#pragma unroll
        for (int d = 16; d; d >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, d));
        const unsigned owner = __ballot_sync(0xffffffffu, first_j == best && first_j != INT_MAX);
        if (m) ref_start = pos + __shfl_sync(0xffffffffu, first_dn, __ffs(owner) - 1);
    } else {
        if (first_j != INT_MAX) atomicMin(&s_first[0], first_j);
        __syncthreads();
        if (first_j != INT_MAX && first_j == s_first[0]) s_first[1] = first_dn;
        __syncthreads();
        if (s_first[0] != INT_MAX) ref_start = pos + s_first[1];
    }
    // ---- soft-clip extents (thread-uniform, short loops)
    const int l_seq = R.l_seq[r];
    int s = 0, e = l_seq;
    for (int c = 0; c < nops; c++) {
        const uint32_t w = R.cigar[c0 + c];
        const int op = w & 15;
        if (op == 5) continue;
        if (op == 4) s += (int)(w >> 4); else break;
    }
    for (int c = nops - 1; c >= 0; c--) {
        const uint32_t w = R.cigar[c0 + c];
        const int op = w & 15;
        if (op == 5) continue;
        if (op == 4) e -= (int)(w >> 4); else break;
    }
    const int aln_len = e - s > 0 ? e - s : 0;
    const int endpos = (!(R.flag[r] & 4) && nops > 0) ? pos + carryH : pos + 1;
    const int64_t so = R.seq_off[r], qo = R.qual_off[r];
    if (carryQ != aln_len || aln_len <= 0 || ref_start != pos || carryR <= 0) dirty = true;
    if (GW == 1) dirty = __any_sync(0xffffffffu, dirty); else dirty = __syncthreads_or(dirty);
    // ---- checkpoints: ckpt[m] = first op j with RP[j+1] >= m << k ; at most nops/2 + 1 of them per read
    int k = 4;
    while ((carryR >> k) > (nops >> 1)) k++;
    const int64_t ckb = (c0 >> 1) + r;
    {
        int cR = 0;
        for (int base = 0; base < nops; base += GT * PK) {
            const int jb = base + gt * PK;
            int dR[PK];
            int sR = 0;
#pragma unroll
            for (int t = 0; t < PK; t++) {
                const uint32_t w = jb + t < nops ? __ldg(R.cigar + c0 + jb + t) : 0u;
                const int op = w & 15;
                dR[t] = op_adv_ref(op) ? (int)(w >> 4) : 0;
                sR += dR[t];
            }
            const int iR = warp_incl_scan(sR, lane);
            int oR = 0, tR = __shfl_sync(0xffffffffu, iR, 31);
            if (GW > 1) {
                if (lane == 31) s_part[wig][0] = iR;
                __syncthreads();
                tR = 0;
#pragma unroll
                for (int q = 0; q < GW; q++) { const int pR = s_part[q][0]; if (q < wig) oR += pR; tR += pR; }
                __syncthreads();
            }
            int a0 = cR + oR + iR - sR;
#pragma unroll
            for (int t = 0; t < PK; t++) {
                if (dR[t] > 0) {
                    const int b0 = a0 + dR[t];
                    for (int m = (a0 >> k) + 1; m <= (b0 >> k); m++) ckpt[ckb + m] = jb + t;
                    a0 = b0;
                }
            }
            cR += tR;
        }
    }
    if (gt == 0) {
        ckpt[ckb] = 0;
        RP[c0 + r + nops] = carryR;
        ReadAux a;
        a.ref_start = ref_start;
        a.endpos = endpos;
        a.aln_start = s;
        a.aln_len = aln_len;
        aux[r] = a;
        const int mq = R.mapq[r];
        const bool contributes = mq >= mapq_cut && ref_start != INT_MIN;
        ReadInfo ri;
        ri.ref_start = ref_start;
        ri.endk = contributes ? endpos : INT_MIN;
        ri.aln_start = s; ri.aln_len = aln_len; ri.l_seq = l_seq;
        ri.mapq28 = mq < 28 ? mq : 28;
        ri.nops = nops;
        ri.flags = (dirty ? 0 : RI_CLEAN) | (k << 8);
        ri.c0 = c0; ri.seq_off = so; ri.qual_off = qo;
        ri.span = carryR; ri.pad = 0;
        info[r] = ri;
        endk[r] = ri.endk;
    }
}

// Per-base code bytes of the fast path, one thread per query base (runs after prep_reads_kernel): entry idx of a
// read pairs the base at ALIGNED index idx (query index idx + leading clip) with the quality at FULL-query index
// idx -- the pairing parseRead makes (Pileup.py:184,197) -- as  own category << 5 | min(quality, mapq, 28).
// A base outside {A,C,G,T,N} (KeyError in the reference) clears the read's clean flag.
__global__ void __launch_bounds__(256) base_codes_kernel(dp_reads R, ReadInfo *__restrict__ info,
                                                         uint8_t *__restrict__ code0) {
    constexpr int BK = 16;                            // consecutive bases per thread: one 16-byte quality load, one 16-byte store
    const int64_t n = R.n_bases;
    const int lane = threadIdx.x & 31;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x * BK;
    // the loop bound is warp-uniform (first base of the warp), so the warp-wide search below always has all 32 lanes
    for (int64_t kw = ((int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31)) * BK; kw < n; kw += stride) {
        const int64_t k0 = kw + lane * BK;
        // read holding the warp's first base (qual_off[r] <= kw < qual_off[r+1]) by a 33-ary search: every lane probes one
        // pivot per round -- 3 rounds for 32 k reads instead of 15 dependent loads per thread
        int lo = 0, hi = R.n_reads;
        while (hi - lo > 1) {
            const int span = hi - lo;
            const int piv = lo + (int)(((int64_t)span * (lane + 1)) / 33);
            const bool le = piv <= lo || __ldg(R.qual_off + piv) <= kw;       // qual_off[lo] <= kw by the invariant; piv < hi
            const unsigned m = __ballot_sync(0xffffffffu, le);               // monotone: a prefix of the lanes
            const int nle = __popc(m);
            const int nlo = nle ? lo + (int)(((int64_t)span * nle) / 33) : lo;
            const int nhi = nle < 32 ? lo + (int)(((int64_t)span * (nle + 1)) / 33) : hi;
            lo = nlo; hi = max(nhi, nlo + 1);
        }
        if (k0 >= n) continue;
        int r = lo;
        while (r + 1 < R.n_reads && k0 >= __ldg(R.qual_off + r + 1)) r++;     // this lane's first read (reads shorter than 512 bases)
        int64_t ro = 0, next = 0;
        int l_seq = 0, aln_start = 0, mq28 = 0;
        const uint8_t *seqp = nullptr;
        bool illegal = false;
        auto load_read = [&]() {
            const ReadInfo *ri = info + r;
            ro = ri->qual_off; l_seq = ri->l_seq; aln_start = ri->aln_start; mq28 = ri->mapq28;
            seqp = R.seq4 + ri->seq_off;
            next = ro + l_seq;
            illegal = false;
        };
        load_read();
        uint32_t qw[BK / 4];
        if (k0 + BK <= n) {
            const uint4 q4 = __ldg(reinterpret_cast<const uint4 *>(R.qual + k0));
            qw[0] = q4.x; qw[1] = q4.y; qw[2] = q4.z; qw[3] = q4.w;
        } else {
#pragma unroll
            for (int t = 0; t < BK / 4; t++) qw[t] = 0;
            for (int t = 0; k0 + t < n; t++) qw[t >> 2] |= (uint32_t)__ldg(R.qual + k0 + t) << ((t & 3) << 3);
        }
        uint32_t out[BK / 4] = {0, 0, 0, 0};
        int last_byte_i = -1, last_byte2_i = -1;
        uint32_t last_byte = 0, last_byte2 = 0;
#pragma unroll
        for (int t = 0; t < BK; t++) {
            const int64_t k = k0 + t;
            if (k >= n) break;
            while (k >= next && r + 1 < R.n_reads) {
                if (illegal) atomicAnd(&info[r].flags, ~RI_CLEAN);
                r++;
                load_read();
                last_byte_i = last_byte2_i = -1;
            }
            const int idx = (int)(k - ro);
            if ((idx >> 1) != last_byte_i) { last_byte_i = idx >> 1; last_byte = __ldg(seqp + last_byte_i); }
            const int nib = (idx & 1) ? (last_byte & 15) : (last_byte >> 4);
            if (!((0x8116u >> nib) & 1u)) illegal = true;
            const int kb = idx + aln_start;
            int cat = 0;
            if (kb < l_seq) {
                int nib2 = nib;
                if (aln_start != 0) {
                    if ((kb >> 1) != last_byte2_i) { last_byte2_i = kb >> 1; last_byte2 = __ldg(seqp + last_byte2_i); }
                    nib2 = (kb & 1) ? (last_byte2 & 15) : (last_byte2 >> 4);
                }
                cat = nib2 == 15 ? CAT_N : ((__ffs(nib2) - 1) & 3);
            }
            const int q = (int)((qw[t >> 2] >> ((t & 3) << 3)) & 0xffu);
            const uint32_t code = (uint32_t)((cat << 5) | min(q, mq28));
            out[t >> 2] |= code << ((t & 3) << 3);
        }
        if (illegal) atomicAnd(&info[r].flags, ~RI_CLEAN);
        if (k0 + BK <= n) *reinterpret_cast<uint4 *>(code0 + k0) = make_uint4(out[0], out[1], out[2], out[3]);
        else for (int t = 0; k0 + t < n; t++) code0[k0 + t] = (uint8_t)(out[t >> 2] >> ((t & 3) << 3));
    }
}

// ------------------------------------------------------------------------------------------------
constexpr int ENC_THREADS = 256;
constexpr int ENC_WARPS = ENC_THREADS / 32;
constexpr int SEG_CAP = 64;        // staged ops per read per round
constexpr int CAND_BLOCK = 256;    // candidate reads filtered per compaction round

struct __align__(16) SegOp {
    int sA;        // column (relative to queryStart) of the op's first reference position
    int n;         // op length
    int qp;        // readPosition at op start (lags by one: -1 + query bases consumed so far)
    short rel;     // relativeIndexRef at op start
    uint8_t op;
    uint8_t pad;
};
struct SegHdr {
    long long seq_off, qual_off;
    int aln_start, aln_len, l_seq, mapq;
    int nops;      // staged ops in this slot (0 = nothing)
    int more;      // read has further relevant ops beyond this round
    int next_j;    // op index (within the read) at which the next round starts
    int next_A, next_qp, next_rel;   // carried state for the next round
    int c0lo, c0hi;                  // cigar base offset (split int64)
};

struct EncArgs {
    dp_reads R;
    const ReadAux *aux;
    const int *RP;
    const int *QP;
    const ReadInfo *info;      // fast path (v3) digest, checkpoints and filtered end positions
    const int *ckpt;
    const int *endk;
    const uint8_t *code0;      // per query base: own category << 5 | min(quality, mapq, 28), indexed like qual[]
    const int32_t *win_qs, *win_lo, *win_hi;
    const uint8_t *ref;
    long long ref_base, ref_len;
    int L, W, H, mapq_cut;
    uint8_t *out_img;
    int32_t *out_rowpos, *out_rowins, *out_status;
    int32_t *out_colrows;      // optional [n_windows][L]: image rows each window column produces (0: unpopulated, else 1 + its
                               // insert rows) -- the full row layout, not cut at H (dp_window_layout)
    int layout_only;           // exact kernel: stop after the row layout
};

// Stage one round (<= SEG_CAP ops) of read r into slot (hdr, ops).  One warp.
// first==true: locate the first relevant op by binary search on the reference-advance prefix; else
// continue from the state saved in hdr.
__device__ void stage_read(const EncArgs &a, int r, int qs, int qe, bool first, SegHdr *hdr, SegOp *ops, int lane) {
    int j0, A0, qp0, rel0;
    long long c0;
    int nops_total;
    if (first) {
        const ReadAux ax = a.aux[r];
        c0 = a.R.cigar_off[r];
        nops_total = (int)(a.R.cigar_off[r + 1] - c0);
        const int refStart = ax.ref_start;
        // first op whose end (refStart + RP[j+1]) >= qs : everything before it emits nothing in-window.
        // RP has nops+1 entries per read (base c0 + r); warp-cooperative 32-ary lower bound on RP[j+1].
        const long long rpb = c0 + r;
        const int T = qs - refStart;
        int lo = 0, hi = nops_total;                          // answer in [lo, hi]
        while (lo < hi) {
            const int len = hi - lo;
            const int step = (len + 31) >> 5;
            const int jk = min(lo + (lane + 1) * step - 1, hi - 1);
            const bool before = a.RP[rpb + jk + 1] < T;
            const int k = __popc(__ballot_sync(0xffffffffu, before));
            if (k == 32) { lo = hi; break; }
            const int nlo = lo + k * step;
            const int nhi = min(lo + (k + 1) * step, hi) - 1;  // block k's probe is not-before => answer <= nhi
            lo = nlo; hi = max(nhi, nlo);
        }
        j0 = lo;
        if (j0 < nops_total) {
            A0 = refStart + a.RP[rpb + j0] - qs;
            qp0 = -1 + a.QP[c0 + j0];
        } else { A0 = 0; qp0 = 0; }
        rel0 = refStart > qs ? refStart - qs - 1 : 0;        // Pileup.py:203-204
        if (lane == 0) {
            hdr->seq_off = a.R.seq_off[r];
            hdr->qual_off = a.R.qual_off[r];
            hdr->aln_start = ax.aln_start;
            hdr->aln_len = ax.aln_len;
            hdr->l_seq = a.R.l_seq[r];
            hdr->mapq = a.R.mapq[r];
            hdr->c0lo = (int)(c0 & 0xffffffffll);
            hdr->c0hi = (int)(c0 >> 32);
        }
    } else {
        j0 = hdr->next_j; A0 = hdr->next_A; qp0 = hdr->next_qp; rel0 = hdr->next_rel;
        c0 = ((long long)hdr->c0hi << 32) | (unsigned int)hdr->c0lo;
        nops_total = (int)(a.R.cigar_off[r + 1] - c0);
    }
    __syncwarp();
    const int Lwin = qe - qs;
    int staged = 0;
    int more = 0;
    int carryA = A0, carryQ = qp0, carryRel = rel0;
    int j = j0;
    for (int round = 0; round < SEG_CAP / 32; round++) {
        const int jj = j + lane;
        const bool valid = jj < nops_total;
        const uint32_t w = valid ? a.R.cigar[c0 + jj] : 0u;
        const int op = w & 15, n = (int)(w >> 4);
        const int dR = op_adv_ref(op) ? n : 0;
        const int dQ = op_adv_read(op) ? n : 0;
        const int iR = warp_incl_scan(dR, lane), iQ = warp_incl_scan(dQ, lane);
        const int sA = carryA + iR - dR;
        const int qp = carryQ + iQ - dQ;
        // in-window M/D steps of this op (advance of relativeIndexRef, Pileup.py:252)
        int inwin = 0;
        if (valid && (op == 0 || op == 2) && n > 0) {
            inwin = (sA >= 0 && sA < Lwin) ? 1 : 0;
            const int lo_i = max(1, 1 - sA);                 // col_i = sA + i - 1 >= 0
            const int hi_i = min(n - 1, Lwin - sA);          // col_i <= Lwin-1
            if (hi_i >= lo_i) inwin += hi_i - lo_i + 1;
        }
        const int iW = warp_incl_scan(inwin, lane);
        const int rel = carryRel + iW - inwin;
        const bool relevant = valid && sA < Lwin;            // ops starting at/after queryEnd emit nothing
        const unsigned mrel = __ballot_sync(0xffffffffu, relevant);
        // relevant lanes form a prefix (sA is monotone)
        const int cnt = __popc(mrel);
        if (relevant) {
            SegOp so;
            so.sA = sA; so.n = n; so.qp = qp; so.rel = (short)min(rel, 32000); so.op = (uint8_t)op; so.pad = 0;
            ops[staged + lane] = so;
        }
        staged += cnt;
        carryA += __shfl_sync(0xffffffffu, iR, 31);
        carryQ += __shfl_sync(0xffffffffu, iQ, 31);
        carryRel += __shfl_sync(0xffffffffu, iW, 31);
        j += 32;
        if (cnt < 32) { more = 0; break; }
        if (j >= nops_total) { more = 0; break; }
        more = 1;                                            // possibly more relevant ops
    }
    if (lane == 0) {
        hdr->nops = staged;
        hdr->more = more;
        hdr->next_j = j; hdr->next_A = carryA; hdr->next_qp = carryQ; hdr->next_rel = carryRel;
    }
    __syncwarp();
}

// Python indexing of readSequence / readQualities (negative index wraps once)
struct ReadCtx {
    const uint8_t *seq4;
    const uint8_t *qual;
    long long seq_off, qual_off;
    int aln_start, aln_len, l_seq, mapq;
};
__device__ __forceinline__ int read_code(const ReadCtx &c, int idx, int &err) {
    if (idx < 0) idx += c.aln_len;
    if (idx < 0 || idx >= c.aln_len) { err = DP_EINDEX; return 15; }
    const int k = idx + c.aln_start;
    const uint8_t b = __ldg(c.seq4 + c.seq_off + (k >> 1));
    return (k & 1) ? (b & 15) : (b >> 4);
}
__device__ __forceinline__ int read_q(const ReadCtx &c, int idx, int &err) {
    if (idx < 0) idx += c.l_seq;
    if (idx < 0 || idx >= c.l_seq) { err = DP_EINDEX; return 0; }
    const int bq = __ldg(c.qual + c.qual_off + idx);
    return c_qlut[min(bq, c.mapq)];
}

// first staged op with key > P, key = (advancing && n>0) ? end column : start column + 1
__device__ __forceinline__ int seg_lower_bound(const SegOp *ops, int nops, int P) {
    int lo = 0, hi = nops;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const SegOp o = ops[mid];
        const bool adv = op_adv_ref(o.op) && o.n > 0;
        const int key = adv ? o.sA + o.n : o.sA + 1;
        if (key <= P) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// MODE 0: mark (populated, max insert length)   MODE 1: count   MODE 2: place
// Anchor rows/columns: kins == 0.  Insert rows: kins = k+1.
template <int MODE>
__device__ __forceinline__ void visit_slot(const EncArgs &a, const SegHdr *hdr, const SegOp *ops, int P, int kins,
                                           const uint8_t *s_ref, int Lwin, int &populated, int &max_ins,
                                           int *cnt, const int *base, int *run, uint8_t *qrow, int Wm1, int &err) {
    const int nops = hdr->nops;
    if (nops == 0) return;
    int j = seg_lower_bound(ops, nops, P);
    if (j >= nops) return;
    ReadCtx rc;
    rc.seq4 = a.R.seq4; rc.qual = a.R.qual;
    rc.seq_off = hdr->seq_off; rc.qual_off = hdr->qual_off;
    rc.aln_start = hdr->aln_start; rc.aln_len = hdr->aln_len; rc.l_seq = hdr->l_seq; rc.mapq = hdr->mapq;
    for (; j < nops; j++) {
        const SegOp o = ops[j];
        if (o.sA > P) break;
        const int op = o.op, n = o.n;
        if (n <= 0) continue;
        if (op == 1) {                                       // insert (Pileup.py:220-230), anchor = sA, needs A > qs
            if (o.sA != P || P <= 0) continue;
            if (MODE == 0) { max_ins = max(max_ins, n); continue; }
            if (kins == 0 || kins > n) continue;
            const int idx = o.qp + kins;                     // readPosition+1+k
            // slices (not indexing): idx >= 0 always; a short slice raises IndexError at [k]
            if (idx >= rc.aln_len || idx >= rc.l_seq) { err = DP_EINDEX; continue; }
            const int code = read_code(rc, idx, err);
            const int cat = cat_of_code(code);
            if (cat < 0) { err = DP_EKEY; continue; }
            if (MODE == 1) cnt[cat]++;
            if (MODE == 2) {
                const int slot = base[cat] + run[cat]++;
                if (slot < Wm1) qrow[1 + slot] = (uint8_t)read_q(rc, idx, err);
            }
            continue;
        }
        if (kins != 0) continue;                             // insert rows only take insert ops
        const bool adv = op_adv_ref(op);
        int i_first, i_count;                                // steps of this op that land on column P
        if (adv) {
            if (P == o.sA) { i_first = 0; i_count = n >= 2 ? 2 : 1; }
            else { i_first = P - o.sA + 1; i_count = (i_first <= n - 1) ? 1 : 0; }
        } else {
            if (o.sA != P) continue;
            i_first = 0; i_count = (op == 4) ? 1 : n;        // soft clip: one step, then break (:257-258)
        }
        if (i_count == 0) continue;
        if (MODE == 0) { populated = 1; continue; }
        const int dq = (op == 0) ? 1 : 0;                    // deltaRead for non-insert ops: only M
        const bool md = (op == 0 || op == 2);
        if (!adv) {
            // i_count identical entries: query index and relativeIndexRef do not move
            const int code = read_code(rc, o.qp, err);
            const int rel = o.rel;
            if (rel < 0 || rel > Lwin) { err = DP_EINDEX; continue; }
            const int fc = s_ref[rel];
            const int rch = ascii_of_code(code);
            const int cat = (rch == fc) ? CAT_M : cat_of_code(code);
            if (cat < 0) { err = DP_EKEY; continue; }
            if (MODE == 1) cnt[cat] += i_count;
            if (MODE == 2) {
                const int q = read_q(rc, o.qp, err);
                for (int t = 0; t < i_count; t++) {
                    const int slot = base[cat] + run[cat]++;
                    if (slot < Wm1) qrow[1 + slot] = (uint8_t)q; else break;
                }
            }
            continue;
        }
        // in-window steps of this op before step i (advance of rel inside the op)
        const int in0 = (o.sA >= 0 && o.sA < Lwin) ? 1 : 0;
        const int lo_i = max(1, 1 - o.sA);
        for (int t = 0; t < i_count; t++) {
            const int i = i_first + t;
            const int idx = o.qp + dq * i;
            int rel = o.rel;
            if (md && i > 0) rel += in0 + max(0, i - lo_i);
            if (rel < 0 || rel > Lwin) { err = DP_EINDEX; continue; }
            int cat;
            if (op == 2) {
                cat = CAT_D;                                 // getPileupEncoding (:128-134)
                if (MODE == 1) {                             // the reference still indexes seq/qual (:240,:243)
                    int e2 = 0; (void)read_code(rc, idx, e2); if (e2) err = e2;
                }
            } else {
                const int code = read_code(rc, idx, err);
                const int fc = s_ref[rel];
                cat = (ascii_of_code(code) == fc) ? CAT_M : cat_of_code(code);
                if (cat < 0) { err = DP_EKEY; continue; }
            }
            if (MODE == 1) cnt[cat]++;
            if (MODE == 2) {
                const int slot = base[cat] + run[cat]++;
                if (slot < Wm1) qrow[1 + slot] = (uint8_t)read_q(rc, idx, err);
            }
        }
    }
}

template <int MODE>
__device__ void run_pass(const EncArgs &a, int qs, int qe, int lo, int hi, SegHdr *s_hdr, SegOp *s_ops, int *s_cand,
                         int *s_ncand, int *s_flags, const uint8_t *s_ref, int P, int kins, bool active,
                         int &populated, int &max_ins, int *cnt, const int *base, int *run, uint8_t *qrow, int Wm1,
                         int &err) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int Lwin = qe - qs;
    for (int cb = lo; cb < hi; cb += CAND_BLOCK) {
        // ---- ordered compaction of the candidates that pass fetch (hts.c:2242) + mapq (:160) + aligned (:187)
        const int r = cb + tid;
        bool keep = false;
        if (r < hi) {
            const ReadAux ax = a.aux[r];
            keep = a.R.pos[r] < qe && ax.endpos > qs && (int)a.R.mapq[r] >= a.mapq_cut && ax.ref_start != INT_MIN;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) s_flags[warp] = __popc(bal);
        __syncthreads();
        int off = 0, total = 0;
        for (int w = 0; w < ENC_WARPS; w++) { const int c = s_flags[w]; if (w < warp) off += c; total += c; }
        if (keep) s_cand[off + __popc(bal & ((1u << lane) - 1))] = r;
        __syncthreads();
        // ---- chunks of ENC_WARPS reads: stage (one warp per read) then visit (one thread per column/row)
        for (int ch = 0; ch < total; ch += ENC_WARPS) {
            const int nslots = min(ENC_WARPS, total - ch);
            if (warp < nslots) stage_read(a, s_cand[ch + warp], qs, qe, true, &s_hdr[warp], &s_ops[warp * SEG_CAP], lane);
            else if (lane == 0) { s_hdr[warp].nops = 0; s_hdr[warp].more = 0; }
            __syncthreads();
            int any_more = 0;
            for (int w = 0; w < nslots; w++) any_more |= s_hdr[w].more;
            if (!any_more) {
                if (active)
                    for (int w = 0; w < nslots; w++)
                        visit_slot<MODE>(a, &s_hdr[w], &s_ops[w * SEG_CAP], P, kins, s_ref, Lwin, populated, max_ins,
                                         cnt, base, run, qrow, Wm1, err);
                __syncthreads();
            } else {
                // rare: a read with more than SEG_CAP relevant ops -> strictly serial over the chunk's reads
                for (int w = 0; w < nslots; w++) {
                    while (true) {
                        if (active)
                            visit_slot<MODE>(a, &s_hdr[w], &s_ops[w * SEG_CAP], P, kins, s_ref, Lwin, populated,
                                             max_ins, cnt, base, run, qrow, Wm1, err);
                        __syncthreads();
                        const int more = s_hdr[w].more;
                        __syncthreads();
                        if (!more) break;
                        if (warp == w) stage_read(a, s_cand[ch + w], qs, qe, false, &s_hdr[w], &s_ops[w * SEG_CAP], lane);
                        __syncthreads();
                    }
                }
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(ENC_THREADS) encode_windows_kernel(EncArgs a, const int *__restrict__ win_count,
                                                                     const int *__restrict__ win_list) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x;
    if (win_list && (int)blockIdx.x >= *win_count) return;      // fallback launch: only the listed windows
    const int w = win_list ? win_list[blockIdx.x] : blockIdx.x;
    const int L = a.L, W = a.W, H = a.H;
    const int qs = a.win_qs[w];
    const int qe = qs + L;
    const int lo = a.win_lo[w], hi = a.win_hi[w];

    // ---- shared memory carve-up
    uint8_t *p = smem;
    SegOp *s_ops = (SegOp *)p;            p += sizeof(SegOp) * ENC_WARPS * SEG_CAP;
    SegHdr *s_hdr = (SegHdr *)p;          p += sizeof(SegHdr) * ENC_WARPS;
    int *s_cand = (int *)p;               p += sizeof(int) * CAND_BLOCK;
    int *s_rowstart = (int *)p;           p += sizeof(int) * (ENC_THREADS + 4);
    int *s_flags = (int *)p;              p += sizeof(int) * 16;
    short *s_rowcol = (short *)p;         p += sizeof(short) * ENC_THREADS;
    short *s_rowk = (short *)p;           p += sizeof(short) * ENC_THREADS;
    unsigned short *s_cum = (unsigned short *)p; p += sizeof(unsigned short) * ENC_THREADS * 8;
    unsigned short *s_maxins = (unsigned short *)p; p += sizeof(unsigned short) * ENC_THREADS;
    uint8_t *s_ref = p;                   p += ((L + 1 + 15) / 16) * 16;
    uint8_t *s_rowref = p;                p += ENC_THREADS;
    uint8_t *s_q = p;                     // [H][W] quality bytes (pixel 0 unused)
    int *s_err = &s_flags[8];
    int *s_nrows = &s_flags[9];
    int *s_ncand = &s_flags[10];

    if (tid == 0) { *s_err = 0; *s_nrows = 0; }
    // window must lie inside the supplied reference: refSequence = ref[qs .. qe] (L+1 chars, Pileup.py:41)
    const long long roff = (long long)qs - a.ref_base;
    const bool ref_ok = roff >= 0 && roff + L + 1 <= a.ref_len;
    for (int i = tid; i < L + 1; i += ENC_THREADS) s_ref[i] = ref_ok ? a.ref[roff + i] : (uint8_t)'N';
    __syncthreads();

    int err = ref_ok ? 0 : DP_EINVAL;
    int cnt[NCAT], base[NCAT], run[NCAT];
#pragma unroll
    for (int c = 0; c < NCAT; c++) { cnt[c] = 0; base[c] = 0; run[c] = 0; }

    // ---- pass 0: one thread per column
    int populated = 0, max_ins = 0;
    {
        const bool active = tid < L;
        run_pass<0>(a, qs, qe, lo, hi, s_hdr, s_ops, s_cand, s_ncand, s_flags, s_ref, tid, 0, active, populated, max_ins,
                    cnt, base, run, nullptr, W - 1, err);
    }
    if (a.out_colrows && tid < L) a.out_colrows[(size_t)w * L + tid] = populated ? 1 + max_ins : 0;
    if (a.layout_only) {                     // dp_window_layout: the label of a window is widened by ALL its insert rows
        if (err) atomicMin(s_err, err);
        __syncthreads();
        if (tid == 0 && a.out_status) a.out_status[w] = *s_err;
        return;
    }
    // ---- row map: rows = populated columns in ascending order, each followed by its insert rows (:295-317)
    {
        const int nrows_here = (tid < L && populated) ? 1 + max_ins : 0;
        s_rowstart[tid] = nrows_here;
        s_maxins[tid] = (unsigned short)min(max_ins, 65535);
        s_rowcol[tid] = -1; s_rowk[tid] = 0; s_rowref[tid] = 'N';
        __syncthreads();
        if (tid == 0) {                      // L <= 256: a serial scan is ~200 adds
            int acc = 0, anchors = 0;
            for (int c = 0; c < L; c++) {
                const int v = s_rowstart[c];
                s_rowstart[c] = acc;
                if (v) {
                    // reference pixel of an anchor row: refSequence[c'] with 'I' spliced at every insert row
                    // (:315) == original ref[#anchor rows before it]
                    if (acc < ENC_THREADS) s_rowref[acc] = s_ref[anchors];
                    anchors++;
                }
                acc += v;
            }
            *s_nrows = acc;
        }
        __syncthreads();
        if (tid < L && populated) {
            const int r0 = s_rowstart[tid];
            for (int k = 0; k <= max_ins; k++) {
                const int y = r0 + k;
                if (y >= H) break;
                s_rowcol[y] = (short)tid; s_rowk[y] = (short)k;
            }
        }
        __syncthreads();
    }
    const int nrows = min(*s_nrows, H);
    // ---- passes 1 and 2: one thread per image row
    {
        const bool active = tid < nrows;
        const int P = active ? s_rowcol[tid] : 0;
        const int kins = active ? s_rowk[tid] : 0;
        int dummy_p = 0, dummy_m = 0;
        run_pass<1>(a, qs, qe, lo, hi, s_hdr, s_ops, s_cand, s_ncand, s_flags, s_ref, P, kins, active, dummy_p, dummy_m, cnt,
                    base, run, nullptr, W - 1, err);
        int acc = 0;
#pragma unroll
        for (int c = 0; c < NCAT; c++) { base[c] = acc; acc += cnt[c]; if (acc < 0) acc = INT_MAX / 2; }
        if (tid < H) {
#pragma unroll
            for (int c = 0; c < NCAT; c++) s_cum[tid * 8 + c] = (unsigned short)min(base[c] + cnt[c], W - 1);
            s_cum[tid * 8 + 7] = 0;
        }
        uint8_t *qrow = s_q + (size_t)tid * W;
        run_pass<2>(a, qs, qe, lo, hi, s_hdr, s_ops, s_cand, s_ncand, s_flags, s_ref, P, kins, active, dummy_p, dummy_m, cnt,
                    base, run, active ? qrow : nullptr, W - 1, err);
    }
    if (err) atomicMin(s_err, err);         // codes are negative; keep the most severe deterministic choice
    __syncthreads();

    // reference pixel must be a legal colour key (KeyError at :281 otherwise)
    // ---- emit: RGBA rows, 16-byte stores
    uint8_t *img = a.out_img + (size_t)w * H * W * 4;
    const int quads_per_row = W / 4;        // W % 4 == 0 enforced by the host
    const int total_quads = H * quads_per_row;
    int bad_ref = 0;
    for (int qd = tid; qd < total_quads; qd += ENC_THREADS) {
        const int y = qd / quads_per_row;
        const int x0 = (qd - y * quads_per_row) * 4;
        uint32_t px[4];
        if (y >= nrows) {
            px[0] = px[1] = px[2] = px[3] = 0xFF000000u;
        } else {
            const bool ins = s_rowk[y] != 0;
            const uint32_t filler = ins ? 0xFFFF00FFu : 0xFF000000u;      // 'I' (255,0,255) / none (0,0,0), alpha 255
            const unsigned short *cum = &s_cum[y * 8];
            const int tot = cum[6];
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const int x = x0 + t;
                uint32_t v;
                if (x == 0) {
                    if (ins) v = 0xFFFF00FFu;
                    else {
                        const int ch = s_rowref[y];
                        const int cat = cat_of_ascii(ch);
                        if (cat < 0) { bad_ref = 1; v = 0xFF000000u; }
                        else v = cat_rgb(cat) | 0xFF000000u;
                    }
                } else {
                    const int e = x - 1;
                    if (e >= tot) v = filler;
                    else {
                        int cat = 0;
#pragma unroll
                        for (int c = 0; c < 6; c++) cat += (e >= cum[c]) ? 1 : 0;
                        v = cat_rgb(cat) | ((uint32_t)s_q[(size_t)y * W + x] << 24);
                    }
                }
                px[t] = v;
            }
        }
        *reinterpret_cast<uint4 *>(img + ((size_t)y * W + x0) * 4) = make_uint4(px[0], px[1], px[2], px[3]);
    }
    if (bad_ref) atomicMin(s_err, (int)DP_EKEY);
    for (int y = tid; y < H; y += ENC_THREADS) {
        if (a.out_rowpos) a.out_rowpos[(size_t)w * H + y] = (y < nrows) ? qs + s_rowcol[y] : -1;
        if (a.out_rowins) a.out_rowins[(size_t)w * H + y] = (y < nrows) ? s_rowk[y] : 0;
    }
    __syncthreads();
    if (tid == 0 && a.out_status) a.out_status[w] = *s_err;
}

// ================================================================================================
// Encoder v3 (fast path).  Same closed form as v1, specialised to `clean` reads (ReadInfo) so that nothing in the
// inner loops can raise, and restricted to the pileup columns that reach the image:
//   scan      one warp per read, lane per CIGAR op (entered through the per-read checkpoints, no search): insert
//             lengths per anchor column and a coverage difference array -> which columns are populated, how many image
//             rows each takes (anchor + insert rows, Pileup.py:295-317) and hence the last column that is still inside
//             the windowCutoff rows (Lgen).  With 3 % insertions only ~70 of the 201 columns are ever drawn.
//   generate  one warp per read: op-parallel part (first step of every M/D op, soft-clip step, inserted bases) and
//             column-parallel part (each lane walks a run of consecutive columns through the staged M/D ops) write one
//             code byte (category << 5 | quality index) per entry into shared-memory planes [kind][read slot][column];
//             per-(category, column) counts are taken with shared-memory atomics on the way.
//   place     one THREAD per column reads its codes in slot (= file) order and puts each at base[category] + running
//             rank -- (read order, step order) without sorting; the 7 running ranks live in two packed registers.
//   emit      filler quads straight from registers, entry quads through a 256-entry code -> RGBA table.
// For clean reads relativeIndexRef has the closed form  rel = column - (step == 0) + shift, where shift = 1 iff an
// M/D op of the read starts exactly at queryStart (Pileup.py:203-204,252 and SURVEY.md A.1 "shifted by +1").
// Anything outside the envelope (a read that is not clean, an illegal reference character, more reads / inserted
// bases than the shared-memory planes hold, insert rows deeper than the image) marks the window for the exact v1
// kernel, which then recomputes it.
constexpr int V3_THREADS = 256;
constexpr int V3_WARPS = V3_THREADS / 32;
constexpr int V3_MAXSLOT = 128;       // most read slots per batch (the pool may hold fewer)
// One shared-memory pool, carved per window: planes [slots][LS] words from the bottom (LS = drawn columns rounded up
// to 16), image-row codes [rows][RS] bytes from the top (RS = deepest row rounded up to whole quads).
constexpr int V3_POOL_BYTES = 36352;  // 4 CTAs per SM (64 registers); deep windows
constexpr int V3_POOL_SMALL = 24576;  // 5 CTAs per SM (48 registers): +3-4 % on 30x windows, -10 % on 100x ones (more of
                                      // them need a second batch of read slots) -- picked per call by the bases per window
constexpr int V3_POOL_DEEP = 55296;   // 3 CTAs per SM: 100x windows (~100 read slots x 48 columns + 150 rows x ~110 codes = 36-40 KB) keep the planes of
                                      // the counting pass for placement instead of generating every read a second time
constexpr int V3_CS = 208;            // column stride of the planes (>= window length)
constexpr int V3_MAXC = 192;          // surviving reads per window
constexpr int V3_MD = 48;             // staged M/D ops per read
constexpr int V3_POOL = 1024;         // inserted-base codes per batch
constexpr int V3_NREC = 253;          // insert records per batch (plane byte = record + 1; 255 = soft-clip step)
constexpr int V3_QMAX = 28;
constexpr uint8_t V3_NONE = 0xFF;
constexpr uint8_t V3_FILL_NONE = (6 << 5) | 29;   // (0,0,0,255)
constexpr uint8_t V3_FILL_INS = (7 << 5) | 29;    // (255,0,255,255)

struct __align__(8) Md3 {
    int qadj;            // M: query index of column c is qadj + c ; D: the frozen query index
    short sA;            // first column of the op (clamped below at -1)
    short endD;          // (one past the last reference column, clamped) << 1 | isD
};

struct V3S {
    uint8_t *q;                  // [rows][RS] codes per image row (pixel 0 = reference)
    uint32_t *pw;                // [slots][LS] per (read slot, column): step>=1 code | step-0 code << 8 | insert record / soft-clip marker << 16
    uint32_t *cnt;               // [NCAT][256/2] 16-bit counters (two per word) per category and (slot group, column)
    uint32_t *rowcg;             // [4][256/4] 8-bit counters (four per word): inserted bases per (slot group, image row)
    Md3 *md;                     // [V3_WARPS][V3_MD]
    uint32_t *lut;               // [256] code -> RGBA
    uint8_t *pool;               // [V3_POOL]
    uint32_t *rec;               // [256] n << 16 | pool offset
    int *cand, *j0;              // [V3_MAXC]
    int *cov, *maxins;           // [V3_CS + 8], [V3_CS]
    uint8_t *refc;               // [224] categories (A,C,G,T,N) of refSequence
    uint8_t *scode;              // [V3_MAXSLOT][2] soft-clip step codes (leading, trailing)
    short *scol;                 // [V3_MAXSLOT][2]
    unsigned short *rowstart;    // [V3_CS] first image row of a drawn column, 0xFFFF otherwise
    int *rowcnt;                 // [256] entries per image row
    int *misc;                   // [0] fallback [1] nrec [2] npool [3] Lgen [4] nrows [5] hasS [8..16] warp sums [16..24] warp sums 2
};

__device__ __forceinline__ void v3_count(const V3S &S, int cat, int gc) {      // gc = slot group * LS + column
    atomicAdd(&S.cnt[(cat * 256 + gc) >> 1], 1u << ((gc & 1) << 4));
}
// scan: one warp; coverage difference array and insert lengths of read `r`; first relevant op -> S.j0[ci]
__device__ void v3_scan_read(const EncArgs &a, const V3S &S, int ci, int qs, int L, int lane) {
    const int r = S.cand[ci];
    const ReadInfo *ri = a.info + r;
    const int ref_start = ri->ref_start, nops = ri->nops;
    const long long c0 = ri->c0;
    const long long rpb = c0 + r;
    int j = S.j0[ci];
    int j0 = -1;
    while (j < nops) {
        const int jj = j + lane;
        const bool valid = jj < nops;
        const uint32_t w = valid ? __ldg(a.R.cigar + c0 + jj) : 0u;
        const int rp = valid ? __ldg(a.RP + rpb + jj) : 0;
        const int op = w & 15, n = (int)(w >> 4);
        const int sA = ref_start + rp - qs;
        const bool md = op == 0 || op == 2;
        const bool relevant = valid && sA < L;
        const bool live = relevant && !(sA + (md ? n : 0) < 0);
        const unsigned lm = __ballot_sync(0xffffffffu, valid && !(sA + (md ? n : 0) < 0));
        if (j0 < 0 && lm) j0 = j + __ffs(lm) - 1;
        if (live) {
            if (op == 1) {
                if (sA >= 1) atomicMax(&S.maxins[sA], min(n, 512));
            } else {
                const int lo = max(sA, 0);
                const int hi = md ? min(max(sA + 1, sA + n - 1), L) : sA + 1;
                if (lo < hi) { atomicAdd(&S.cov[lo], 1); atomicAdd(&S.cov[hi], -1); }
            }
        }
        if (__popc(__ballot_sync(0xffffffffu, relevant)) < 32) break;
        j += 32;
    }
    if (lane == 0) S.j0[ci] = j0 < 0 ? nops : j0;
}

// generate: one warp; every entry of read S.cand[ci] on columns [0, Lgen) -> planes of `slot`
__device__ void v3_gen_read(const EncArgs &a, const V3S &S, int ci, int slot, int grp, int qs, int Lgen, int LS, int H,
                            bool count, int lane, int warp) {
    const int r = S.cand[ci];
    const ReadInfo *ri = a.info + r;
    const int ref_start = ri->ref_start, nops = ri->nops;
    const int aln_len = ri->aln_len, l_seq = ri->l_seq;
    const long long c0 = ri->c0;
    const long long rpb = c0 + r;
    const uint8_t *__restrict__ codep = a.code0 + ri->qual_off;
    Md3 *md = S.md + warp * V3_MD;
    uint8_t *pl = reinterpret_cast<uint8_t *>(S.pw + slot * LS);      // byte 0: step>=1, 1: step 0, 2: insert / soft clip
    const unsigned lt = (1u << lane) - 1u;
    int nmd = 0, shift = 0;
    bool bad = false;
    int j = S.j0[ci];
    for (int round = 0; j < nops; round++, j += 32) {
        const int jj = j + lane;
        const bool valid = jj < nops;
        const uint32_t w = valid ? __ldg(a.R.cigar + c0 + jj) : 0u;
        const int rp = valid ? __ldg(a.RP + rpb + jj) : 0;
        const int qp = valid ? __ldg(a.QP + c0 + jj) - 1 : 0;
        const int op = w & 15, n = (int)(w >> 4);
        const int sA = ref_start + rp - qs;
        const bool relevant = valid && sA < Lgen;
        const bool isMD = relevant && (op == 0 || op == 2);
        const unsigned z = __ballot_sync(0xffffffffu, isMD && sA == 0);
        if (round == 0) shift = z ? 1 : 0; else if (z) bad = true;
        // ---- staged M/D ops (column-parallel part below)
        const unsigned mdm = __ballot_sync(0xffffffffu, isMD);
        if (isMD) {
            const int k = nmd + __popc(mdm & lt);
            if (k < V3_MD) {
                Md3 m;
                m.qadj = op == 0 ? qp - sA + 1 : qp;
                m.sA = (short)max(sA, -1);
                m.endD = (short)((min(sA + n, 400) << 1) | (op == 2 ? 1 : 0));
                md[k] = m;
            } else bad = true;
        }
        nmd += __popc(mdm);
        // ---- op-parallel part
        if (relevant && sA >= 0) {
            if (op != 1) {
                // step 0 of an M/D op, or the single step of a soft clip: previous query base (index -1 wraps like
                // Python: last aligned base, last quality of the full query), reference character one to the left
                const int rel = sA - 1 + shift;
                if (rel < 0) bad = true;
                else {
                    int cb;
                    if (qp >= 0) cb = __ldg(codep + qp);
                    else cb = (__ldg(codep + aln_len - 1) & 0xE0) | (__ldg(codep + l_seq - 1) & 31);
                    int cat = cb >> 5;
                    if (op == 2) cat = CAT_D; else if (cat == S.refc[rel]) cat = CAT_M;
                    const uint8_t code = (uint8_t)((cat << 5) | (cb & 31));
                    if (op == 4) {
                        const int which = jj == 0 ? 0 : 1;
                        S.scode[slot * 2 + which] = code;
                        S.scol[slot * 2 + which] = (short)sA;
                        pl[sA * 4 + 2] = 255;
                    } else pl[sA * 4 + 1] = code;
                    if (count) v3_count(S, cat, grp * LS + sA);
                }
            } else if (sA >= 1) {
                // insertion anchored at column sA (Pileup.py:220-230): bases qp+1 .. qp+n
                const int rec = atomicAdd(&S.misc[1], 1);
                const int off = atomicAdd(&S.misc[2], n);
                if (rec >= V3_NREC || off + n > V3_POOL || n > 0xffff) bad = true;
                else {
                    S.rec[rec] = ((uint32_t)n << 16) | (uint32_t)off;
                    pl[sA * 4 + 2] = (uint8_t)(rec + 1);
                    for (int k2 = 0; k2 < n; k2++) S.pool[off + k2] = __ldg(codep + qp + 1 + k2);
                    const int y0 = S.rowstart[sA];
                    if (count && y0 != 0xFFFF)
                        for (int k2 = 0; k2 < n && y0 + 1 + k2 < H; k2++) {
                            const int y = y0 + 1 + k2;
                            atomicAdd(&S.rowcg[(grp * 256 + y) >> 2], 1u << ((y & 3) << 3));
                        }
                }
            }
        }
        if (__popc(__ballot_sync(0xffffffffu, relevant)) < 32) break;
    }
    __syncwarp();
    // ---- column-parallel part: steps >= 1 of the M/D ops; lane owns columns [cA, cA + cpl)
    const int cpl = (Lgen + 31) >> 5;
    const int cA = lane * cpl;
    if (nmd > V3_MD) nmd = V3_MD;
    if (nmd > 0 && cA < Lgen) {
        int jm = 0;                                   // last staged op with sA <= cA (0 if none)
#pragma unroll
        for (int step = 32; step; step >>= 1) {
            const int t = jm + step;
            if (t < nmd && md[t].sA <= cA) jm = t;
        }
        Md3 m = md[jm];
        int end = m.endD >> 1;
        const int cEnd = min(cA + cpl, Lgen);
        const uint8_t *refs = S.refc + shift;
        for (int c = cA; c < cEnd; c++) {
            if (c >= end && jm + 1 < nmd) { jm++; m = md[jm]; end = m.endD >> 1; }
            if (c >= m.sA && c < end - 1) {
                const bool isD = m.endD & 1;
                const int cb = __ldg(codep + m.qadj + (isD ? 0 : c));
                int cat = cb >> 5;
                if (isD) cat = CAT_D; else if (cat == refs[c]) cat = CAT_M;
                pl[c * 4] = (uint8_t)((cat << 5) | (cb & 31));
                if (count) v3_count(S, cat, grp * LS + c);
            }
        }
    }
    if (bad) S.misc[0] = 1;
}

// place: thread `c` owns column c (image row my_row); codes of the batch's slots in file order
__device__ __forceinline__ void v3_place(uint32_t code, uint32_t &lo, uint32_t &hi, uint8_t *qrow, int Wm1) {
    const uint32_t cat = code >> 5;
    const int pos = (int)(__byte_perm(lo, hi, cat) & 0xffu);
    if (pos < Wm1) {
        qrow[pos] = (uint8_t)code;
        const uint32_t inc = 1u << ((cat & 3) << 3);
        if (cat < 4) lo += inc; else hi += inc;
    }
}

__device__ __forceinline__ void v3_walk(const V3S &S, int c, int my_row, int s0, int s1, int LS, int RS, int W, int H,
                                        uint32_t &lo, uint32_t &hi, uint8_t *rowfill) {
    uint8_t *qrow = S.q + my_row * RS + 1;
    const uint32_t *pw = S.pw + c;
    const int Wm1 = W - 1;
    for (int s = s0; s < s1; s++) {
        const uint32_t v = pw[s * LS];
        const uint32_t c1 = v & 0xffu, c0 = (v >> 8) & 0xffu, ir = (v >> 16) & 0xffu;
        if (ir == 255u) {
            const int which = S.scol[s * 2] == c ? 0 : 1;
            v3_place(S.scode[s * 2 + which], lo, hi, qrow, Wm1);
        }
        if (c0 != V3_NONE) v3_place(c0, lo, hi, qrow, Wm1);
        if (c1 != V3_NONE) v3_place(c1, lo, hi, qrow, Wm1);
        if (ir != 0u && ir != 255u) {
            const uint32_t rc = S.rec[ir - 1];
            const int n = rc >> 16, off = rc & 0xffff;
            for (int k = 0; k < n; k++) {
                const int y = my_row + 1 + k;
                if (y >= H) break;
                const int pos = rowfill[y]++;
                if (pos < Wm1 && pos < RS - 1) S.q[y * RS + 1 + pos] = S.pool[off + k];
            }
        }
    }
}

// finish: one thread per image row sorts an insert row, adds pixel 0 and filler codes up to the quad boundary and
// writes the row's entry quads (16-byte stores through the code -> RGBA table)
__device__ __forceinline__ void v3_finish_row(const V3S &S, int y, int tot, bool ins, uint8_t ref_code, int RS, int W,
                                              uint8_t *img) {
    uint8_t *row = S.q + y * RS;
    if (ins) {                                        // stable sort by category (A,C,G,T,N) of codes appended in read order
        for (int i = 2; i <= tot; i++) {
            const uint8_t c = row[i];
            int j = i - 1;
            while (j >= 1 && (row[j] >> 5) > (c >> 5)) { row[j + 1] = row[j]; j--; }
            row[j + 1] = c;
        }
    }
    row[0] = ins ? V3_FILL_INS : ref_code;
    for (int x = tot + 1; x & 3; x++) row[x] = ins ? V3_FILL_INS : V3_FILL_NONE;
    uint8_t *orow = img + ((size_t)y * W << 2);
    for (int x0 = 0; x0 <= tot; x0 += 4) {
        const uint32_t codes = *reinterpret_cast<const uint32_t *>(row + x0);
        const uint32_t p0 = S.lut[codes & 0xffu], p1 = S.lut[(codes >> 8) & 0xffu];
        const uint32_t p2 = S.lut[(codes >> 16) & 0xffu], p3 = S.lut[codes >> 24];
        *reinterpret_cast<uint4 *>(orow + (x0 << 2)) = make_uint4(p0, p1, p2, p3);
    }
}

template <int MINB, int POOL_BYTES>
__global__ void __launch_bounds__(V3_THREADS, MINB)
encode_windows_v3_kernel(EncArgs a, int *__restrict__ fb_count, int *__restrict__ fb_list, int emit_mode,
                         unsigned long long *__restrict__ dbg) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int w = blockIdx.x;
    const int L = a.L, W = a.W, H = a.H;
    const int qs = a.win_qs[w], qe = qs + L;
    const int lo = a.win_lo[w], hi = a.win_hi[w];
    // phase clocks (debugging aid, DP_ENC_TIMING): thread 0 adds the cycles between marks to dbg[phase]
    long long tmark = dbg ? clock64() : 0;
#define V3_MARK(ph) do { if (dbg && tid == 0) { const long long t_ = clock64(); atomicAdd(&dbg[ph], (unsigned long long)(t_ - tmark)); tmark = t_; } } while (0)

    // ---- shared memory carve-up (16-byte aligned sections first)
    uint8_t *p = smem;
    V3S S;
    uint8_t *s_pool0 = p;                          p += POOL_BYTES;
    S.pw = (uint32_t *)s_pool0;
    S.q = s_pool0;                                 // placed at the top of the pool once the row stride is known
    S.cnt = (uint32_t *)p;                         p += 4 * (NCAT * 256 / 2);
    S.rowcg = (uint32_t *)p;                       p += 4 * 256;
    S.md = (Md3 *)p;                               p += sizeof(Md3) * V3_WARPS * V3_MD;
    S.lut = (uint32_t *)p;                         p += 4 * 256;
    S.rec = (uint32_t *)p;                         p += 4 * 256;
    S.pool = p;                                    p += V3_POOL;
    S.cand = (int *)p;                             p += 4 * V3_MAXC;
    S.j0 = (int *)p;                               p += 4 * V3_MAXC;
    S.cov = (int *)p;                              p += 4 * (V3_CS + 8);
    S.maxins = (int *)p;                           p += 4 * V3_CS;
    S.misc = (int *)p;                             p += 4 * 32;
    uint32_t *s_fill = (uint32_t *)p;              p += 4 * 2 * 256;       // two constant filler rows (none / insert)
    int *s_rowcnt = (int *)p;                      p += 4 * 256;           // entries per image row
    S.rowstart = (unsigned short *)p;              p += 2 * V3_CS;         // first image row of a drawn column (0xFFFF: none)
    short *s_rowcol = (short *)p;                  p += 2 * 256;
    short *s_rowk = (short *)p;                    p += 2 * 256;
    uint8_t *s_rowfill = p;                        p += 4 * 256;           // [slot group][row] running fill of insert rows
    uint8_t *s_rowref = p;                         p += 256;               // reference-pixel code of an anchor row
    S.scol = (short *)p;                           p += 2 * V3_MAXSLOT * 2;
    S.scode = p;                                   p += V3_MAXSLOT * 2;
    S.refc = p;                                    p += 224;
    int *s_wsum = &S.misc[8], *s_wsum2 = &S.misc[16];
    S.rowcnt = s_rowcnt;
    // packed bases of slot groups 1..3 [group-1][lo|hi][LS] (G * LS <= 256, so <= 2 KB): handed from a column's group-0 thread to the others
    // through the staged-op area (3 KB), which is idle between the counting pass and placement -- except in the
    // multi-batch case, where only group 0 places and the hand-over is not needed
    uint32_t *s_base = reinterpret_cast<uint32_t *>(S.md);

    // ---- setup
    const long long roff = (long long)qs - a.ref_base;
    const bool ref_ok = roff >= 0 && roff + L + 1 <= a.ref_len;
    if (tid < 32) S.misc[tid] = 0;
    __syncthreads();
    if (tid == 0 && !ref_ok) S.misc[0] = 1;
    for (int i = tid; i < L + 1; i += V3_THREADS) {
        const int ch = ref_ok ? a.ref[roff + i] : 'N';
        int code = 7;
        switch (ch) { case 'A': code = CAT_A; break; case 'C': code = CAT_C; break; case 'G': code = CAT_G; break;
                      case 'T': code = CAT_T; break; case 'N': code = CAT_N; break; default: break; }
        if (code == 7) S.misc[0] = 1;                // KeyError at Pileup.py:281 if it becomes a reference pixel
        S.refc[i] = (uint8_t)code;
    }
    {
        // code -> RGBA (SNPtoRGB, Pileup.py:56-64); category 7 = insert colour; index 29.. = alpha 255
        const uint32_t rgb[8] = {0x000000FFu, 0x0000FFFFu, 0x0000FF00u, 0x00FF0000u, 0x00FFFF00u, 0x00FFFFFFu, 0x00000000u, 0x00FF00FFu};
        const int qi = tid & 31;
        uint32_t col = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) if ((tid >> 5) == k) col = rgb[k];
        S.lut[tid] = col | ((uint32_t)(qi <= V3_QMAX ? c_qlut[qi] : 255) << 24);
    }
    for (int i = tid; i < NCAT * 256 / 2; i += V3_THREADS) S.cnt[i] = 0;
    S.rowcg[tid] = 0;
    for (int i = tid; i < V3_CS + 8; i += V3_THREADS) S.cov[i] = 0;
    for (int i = tid; i < V3_CS; i += V3_THREADS) { S.maxins[i] = 0; S.rowstart[i] = 0xFFFF; }
    s_rowcnt[tid] = 0; s_rowcol[tid] = -1; s_rowk[tid] = 0; s_rowref[tid] = V3_FILL_NONE;
    s_fill[tid] = 0xFF000000u; s_fill[256 + tid] = 0xFFFF00FFu;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");

    // ---- ordered compaction of the reads that pysam.fetch would return and parseRead would keep
    int ncand = 0;
    for (int cb = lo; cb < hi; cb += V3_THREADS) {
        const int r = cb + tid;
        const bool keep = r < hi && a.R.pos[r] < qe && a.endk[r] > qs;
        if (keep) asm volatile("prefetch.global.L1 [%0];" ::"l"(a.info + r));      // read by the digest below
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        __syncthreads();
        if (lane == 0) s_wsum[warp] = __popc(bal);
        __syncthreads();
        int off = ncand, tot = 0;
#pragma unroll
        for (int k = 0; k < V3_WARPS; k++) { const int c = s_wsum[k]; if (k < warp) off += c; tot += c; }
        const int idx = off + __popc(bal & ((1u << lane) - 1u));
        if (keep && idx < V3_MAXC) S.cand[idx] = r;
        ncand += tot;
    }
    __syncthreads();
    if (ncand > V3_MAXC) { if (tid == 0) S.misc[0] = 1; ncand = 0; }
    V3_MARK(0);
    // ---- per read: clean?  first op to look at (checkpoint of the read's CIGAR)
    if (tid < ncand) {
        const int r = S.cand[tid];
        const ReadInfo *ri = a.info + r;
        const int flags = ri->flags;
        if (!(flags & RI_CLEAN)) S.misc[0] = 1;
        const int T = qs - ri->ref_start;
        int jc = 0;
        if (T > 0 && (flags & RI_CLEAN)) {
            const int k = (flags >> 8) & 31;
            jc = __ldg(a.ckpt + (ri->c0 >> 1) + r + (min(T, ri->span) >> k));
        }
        S.j0[tid] = jc;
        if ((flags & RI_CLEAN) && jc < ri->nops) {    // the scan and generate passes start here: pull the lines towards L1
            const long long c0 = ri->c0;
            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.R.cigar + c0 + jc));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.RP + c0 + r + jc));
            asm volatile("prefetch.global.L1 [%0];" ::"l"(a.QP + c0 + jc));
        }
    }
    __syncthreads();
    if (S.misc[0]) ncand = 0;                         // bail out early: the exact kernel takes the window
    V3_MARK(1);
    // ---- scan: coverage and insert lengths -> row map
    for (;;) {                                        // warps draw reads from a shared counter (uneven CIGAR density)
        int ci = 0;
        if (lane == 0) ci = atomicAdd(&S.misc[6], 1);
        ci = __shfl_sync(0xffffffffu, ci, 0);
        if (ci >= ncand) break;
        v3_scan_read(a, S, ci, qs, L, lane);
    }
    __syncthreads();
    V3_MARK(2);
    {
        const int cv = tid < L ? S.cov[tid] : 0;
        const int inc1 = warp_incl_scan(cv, lane);
        if (lane == 31) s_wsum[warp] = inc1;
        __syncthreads();
        int b1 = 0;
        for (int k = 0; k < warp; k++) b1 += s_wsum[k];
        const bool pop = tid < L && (b1 + inc1) > 0;
        const int mi = pop ? S.maxins[tid] : 0;
        const int v = pop ? min(1 + mi, 300) | (1 << 16) : 0;          // rows here | populated (two scans in one word)
        const int inc2 = warp_incl_scan(v, lane);
        if (lane == 31) s_wsum2[warp] = inc2;
        __syncthreads();
        int b2 = 0;
        for (int k = 0; k < warp; k++) b2 += s_wsum2[k];
        const int excl = b2 + inc2 - v;
        const int rstart = excl & 0xffff, ord = excl >> 16;
        if (tid == V3_THREADS - 1) S.misc[4] = (b2 + inc2) & 0xffff;
        if (pop && rstart < H) {
            const int nown = min(1 + mi, H - rstart);
            S.rowstart[tid] = (unsigned short)rstart;
            atomicMax(&S.misc[3], tid + 1);
            s_rowref[rstart] = (uint8_t)((S.refc[ord] << 5) | 29);
            for (int k = 0; k < nown; k++) { s_rowcol[rstart + k] = (short)tid; s_rowk[rstart + k] = (short)k; }
        }
        __syncthreads();
    }
    V3_MARK(3);
    const int Lgen = S.misc[3];
    const int nrows = min(S.misc[4], H);
    // plane geometry for this window: stride = drawn columns rounded up, as many read slots as the pool holds
    const int LS = max(16, (Lgen + 15) & ~15);
    int slots = min(V3_MAXSLOT, POOL_BYTES / (4 * LS));   // counting pass: the whole pool is planes
    int nb = (ncand + slots - 1) / slots;
    // placement runs on G slot groups per column in parallel (thread = group * LS + column): counts are kept per group
    const int G = nb == 1 ? max(1, min(4, V3_THREADS / LS)) : 1;
    const int GS = max(1, (ncand + G - 1) / G);
    uint8_t *img = a.out_img + (size_t)w * H * W * 4;

    // ---- generate (counting).  One batch: its planes are kept for placement.  More: count over all batches first,
    // then each batch is generated again and placed.
    for (int b = 0; b < nb; b++) {
        const int nslots = min(slots, ncand - b * slots);
        {
            uint4 *q4 = reinterpret_cast<uint4 *>(S.pw);
            const int n16 = nslots * LS / 4;
            for (int i = tid; i < n16; i += V3_THREADS) q4[i] = make_uint4(0xFFFFu, 0xFFFFu, 0xFFFFu, 0xFFFFu);
            if (tid == 0) { S.misc[1] = 0; S.misc[2] = 0; S.misc[7] = 0; }
            if (tid < nslots * 2) S.scol[tid] = -1;
        }
        __syncthreads();
        for (;;) {
            int s = 0;
            if (lane == 0) s = atomicAdd(&S.misc[7], 1);
            s = __shfl_sync(0xffffffffu, s, 0);
            if (s >= nslots) break;
            v3_gen_read(a, S, b * slots + s, s, s / GS, qs, Lgen, LS, H, true, lane, warp);
        }
        __syncthreads();
    }
    V3_MARK(4);
    // ---- counts -> first slot of each category (sortingKey order A,C,G,T,D,M,N) for every (group, column) thread,
    // saturated at W; totals per image row; row stride of S.q
    const int wg = tid / LS, wc = tid - wg * LS;
    int my_row = -1, nown = 0;
    if (wg < G && wc < Lgen && S.rowstart[wc] != 0xFFFF) {
        my_row = S.rowstart[wc];
        nown = min(1 + S.maxins[wc], H - my_row);
    }
    const bool walker = my_row >= 0;
    uint32_t run_lo = 0, run_hi = 0;
    if (walker && wg == 0) {
        // the column's group-0 thread derives the packed bases of every group and the totals of the column's rows
        uint32_t blo[4] = {0, 0, 0, 0}, bhi[4] = {0, 0, 0, 0};
        int acc = 0;
#pragma unroll
        for (int c = 0; c < NCAT; c++) {
            int before = 0;
#pragma unroll
            for (int g = 0; g < 4; g++) {
                if (g < G) {
                    const int gc = c * 256 + g * LS + wc;
                    const int n = (S.cnt[gc >> 1] >> ((gc & 1) << 4)) & 0xffff;
                    const uint32_t bv = (uint32_t)min(acc + before, W - 1);
                    if (c < 4) blo[g] |= bv << (c << 3); else bhi[g] |= bv << ((c - 4) << 3);
                    before += n;
                }
            }
            acc += before;
        }
        run_lo = blo[0]; run_hi = bhi[0];
#pragma unroll
        for (int g = 1; g < 4; g++)
            if (g < G) { s_base[(g - 1) * 2 * LS + wc] = blo[g]; s_base[((g - 1) * 2 + 1) * LS + wc] = bhi[g]; }
        int mx = min(acc, W - 1);
        s_rowcnt[my_row] = mx;
        for (int k = 1; k < nown; k++) {             // insert rows of this column
            const int y = my_row + k;
            int tot = 0;
            for (int g = 0; g < G; g++) {
                s_rowfill[g * 256 + y] = (uint8_t)tot;
                tot += (S.rowcg[(g * 256 + y) >> 2] >> ((y & 3) << 3)) & 0xff;
            }
            s_rowcnt[y] = tot;
            if (tot > W - 1) S.misc[0] = 1;           // deeper than the image: the category sort needs every entry
            mx = max(mx, min(tot, W - 1));
        }
        atomicMax(&S.misc[9], mx);
    }
    __syncthreads();
    if (walker && wg > 0) { run_lo = s_base[(wg - 1) * 2 * LS + wc]; run_hi = s_base[((wg - 1) * 2 + 1) * LS + wc]; }
    const int RS = (S.misc[9] + 4) & ~3;             // pixel 0 + entries, rounded up to whole quads
    const int qbytes = (nrows * RS + 15) & ~15;
    S.q = s_pool0 + POOL_BYTES - qbytes;
    // the counting pass's planes can be placed from directly when every read was in them and they clear the rows
    const bool multi = !(nb == 1 && ncand * 4 * LS + qbytes <= POOL_BYTES);
    if (multi) {
        slots = min(V3_MAXSLOT, (POOL_BYTES - qbytes) / (4 * LS));
        if (slots < 8 && slots < ncand) { if (tid == 0) S.misc[0] = 1; slots = 8; }
        nb = (ncand + slots - 1) / slots;
    }
    __syncthreads();
    if (S.misc[0]) {
        if (tid == 0) {
            const int k = atomicAdd(fb_count, 1);
            fb_list[k] = w;
            if (a.out_status) a.out_status[w] = 0;
        }
        return;                                      // the exact kernel recomputes this window
    }
    V3_MARK(5);
    // ---- emit 1: the filler tail of every row (and blank rows) -- one bulk copy (TMA) per row out of two constant
    // filler rows in shared memory, issued by the high threads while the low ones place; emit_mode 1 = plain stores
    if (emit_mode == 0) {
        const int y = V3_THREADS - 1 - tid;
        if (y < H) {
            const int tot = y < nrows ? s_rowcnt[y] : -1;
            const int xq = ((tot + 4) >> 2) << 2;    // first pixel of the first all-filler quad
            const int bytes = (W - xq) << 2;
            if (bytes > 0) {
                const uint32_t src = (uint32_t)__cvta_generic_to_shared(s_fill + ((y < nrows && s_rowk[y] != 0) ? 256 : 0));
                uint8_t *dst = img + (((size_t)y * W + xq) << 2);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             :: "l"(dst), "r"(src), "r"(bytes) : "memory");
            }
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    } else {
        const int qpr = W >> 2;
        const int rpi = V3_THREADS / qpr;            // rows per iteration
        const int ty = tid / qpr, tx = tid - ty * qpr;
        if (ty < rpi) {
            const int x0 = tx << 2;
            for (int y = ty; y < H; y += rpi) {
                int tot = -1;
                uint32_t fill = 0xFF000000u;
                if (y < nrows) { tot = s_rowcnt[y]; if (s_rowk[y] != 0) fill = 0xFFFF00FFu; }
                if (x0 > tot) *reinterpret_cast<uint4 *>(img + (((size_t)y * W + x0) << 2)) = make_uint4(fill, fill, fill, fill);
            }
        }
    }
    V3_MARK(6);
    // ---- place
    if (!multi) {
        if (walker) v3_walk(S, wc, my_row, wg * GS, min((wg + 1) * GS, ncand), LS, RS, W, H, run_lo, run_hi, s_rowfill + wg * 256);
    } else {
        for (int b = 0; b < nb; b++) {
            const int nslots = min(slots, ncand - b * slots);
            __syncthreads();
            {
                uint4 *q4 = reinterpret_cast<uint4 *>(S.pw);
                const int n16 = nslots * LS / 4;
                for (int i = tid; i < n16; i += V3_THREADS) q4[i] = make_uint4(0xFFFFu, 0xFFFFu, 0xFFFFu, 0xFFFFu);
                if (tid == 0) { S.misc[1] = 0; S.misc[2] = 0; S.misc[7] = 0; }
                if (tid < nslots * 2) S.scol[tid] = -1;
            }
            __syncthreads();
            for (;;) {
                int s = 0;
                if (lane == 0) s = atomicAdd(&S.misc[7], 1);
                s = __shfl_sync(0xffffffffu, s, 0);
                if (s >= nslots) break;
                v3_gen_read(a, S, b * slots + s, s, 0, qs, Lgen, LS, H, false, lane, warp);
            }
            __syncthreads();
            if (walker && wg == 0) v3_walk(S, wc, my_row, 0, nslots, LS, RS, W, H, run_lo, run_hi, s_rowfill);
        }
    }
    __syncthreads();
    V3_MARK(7);
    // ---- finish: one thread per image row
    if (tid < nrows) v3_finish_row(S, tid, s_rowcnt[tid], s_rowk[tid] != 0, s_rowref[tid], RS, W, img);
    for (int y = tid; y < H; y += V3_THREADS) {
        if (a.out_rowpos) a.out_rowpos[(size_t)w * H + y] = (y < nrows) ? qs + s_rowcol[y] : -1;
        if (a.out_rowins) a.out_rowins[(size_t)w * H + y] = (y < nrows) ? s_rowk[y] : 0;
    }
    if (tid == 0 && a.out_status) a.out_status[w] = 0;
    V3_MARK(8);
    if (emit_mode == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // shared memory must outlive the copies
    V3_MARK(9);
#undef V3_MARK
}

// ================================================================================================
// v4 fast path: ONE WARP PER WINDOW, column sweep.  (v3 above spends ~38 k warp instructions and a dozen CTA barriers on
// a window whose real content is ~2 300 entries: the passes over shared-memory planes, not the data, set its pace.)
// Here lane = read (NR rounds of 32 reads); all lanes walk the window's columns left to right in lock step, each with
// the state of its read's CIGAR (current M/D op, the op after it, query index) in registers.  At a column every lane
// knows its own entries -- soft-clip step, step 0 of an op that starts here, step >= 1 of the op in progress
// (parseRead's closed form, header of this file) -- and a packed warp prefix sum over the 7 categories turns them into
// row positions: base[category] + entries of that category in earlier reads (+ earlier steps of the same read), i.e.
// sortingKey order, then read order, then step order (Pileup.py:66-73, 268-292) with no sort, no atomics and no barrier
// wider than a warp.  The row is assembled as code bytes in a 256-byte shared-memory line, turned into RGBA through the
// code table and stored with 16-byte stores; the all-filler tail of the row leaves the SM as one bulk copy from a
// constant row.  Insert rows of an anchor column follow immediately (Pileup.py:305-317), and the sweep stops when the
// image is full (150 rows: ~70 of the 201 columns at 30x with 3 % insertions) -- the row map needs no pass of its own.
// Windows outside the envelope (a non-clean read, more than 32 * NR reads, an illegal reference character) go to the
// list of the exact kernel, as in v3.
constexpr int V4_WARPS = 8;                    // windows per CTA
constexpr int V4_ROWB = 256;                   // bytes of a row-code line (W <= 256)

struct V4Read {                                // per (lane, round) walk state
    const uint32_t *cig;                       // next op word to fetch
    const uint8_t *codep;                      // code bytes of the read (code0 + qual_off)
    uint32_t wnext;                            // the op after the current one (prefetched); valid while nleft > 0
    int nleft;                                 // ops not yet started (wnext is the first of them)
    int cs, ce;                                // current M/D op covers columns [cs, ce); next op starts at column ce
    int cq;                                    // query index before the current op (parseRead's lagging readPos)
    int qnext;                                 // query index before the next op
    int flags;                                 // bit 0: current op is D; bit 1: shift; bit 2: read present
    int aln_len, l_seq;
};

__device__ __forceinline__ uint64_t v4_scan_packed(uint64_t v, int lane) {      // inclusive warp prefix sum, 7 x 8-bit fields
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint64_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}
__device__ __forceinline__ int v4_field(uint64_t v, int cat) { return (int)((v >> (cat << 3)) & 0xffu); }
// per-byte saturating add of packed counters (fields stay <= 255: anything >= W - 1 is off the image anyway)
__device__ __forceinline__ uint64_t v4_addsat(uint64_t a, uint64_t b) {
    const uint32_t lo = __vaddus4((uint32_t)a, (uint32_t)b), hi = __vaddus4((uint32_t)(a >> 32), (uint32_t)(b >> 32));
    return ((uint64_t)hi << 32) | lo;
}

template <int NR>
__global__ void __launch_bounds__(V4_WARPS * 32)
encode_windows_v4_kernel(EncArgs a, int n_windows, int *__restrict__ fb_count, int *__restrict__ fb_list) {
    __shared__ uint32_t s_lut[256];                          // code -> RGBA
    __shared__ __align__(16) uint32_t s_fill[2][256];        // constant filler rows: (0,0,0,255) / insert colour
    __shared__ __align__(16) uint8_t s_row[V4_WARPS][2][V4_ROWB];
    __shared__ uint8_t s_refc[V4_WARPS][224];
    __shared__ int s_cand[V4_WARPS][NR * 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int L = a.L, W = a.W, H = a.H;
    {
        const uint32_t rgb[8] = {0x000000FFu, 0x0000FFFFu, 0x0000FF00u, 0x00FF0000u, 0x00FFFF00u, 0x00FFFFFFu, 0x00000000u, 0x00FF00FFu};
        uint32_t col = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) if ((tid >> 5) == k) col = rgb[k];
        const int qi = tid & 31;
        s_lut[tid] = col | ((uint32_t)(qi <= V3_QMAX ? c_qlut[qi] : 255) << 24);
        s_fill[0][tid] = 0xFF000000u; s_fill[1][tid] = 0xFFFF00FFu;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int w = blockIdx.x * V4_WARPS + warp;
    if (w >= n_windows) return;                              // warp-uniform; no CTA barrier below
    const unsigned lt = (1u << lane) - 1u;
    const int qs = a.win_qs[w], qe = qs + L;
    const int lo = a.win_lo[w], hi = a.win_hi[w];
    uint8_t *refc = s_refc[warp];
    int *cand = s_cand[warp];
    bool bad = false;
    // ---- reference categories of the window (Pileup.py:41: L + 1 characters)
    const long long roff = (long long)qs - a.ref_base;
    const bool ref_ok = roff >= 0 && roff + L + 1 <= a.ref_len;
    if (!ref_ok) bad = true;
    for (int i = lane; i < L + 1; i += 32) {
        const int ch = ref_ok ? a.ref[roff + i] : 'N';
        int code = 7;
        switch (ch) { case 'A': code = CAT_A; break; case 'C': code = CAT_C; break; case 'G': code = CAT_G; break;
                      case 'T': code = CAT_T; break; case 'N': code = CAT_N; break; default: break; }
        if (code == 7) bad = true;                           // KeyError at Pileup.py:281 if it becomes a reference pixel
        refc[i] = (uint8_t)code;
    }
    // ---- ordered compaction of the reads pysam.fetch returns and parseRead keeps
    int ncand = 0;
    for (int cb = lo; cb < hi; cb += 32) {
        const int r = cb + lane;
        const bool keep = r < hi && a.R.pos[r] < qe && a.endk[r] > qs;
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        const int idx = ncand + __popc(bal & lt);
        if (keep && idx < NR * 32) cand[idx] = r;
        ncand += __popc(bal);
    }
    if (ncand > NR * 32) bad = true;
    __syncwarp();
    // ---- per read: digest, entry op (checkpoint + short walk), walk state
    V4Read rd[NR];
#pragma unroll
    for (int r_ = 0; r_ < NR; r_++) {
        V4Read &R = rd[r_];
        R.flags = 0; R.nleft = 0; R.cs = INT_MAX / 2; R.ce = INT_MAX; R.cq = 0; R.qnext = 0; R.wnext = 0;
        R.cig = nullptr; R.codep = nullptr; R.aln_len = 0; R.l_seq = 0;
        const int ci = r_ * 32 + lane;
        if (!bad && ci < ncand) {
            const int r = cand[ci];
            const ReadInfo *ri = a.info + r;
            const int flags = ri->flags;
            if (!(flags & RI_CLEAN)) { bad = true; continue; }
            const int ref_start = ri->ref_start, nops = ri->nops;
            const long long c0 = ri->c0;
            R.codep = a.code0 + ri->qual_off;
            R.aln_len = ri->aln_len; R.l_seq = ri->l_seq;
            const int T = qs - ref_start;
            int j = 0;
            if (T > 0) j = __ldg(a.ckpt + (c0 >> 1) + r + (min(T, ri->span) >> ((flags >> 8) & 31)));
            // first op that is not wholly left of the window: sA + (M/D ? n : 0) >= 0
            int sA = 0, n = 0, op = 0;
            uint32_t wd = 0;
            for (; j < nops; j++) {
                wd = __ldg(a.R.cigar + c0 + j);
                op = wd & 15; n = (int)(wd >> 4);
                sA = ref_start + __ldg(a.RP + c0 + r + j) - qs;
                if (sA + ((op == 0 || op == 2) ? n : 0) >= 0) break;
            }
            if (j >= nops) continue;                         // nothing of this read reaches the window (flags stay 0)
            const int qp = __ldg(a.QP + c0 + j) - 1;
            R.flags = 4;
            if ((op == 0 || op == 2) && sA < 0) {            // an M/D op in progress at the window's left edge
                R.cs = sA; R.ce = sA + n; R.cq = qp;
                R.flags |= (op == 2) ? 1 : 0;
                R.qnext = qp + (op == 0 ? n : 0);
                R.nleft = nops - j - 1;
                R.cig = a.R.cigar + c0 + j + 1;
            } else {                                         // the op starts inside the window: it is the "next" one
                R.ce = sA; R.qnext = qp;
                R.nleft = nops - j;
                R.cig = a.R.cigar + c0 + j;
            }
            if (R.nleft > 0) { R.wnext = __ldg(R.cig); R.cig++; }
        }
    }
    bad = __any_sync(0xffffffffu, bad);
    if (bad) {
        if (lane == 0) {
            const int k = atomicAdd(fb_count, 1);
            fb_list[k] = w;
            if (a.out_status) a.out_status[w] = 0;
        }
        return;                                              // the exact kernel recomputes this window
    }
    uint8_t *img = a.out_img + (size_t)w * H * W * 4;
    int32_t *rowpos = a.out_rowpos ? a.out_rowpos + (size_t)w * H : nullptr;
    int32_t *rowins = a.out_rowins ? a.out_rowins + (size_t)w * H : nullptr;
    const int Wm1 = W - 1;
    int y = 0, ord = 0, rb = 0;

    // emits image row y from the code line `line` holding pixel 0 + `tot` entries (tot <= W - 1)
    auto emit_row = [&](const uint8_t *line, int tot, bool ins) {
        __syncwarp();
        const int nq = (tot + 4) >> 2;                       // quads holding pixel 0 .. tot
        const uint8_t fillc = ins ? V3_FILL_INS : V3_FILL_NONE;
        uint8_t *orow = img + ((size_t)y * W << 2);
        for (int q0 = 0; q0 < nq; q0 += 32) {
            const int q = q0 + lane;
            if (q < nq) {
                uint32_t codes = *reinterpret_cast<const uint32_t *>(line + (q << 2));
                const int over = (q << 2) + 3 - tot;         // bytes of this quad beyond the last entry
                if (over > 0) {
                    const uint32_t keep = 0xFFFFFFFFu >> (over << 3);
                    codes = (codes & keep) | ((fillc * 0x01010101u) & ~keep);
                }
                const uint32_t p0 = s_lut[codes & 0xffu], p1 = s_lut[(codes >> 8) & 0xffu];
                const uint32_t p2 = s_lut[(codes >> 16) & 0xffu], p3 = s_lut[codes >> 24];
                *reinterpret_cast<uint4 *>(orow + (q << 4)) = make_uint4(p0, p1, p2, p3);
            }
        }
        if (lane == 0) {
            const int bytes = (W - (nq << 2)) << 2;
            if (bytes > 0) {
                const uint32_t src = (uint32_t)__cvta_generic_to_shared(&s_fill[ins ? 1 : 0][0]);
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             :: "l"(orow + (nq << 4)), "r"(src), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
    };

    for (int c = 0; c < L && y < H; c++) {
        // ---- every read's entries at column c
        uint32_t eS[NR], e0[NR], e1[NR];                     // soft-clip step, step 0, step >= 1 (V3_NONE: none)
        int insn[NR], insq[NR];                              // insertion anchored here: length, first query index
        uint64_t mine[NR];                                   // packed per-category counts of this lane's entries
        uint64_t total = 0;
        int anyins = 0;
#pragma unroll
        for (int r_ = 0; r_ < NR; r_++) {
            V4Read &R = rd[r_];
            eS[r_] = e0[r_] = e1[r_] = V3_NONE; insn[r_] = 0; insq[r_] = 0; mine[r_] = 0;
            if (R.flags & 4) {
                if (c == R.ce) {
                    // ops that start at this column, in CIGAR order: [S] [I] then an M/D op, or a trailing S
                    bool started = false;
#pragma unroll 1
                    while (R.nleft > 0 && !started) {
                        const uint32_t wd = R.wnext;
                        const int op = wd & 15, n = (int)(wd >> 4);
                        const int qp = R.qnext;
                        R.nleft--;
                        if (R.nleft > 0) { R.wnext = __ldg(R.cig); R.cig++; }
                        if (op == 1) {                       // insertion anchored at c (Pileup.py:220-230: needs A > queryStart)
                            if (c >= 1) { insn[r_] = n; insq[r_] = qp + 1; }
                            R.qnext = qp + n;
                            continue;
                        }
                        // an M/D op starts exactly at queryStart (a leading soft clip at column 0 is followed by one): the
                        // read's reference characters are shifted by one for the rest of the window
                        if (c == 0 && (op == 0 || op == 2 || (op == 4 && R.nleft > 0))) R.flags |= 2;
                        // step 0 of an M/D op / the single step of a soft clip: previous query base (index -1 wraps like
                        // Python), reference character one to the left (+ shift)
                        const int rel = c - 1 + ((R.flags >> 1) & 1);
                        if (rel < 0) { bad = true; break; }
                        int cb;
                        if (qp >= 0) cb = __ldg(R.codep + qp);
                        else cb = (__ldg(R.codep + R.aln_len - 1) & 0xE0) | (__ldg(R.codep + R.l_seq - 1) & 31);
                        int cat = cb >> 5;
                        if (op == 2) cat = CAT_D; else if (cat == refc[rel]) cat = CAT_M;
                        const uint32_t code = (uint32_t)((cat << 5) | (cb & 31));
                        if (op == 4) { eS[r_] = code; continue; }              // S advances nothing; a following M starts here too
                        e0[r_] = code;
                        R.cs = c; R.ce = c + n; R.cq = qp;
                        R.flags = (R.flags & ~1) | (op == 2 ? 1 : 0);
                        R.qnext = qp + (op == 0 ? n : 0);
                        started = true;
                    }
                    if (!started) R.ce = INT_MAX;            // no further M/D op: the read ends here
                }
                if (c >= R.cs && c < R.ce - 1 && R.ce != INT_MAX) {            // step >= 1 of the op in progress
                    const bool isD = R.flags & 1;
                    const int cb = __ldg(R.codep + (isD ? R.cq : R.cq + 1 + (c - R.cs)));
                    int cat = cb >> 5;
                    if (isD) cat = CAT_D; else if (cat == refc[c + ((R.flags >> 1) & 1)]) cat = CAT_M;
                    e1[r_] = (uint32_t)((cat << 5) | (cb & 31));
                }
                uint64_t m = 0;
                if (eS[r_] != V3_NONE) m += 1ull << ((eS[r_] >> 5) << 3);
                if (e0[r_] != V3_NONE) m += 1ull << ((e0[r_] >> 5) << 3);
                if (e1[r_] != V3_NONE) m += 1ull << ((e1[r_] >> 5) << 3);
                mine[r_] = m;
            }
            anyins |= insn[r_];
        }
        if (__any_sync(0xffffffffu, bad)) break;
        // ---- ranks: inclusive packed scan per round, rounds chained by the (saturating) totals
        uint64_t incl[NR], before[NR];
#pragma unroll
        for (int r_ = 0; r_ < NR; r_++) {
            incl[r_] = v4_scan_packed(mine[r_], lane);
            before[r_] = total;
            total = v4_addsat(total, __shfl_sync(0xffffffffu, incl[r_], 31));
        }
        int tot = 0;
        uint64_t base = 0;                                   // first position of each category
#pragma unroll
        for (int k = 0; k < NCAT; k++) {
            base |= (uint64_t)min(tot, 255) << (k << 3);
            tot += v4_field(total, k);
        }
        const bool haveins = __any_sync(0xffffffffu, anyins != 0);
        if (tot == 0) continue;                              // unpopulated column: it vanishes, its insertions with it
        // ---- anchor row
        uint8_t *line = s_row[warp][rb];
        rb ^= 1;
#pragma unroll
        for (int r_ = 0; r_ < NR; r_++) {
            if (mine[r_] == 0) continue;
            const uint64_t ex = v4_addsat(before[r_], incl[r_] - mine[r_]);     // same-category entries of earlier reads
            int extra[NCAT];
#pragma unroll
            for (int k = 0; k < NCAT; k++) extra[k] = 0;
            const uint32_t es[3] = {eS[r_], e0[r_], e1[r_]};
#pragma unroll
            for (int t = 0; t < 3; t++) {
                if (es[t] == V3_NONE) continue;
                const int cat = es[t] >> 5;
                int own = 0;
#pragma unroll
                for (int k = 0; k < NCAT; k++) if (k == cat) { own = extra[k]; extra[k]++; }
                const int pos = v4_field(base, cat) + v4_field(ex, cat) + own;
                if (pos < Wm1) line[1 + pos] = (uint8_t)es[t];
            }
        }
        if (lane == 0) {
            line[0] = (uint8_t)((refc[ord] << 5) | 29);
            if (rowpos) rowpos[y] = qs + c;
            if (rowins) rowins[y] = 0;
        }
        emit_row(line, min(tot, Wm1), false);
        y++; ord++;
        // ---- insert rows of this anchor (Pileup.py:305-317): row k holds base k of every insertion longer than k,
        // categories A,C,G,T,N, read order inside a category; reference pixel and filler in the insert colour
        if (haveins) {
            int mx = 0;
#pragma unroll
            for (int r_ = 0; r_ < NR; r_++) mx = max(mx, insn[r_]);
            mx = __reduce_max_sync(0xffffffffu, mx);
            for (int k = 0; k < mx && y < H; k++) {
                uint32_t ec[NR];
                uint64_t im[NR], iincl[NR], ibefore[NR];
                uint64_t itotal = 0;
#pragma unroll
                for (int r_ = 0; r_ < NR; r_++) {
                    ec[r_] = V3_NONE; im[r_] = 0;
                    if (insn[r_] > k) {
                        ec[r_] = __ldg(rd[r_].codep + insq[r_] + k);
                        im[r_] = 1ull << ((ec[r_] >> 5) << 3);
                    }
                    iincl[r_] = v4_scan_packed(im[r_], lane);
                    ibefore[r_] = itotal;
                    itotal = v4_addsat(itotal, __shfl_sync(0xffffffffu, iincl[r_], 31));
                }
                int itot = 0;
                uint64_t ibase = 0;
#pragma unroll
                for (int kk = 0; kk < NCAT; kk++) {
                    ibase |= (uint64_t)min(itot, 255) << (kk << 3);
                    itot += v4_field(itotal, kk);
                }
                uint8_t *iline = s_row[warp][rb];
                rb ^= 1;
#pragma unroll
                for (int r_ = 0; r_ < NR; r_++) {
                    if (ec[r_] == V3_NONE) continue;
                    const int cat = ec[r_] >> 5;
                    const uint64_t ex = v4_addsat(ibefore[r_], iincl[r_] - im[r_]);
                    const int pos = v4_field(ibase, cat) + v4_field(ex, cat);
                    if (pos < Wm1) iline[1 + pos] = (uint8_t)ec[r_];
                }
                if (lane == 0) {
                    iline[0] = V3_FILL_INS;
                    if (rowpos) rowpos[y] = qs + c;
                    if (rowins) rowins[y] = k + 1;
                }
                emit_row(iline, min(itot, Wm1), true);
                y++;
            }
        }
    }
    if (__any_sync(0xffffffffu, bad)) {
        // a rule of the closed form broke mid-sweep (reference index -1): the rows written so far are overwritten by the
        // exact kernel, which takes the whole window
        if (lane == 0) {
            const int k = atomicAdd(fb_count, 1);
            fb_list[k] = w;
            if (a.out_status) a.out_status[w] = 0;
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        return;
    }
    // ---- blank rows below the pileup: (0,0,0,255)
    for (int yy = y + lane; yy < H; yy += 32) {
        const uint32_t src = (uint32_t)__cvta_generic_to_shared(&s_fill[0][0]);
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     :: "l"(img + ((size_t)yy * W << 2)), "r"(src), "r"(W << 2) : "memory");
        if (rowpos) rowpos[yy] = -1;
        if (rowins) rowins[yy] = 0;
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    if (lane == 0 && a.out_status) a.out_status[w] = 0;
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");      // the constant rows must outlive the copies
}

static size_t encoder_v3_smem_bytes(const dp_encode_params *p, int pool_bytes) {
    (void)p;
    size_t s = 0;
    s += pool_bytes;
    s += 4 * (NCAT * 256 / 2) + 4 * 256;                                   // cnt, rowcg
    s += sizeof(Md3) * V3_WARPS * V3_MD;
    s += 4 * 256 * 2 + V3_POOL + 4 * V3_MAXC * 2 + 4 * (V3_CS + 8) + 4 * V3_CS + 4 * 32;
    s += 4 * 2 * 256 + 4 * 256 + 2 * V3_CS + 2 * 256 * 2 + 4 * 256 + 256;  // fill, rowcnt, rowstart, rowcol/rowk, rowfill, rowref
    s += 2 * V3_MAXSLOT * 2 + V3_MAXSLOT * 2 + 224;
    return s + 16;
}

static size_t encoder_smem_bytes(const dp_encode_params *p) {
    const int L = 2 * p->flank + 1;
    size_t s = 0;
    s += sizeof(SegOp) * ENC_WARPS * SEG_CAP;
    s += sizeof(SegHdr) * ENC_WARPS;
    s += sizeof(int) * CAND_BLOCK;
    s += sizeof(int) * (ENC_THREADS + 4);
    s += sizeof(int) * 16;
    s += sizeof(short) * ENC_THREADS * 2;
    s += sizeof(unsigned short) * ENC_THREADS * 8;
    s += sizeof(unsigned short) * ENC_THREADS;
    s += ((L + 1 + 15) / 16) * 16;
    s += ENC_THREADS;
    s += (size_t)p->window_cutoff * p->coverage_cutoff;
    return s + 16;
}

}  // namespace dp

// ================================================================================================
namespace dp {

// Workspace of one read batch.  Everything up to `fblist` depends on the READS only, so a batch prepared once
// (dp_encode_prepare) serves any number of dp_encode_prepared / dp_window_ranges calls of up to `n_windows` windows.
struct WsLayout {
    int64_t hdr, aux, RP, QP, info, ckpt, endk, code0, dbg, pmax, tmax, fblist, total;
};
constexpr int PM_THREADS = 256, PM_PER = 4, PM_TILE = PM_THREADS * PM_PER;     // prefix-maximum tiles of 1024 reads

static WsLayout ws_layout(int32_t n_reads, int64_t n_cigar, int64_t n_bases, int32_t n_windows) {
    WsLayout l;
    const int64_t nr = n_reads > 0 ? n_reads : 1;
    int64_t b = 0;
    l.hdr = b;    b += 256;                                                                   /* word 0: fallback count */
    l.aux = b;    b += align_up((int64_t)sizeof(ReadAux) * nr, 256);
    l.RP = b;     b += align_up((int64_t)sizeof(int) * (n_cigar + n_reads + 1), 256);          /* nops+1 per read */
    l.QP = b;     b += align_up((int64_t)sizeof(int) * (n_cigar + 1), 256);
    l.info = b;   b += align_up((int64_t)sizeof(ReadInfo) * nr, 256);
    l.ckpt = b;   b += align_up((int64_t)sizeof(int) * ((n_cigar >> 1) + n_reads + 2), 256);   /* checkpoints */
    l.endk = b;   b += align_up((int64_t)sizeof(int) * nr, 256);                              /* filtered endpos */
    l.code0 = b;  b += align_up((n_bases > 0 ? n_bases : 0) + 16, 256);                       /* per-base code bytes */
    l.dbg = b;    b += 256;                                                                   /* phase clocks (debug) */
    l.pmax = b;   b += align_up((int64_t)sizeof(int) * nr, 256);                              /* running max of endpos */
    l.tmax = b;   b += align_up((int64_t)sizeof(int) * ((nr + PM_TILE - 1) / PM_TILE + 1), 256);
    l.fblist = b; b += align_up((int64_t)sizeof(int) * ((n_windows > 0 ? n_windows : 0) + 64), 256);
    l.total = b;
    return l;
}

// ---- candidate read range per window on the device (reads.ReadBatch.window_ranges: lo = number of reads whose RUNNING
// MAXIMUM of bam_endpos is <= queryStart, hi = number of reads with pos < queryEnd; a superset of what pysam.fetch returns,
// dep/htslib/hts.c:2242 -- the encoder re-applies the exact rule).  Tile-local inclusive maximum scan + tile maxima ...
__global__ void __launch_bounds__(PM_THREADS) prefmax_tiles_kernel(const ReadAux *__restrict__ aux, const int32_t *__restrict__ pos,
                                                                   int n, int *__restrict__ pmax, int *__restrict__ tmax,
                                                                   int32_t *__restrict__ unsorted) {
    __shared__ int s_w[PM_THREADS / 32];
    const int base = blockIdx.x * PM_TILE + threadIdx.x * PM_PER;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int v[PM_PER];
    int run = INT_MIN;
    bool bad = false;
#pragma unroll
    for (int k = 0; k < PM_PER; k++) {
        const int i = base + k;
        int e = INT_MIN;
        if (i < n) {
            e = aux[i].endpos;
            if (i + 1 < n && pos[i + 1] < pos[i]) bad = true;    // pysam.fetch needs a coordinate-sorted file
        }
        run = max(run, e);
        v[k] = run;
    }
    if (bad) *unsorted = 1;
    int incl = run;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl = max(incl, t);
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    int before = INT_MIN;                                        // maximum of all earlier threads of the tile
    for (int w = 0; w < warp; w++) before = max(before, s_w[w]);
    const int prev = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane > 0) before = max(before, prev);
#pragma unroll
    for (int k = 0; k < PM_PER; k++)
        if (base + k < n) pmax[base + k] = max(v[k], before);
    if (threadIdx.x == PM_THREADS - 1) tmax[blockIdx.x] = max(incl, before);
}

// ... exclusive maximum scan of the tile maxima (one block; tmax[t] becomes the maximum of all tiles before t)
__global__ void __launch_bounds__(1024) prefmax_spine_kernel(int *__restrict__ tmax, int tiles) {
    __shared__ int s_w[32];
    __shared__ int s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = INT_MIN;
    __syncthreads();
    for (int t0 = 0; t0 < tiles; t0 += 1024) {
        const int i = t0 + threadIdx.x;
        const int own = i < tiles ? tmax[i] : INT_MIN;
        int incl = own;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl = max(incl, t);
        }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        int before = s_carry;
        for (int w = 0; w < warp; w++) before = max(before, s_w[w]);
        const int prev = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane > 0) before = max(before, prev);
        if (i < tiles) tmax[i] = before;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = max(before, own);
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) window_ranges_kernel(const int32_t *__restrict__ pos, const int *__restrict__ pmax,
                                                            const int *__restrict__ tmax, int n,
                                                            const int32_t *__restrict__ win_qs, int n_windows, int L,
                                                            int32_t *__restrict__ win_lo, int32_t *__restrict__ win_hi) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_windows) return;
    const int qs = win_qs[w], qe = qs + L;
    int lo = 0, hi = n;                                          // first i with runmax(endpos)[i] > qs
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const int m = max(pmax[mid], tmax[mid / PM_TILE]);
        if (m <= qs) lo = mid + 1; else hi = mid;
    }
    const int first = lo;
    lo = 0; hi = n;                                              // first i with pos[i] >= qe
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (pos[mid] < qe) lo = mid + 1; else hi = mid;
    }
    win_lo[w] = first;
    win_hi[w] = max(lo, first);
}

}  // namespace dp

extern "C" {

const char *dp_last_error(void) { return dp::g_err; }
int dp_version(void) { return 100; }
int dp_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int dp_set_quality_lut(const uint8_t lut[256]) {
    if (!lut) return DP_EINVAL;
    DP_CUDA_CHECK(cudaMemcpyToSymbol(dp::c_qlut, lut, 256));
    int dev = 0;
    DP_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 64) dp::g_lut_ready[dev] = true;
    bool flat = true;
    for (int i = 28; i < 256; i++) flat = flat && lut[i] == lut[28];
    dp::g_lut_flat = flat;
    return DP_OK;
}

int64_t dp_encode_workspace_bytes(int32_t n_reads, int64_t n_cigar, int64_t n_bases, int32_t n_windows,
                                  const dp_encode_params *p) {
    (void)p;
    return dp::ws_layout(n_reads, n_cigar, n_bases, n_windows).total;
}

int dp_encode_fallback_count(const void *workspace, int32_t n_reads, int64_t n_cigar, int32_t *out_count, void *stream) {
    DP_REQUIRE(workspace && out_count && n_reads >= 0 && n_cigar >= 0, "bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    DP_CUDA_CHECK(cudaMemcpyAsync(out_count, workspace, sizeof(int), cudaMemcpyDeviceToHost, st));   /* header word 0 */
    DP_CUDA_CHECK(cudaStreamSynchronize(st));
    return DP_OK;
}

}  // extern "C"

namespace dp {

// per-read digests + per-base code bytes of every read of `reads` (depends on the reads only, not on the windows)
static int encode_prepare(const dp_reads *reads, const dp_encode_params *p, uint8_t *ws, const WsLayout &lay, cudaStream_t st,
                          cudaEvent_t *ev_mid) {
    int rc = ensure_lut();
    if (rc) return rc;
    ReadAux *aux = (ReadAux *)(ws + lay.aux);
    int *RP = (int *)(ws + lay.RP), *QP = (int *)(ws + lay.QP);
    ReadInfo *info = (ReadInfo *)(ws + lay.info);
    int *ckpt = (int *)(ws + lay.ckpt), *endk = (int *)(ws + lay.endk);
    uint8_t *code0 = ws + lay.code0;
    if (reads->n_reads > 0) {
        if (reads->n_cigar / reads->n_reads >= 96) {           /* long CIGARs (long reads): one block per read */
            prep_reads_kernel<8><<<reads->n_reads, 256, 0, st>>>(*reads, aux, RP, QP, info, ckpt, endk, p->map_quality_cutoff);
        } else {
            const int warps_per_block = 8;
            const int blocks = (reads->n_reads + warps_per_block - 1) / warps_per_block;
            prep_reads_kernel<1><<<blocks, 256, 0, st>>>(*reads, aux, RP, QP, info, ckpt, endk, p->map_quality_cutoff);
        }
        DP_CUDA_CHECK(cudaGetLastError());
        if (ev_mid) cudaEventRecord(*ev_mid, st);
        if (reads->n_bases > 0) {
            const int64_t nb = (reads->n_bases + 256 * 16 - 1) / (256 * 16);
            base_codes_kernel<<<(unsigned)(nb < 148 * 64 ? nb : 148 * 64), 256, 0, st>>>(*reads, info, code0);
            DP_CUDA_CHECK(cudaGetLastError());
        }
    } else if (ev_mid) {
        cudaEventRecord(*ev_mid, st);
    }
    return DP_OK;
}

}  // namespace dp

extern "C" {

static int dp_encode_impl(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs, const int32_t *win_lo,
                          const int32_t *win_hi, const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                          const dp_encode_params *p, uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins,
                          int32_t *out_status, void *workspace, int64_t workspace_bytes, int32_t ws_windows,
                          int32_t bases_per_window, int prepared, void *stream) {
    DP_REQUIRE(reads && p && out_img, "null argument");
    DP_REQUIRE(n_windows >= 0 && reads->n_reads >= 0, "negative extent");
    if (n_windows == 0) return DP_OK;
    DP_REQUIRE(win_qs && win_lo && win_hi && ref, "null argument");
    const int L = 2 * p->flank + 1;
    DP_REQUIRE(p->flank >= 0 && L <= dp::ENC_THREADS, "flankLength must satisfy 2*flank+1 <= 256");
    DP_REQUIRE(p->window_cutoff >= 1 && p->window_cutoff <= dp::ENC_THREADS, "windowCutoff must be in [1,256]");
    DP_REQUIRE(p->coverage_cutoff >= 4 && p->coverage_cutoff % 4 == 0 && p->coverage_cutoff <= 1024,
               "coverageCutoff must be a multiple of 4 in [4,1024]");
    cudaStream_t st = (cudaStream_t)stream;
    int rc = dp::ensure_lut();
    if (rc) return rc;

    const int64_t n_cigar = reads->n_cigar;
    DP_REQUIRE(n_cigar >= 0, "negative n_cigar");
    DP_REQUIRE(reads->n_bases >= 0, "negative n_bases");
    DP_REQUIRE(ws_windows >= n_windows, "more windows than the workspace was sized for");
    const dp::WsLayout lay = dp::ws_layout(reads->n_reads, n_cigar, reads->n_bases, ws_windows);
    DP_REQUIRE(workspace && workspace_bytes >= lay.total, "workspace too small (dp_encode_workspace_bytes)");
    uint8_t *wsb = (uint8_t *)workspace;
    dp::ReadAux *aux = (dp::ReadAux *)(wsb + lay.aux);
    int *RP = (int *)(wsb + lay.RP), *QP = (int *)(wsb + lay.QP);
    int *fb = (int *)(wsb + lay.hdr);                          /* fb[0] = count of windows listed for the exact kernel */
    int *fblist = (int *)(wsb + lay.fblist);
    dp::ReadInfo *info = (dp::ReadInfo *)(wsb + lay.info);
    int *ckpt = (int *)(wsb + lay.ckpt), *endk = (int *)(wsb + lay.endk);
    uint8_t *code0 = wsb + lay.code0;
    unsigned long long *dbg = (unsigned long long *)(wsb + lay.dbg);

    /* DP_ENC_TIMING=1: CUDA events around each kernel of this call, printed to stderr (synchronises; debugging aid) */
    static int timing = -1;
    if (timing < 0) { const char *e = getenv("DP_ENC_TIMING"); timing = (e && e[0] == '1') ? 1 : 0; }
    cudaEvent_t ev[5];
    if (timing) { for (int i = 0; i < 5; i++) cudaEventCreate(&ev[i]); cudaEventRecord(ev[0], st); }
    if (!prepared) {
        rc = dp::encode_prepare(reads, p, wsb, lay, st, timing ? &ev[1] : nullptr);
        if (rc) return rc;
    } else if (timing) {
        cudaEventRecord(ev[1], st);
    }
    if (timing) cudaEventRecord(ev[2], st);
    dp::EncArgs a;
    a.R = *reads; a.aux = aux; a.RP = RP; a.QP = QP; a.info = info; a.ckpt = ckpt; a.endk = endk; a.code0 = code0;
    a.win_qs = win_qs; a.win_lo = win_lo; a.win_hi = win_hi;
    a.ref = ref; a.ref_base = ref_base; a.ref_len = ref_len;
    a.L = L; a.W = p->coverage_cutoff; a.H = p->window_cutoff; a.mapq_cut = p->map_quality_cutoff;
    a.out_img = out_img; a.out_rowpos = out_rowpos; a.out_rowins = out_rowins; a.out_status = out_status;
    a.out_colrows = nullptr; a.layout_only = 0;
    const size_t smem = dp::encoder_smem_bytes(p);
    DP_REQUIRE(smem <= 227 * 1024, "coverageCutoff*windowCutoff too large for shared-memory staging");
    DP_CUDA_CHECK(cudaFuncSetAttribute(dp::encode_windows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // fast path (v3) when the geometry fits its shared-memory planes; windows it cannot take are listed for v1
    int force_v1 = 0, force_v3 = 0, force_v4 = 0;             /* DP_ENCODER = v1 | v3 | v4 pins the kernel (read per call: tests switch it) */
    {
        const char *e = getenv("DP_ENCODER");
        force_v1 = (e && !strcmp(e, "v1")) ? 1 : 0;
        force_v3 = (e && !strcmp(e, "v3")) ? 1 : 0;
        force_v4 = (e && !strcmp(e, "v4")) ? 1 : 0;
    }
    // shallow windows (30x long reads: ~3 k read bases per window at stride 100) run 5 CTAs per SM on a smaller pool, deep ones
    // 3 CTAs on a larger one; thresholds from a coverage sweep (45 / 60 / 75 / 90x: the 4-CTA build wins at 45-75x, the 3-CTA one from 90x)
    static int force_pool = -1;                                /* DP_ENC_POOL = small | large | deep overrides the choice */
    if (force_pool < 0) { const char *e = getenv("DP_ENC_POOL"); force_pool = !e ? 0 : !strcmp(e, "small") ? 1 : !strcmp(e, "large") ? 2 : !strcmp(e, "deep") ? 3 : 0; }
    // depth of the pileup: read bases over an average reference position (the caller's figure in prepared mode, where the read
    // arrays may cover more than the reference slice), and reads overlapping an average window
    const double cover = bases_per_window > 0 ? (double)bases_per_window / L : (double)reads->n_bases / (double)(ref_len > 0 ? ref_len : 1);
    const double reads_per_window = reads->n_bases > 0 ? cover * (1.0 + (double)L * reads->n_reads / (double)reads->n_bases) : 0.0;
    const bool small_pool = force_pool == 1 || (force_pool == 0 && cover < 40.0);
    const bool deep_pool = force_pool == 3 || (force_pool == 0 && cover >= 87.0);
    const size_t smem3 = dp::encoder_v3_smem_bytes(p, small_pool ? dp::V3_POOL_SMALL : deep_pool ? dp::V3_POOL_DEEP : dp::V3_POOL_BYTES);
    const bool use_v3 = !force_v1 && dp::g_lut_flat && L <= dp::V3_CS && p->window_cutoff <= 256 &&
                        p->coverage_cutoff >= 8 && p->coverage_cutoff <= 256 && smem3 <= 100 * 1024;
    if (!use_v3) {
        DP_CUDA_CHECK(cudaMemsetAsync(fb, 0xFF, sizeof(int), st));   /* fallback count -1: fast path not used */
        dp::encode_windows_kernel<<<n_windows, dp::ENC_THREADS, smem, st>>>(a, nullptr, nullptr);
        DP_CUDA_CHECK(cudaGetLastError());
        return DP_OK;
    }
    DP_CUDA_CHECK(cudaMemsetAsync(fb, 0, sizeof(int), st));
    // v4 (a warp per window, lane = read) takes the call when 32 * NR lanes hold the reads of nearly every window; the few
    // windows beyond that, like windows with a non-clean read, go to the exact kernel's list
    int v4_nr = reads_per_window <= 18.0 ? 1 : reads_per_window <= 46.0 ? 2 : reads_per_window <= 100.0 ? 4 : 0;
    if (force_v4 && v4_nr == 0) v4_nr = 4;
    // v4 is OPT-IN (DP_ENCODER=v4): bit-exact like v3, but measured slower on B200 (0.53 vs 0.23 ms per 4096 windows at 30x):
    // one warp walks ~70 columns as one dependent chain of ~1 000 instructions each (64-bit packed counters, three entry slots
    // per lane and round), at 16 warps per SM -- see DESIGN.md section 3
    if (force_v4 && !force_v3 && v4_nr > 0 && p->coverage_cutoff <= 256 && L + 1 <= 224) {
        const unsigned grid = (unsigned)((n_windows + dp::V4_WARPS - 1) / dp::V4_WARPS);
        if (v4_nr == 1) dp::encode_windows_v4_kernel<1><<<grid, dp::V4_WARPS * 32, 0, st>>>(a, n_windows, fb, fblist);
        else if (v4_nr == 2) dp::encode_windows_v4_kernel<2><<<grid, dp::V4_WARPS * 32, 0, st>>>(a, n_windows, fb, fblist);
        else dp::encode_windows_v4_kernel<4><<<grid, dp::V4_WARPS * 32, 0, st>>>(a, n_windows, fb, fblist);
        DP_CUDA_CHECK(cudaGetLastError());
        if (timing) cudaEventRecord(ev[3], st);
        dp::encode_windows_kernel<<<n_windows, dp::ENC_THREADS, smem, st>>>(a, fb, fblist);
        DP_CUDA_CHECK(cudaGetLastError());
        if (timing) {
            cudaEventRecord(ev[4], st);
            cudaEventSynchronize(ev[4]);
            float t[4] = {0, 0, 0, 0};
            cudaEventElapsedTime(&t[0], ev[0], ev[1]); cudaEventElapsedTime(&t[1], ev[1], ev[2]);
            cudaEventElapsedTime(&t[2], ev[2], ev[3]); cudaEventElapsedTime(&t[3], ev[3], ev[4]);
            int nfb = 0;
            cudaMemcpy(&nfb, fb, sizeof(int), cudaMemcpyDeviceToHost);
            fprintf(stderr, "[dp_encode_windows] %d windows (v4, %d rounds): prep %.3f ms, base codes %.3f ms, fast kernel %.3f ms, "
                            "exact kernel %.3f ms (%d windows)\n", n_windows, v4_nr, t[0], t[1], t[2], t[3], nfb);
            for (int i = 0; i < 5; i++) cudaEventDestroy(ev[i]);
        }
        return DP_OK;
    }
    DP_CUDA_CHECK(cudaFuncSetAttribute(dp::encode_windows_v3_kernel<4, dp::V3_POOL_BYTES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)dp::encoder_v3_smem_bytes(p, dp::V3_POOL_BYTES)));
    DP_CUDA_CHECK(cudaFuncSetAttribute(dp::encode_windows_v3_kernel<5, dp::V3_POOL_SMALL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)dp::encoder_v3_smem_bytes(p, dp::V3_POOL_SMALL)));
    DP_CUDA_CHECK(cudaFuncSetAttribute(dp::encode_windows_v3_kernel<3, dp::V3_POOL_DEEP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)dp::encoder_v3_smem_bytes(p, dp::V3_POOL_DEEP)));
    static int emit_mode = -1;                                 /* DP_ENC_EMIT=stg: plain stores instead of bulk copies */
    if (emit_mode < 0) { const char *e = getenv("DP_ENC_EMIT"); emit_mode = (e && !strcmp(e, "stg")) ? 1 : 0; }
    if (timing) cudaMemsetAsync(dbg, 0, 256, st);
    if (small_pool)
        dp::encode_windows_v3_kernel<5, dp::V3_POOL_SMALL><<<n_windows, dp::V3_THREADS, smem3, st>>>(a, fb, fblist, emit_mode, timing ? dbg : nullptr);
    else if (deep_pool)
        dp::encode_windows_v3_kernel<3, dp::V3_POOL_DEEP><<<n_windows, dp::V3_THREADS, smem3, st>>>(a, fb, fblist, emit_mode, timing ? dbg : nullptr);
    else
        dp::encode_windows_v3_kernel<4, dp::V3_POOL_BYTES><<<n_windows, dp::V3_THREADS, smem3, st>>>(a, fb, fblist, emit_mode, timing ? dbg : nullptr);
    DP_CUDA_CHECK(cudaGetLastError());
    if (timing) cudaEventRecord(ev[3], st);
    dp::encode_windows_kernel<<<n_windows, dp::ENC_THREADS, smem, st>>>(a, fb, fblist);
    DP_CUDA_CHECK(cudaGetLastError());
    if (timing) {
        cudaEventRecord(ev[4], st);
        cudaEventSynchronize(ev[4]);
        float t[4] = {0, 0, 0, 0};
        if (reads->n_reads > 0 && reads->n_bases > 0) { cudaEventElapsedTime(&t[0], ev[0], ev[1]); cudaEventElapsedTime(&t[1], ev[1], ev[2]); }
        cudaEventElapsedTime(&t[2], ev[2], ev[3]);
        cudaEventElapsedTime(&t[3], ev[3], ev[4]);
        fprintf(stderr, "[dp_encode_windows] %d windows: prep %.3f ms, base codes %.3f ms, fast kernel %.3f ms, exact kernel %.3f ms\n",
                n_windows, t[0], t[1], t[2], t[3]);
        for (int i = 0; i < 5; i++) cudaEventDestroy(ev[i]);
        unsigned long long hd[16];
        cudaMemcpy(hd, dbg, sizeof(hd), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[dp_encode_windows] fast kernel cycles/window by phase (setup+compact, info, scan, rowmap, gen, bases, emit1, place, finish, drain):");
        for (int i = 0; i < 10; i++) fprintf(stderr, " %llu", hd[i] / (unsigned long long)n_windows);
        fprintf(stderr, "\n");
    }
    return DP_OK;
}

int dp_encode_windows(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs, const int32_t *win_lo,
                      const int32_t *win_hi, const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                      const dp_encode_params *p, uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins,
                      int32_t *out_status, void *workspace, int64_t workspace_bytes, void *stream) {
    return dp_encode_impl(reads, n_windows, win_qs, win_lo, win_hi, ref, ref_base, ref_len, p, out_img, out_rowpos, out_rowins,
                          out_status, workspace, workspace_bytes, n_windows, 0, 0, stream);
}

int dp_encode_prepare(const dp_reads *reads, const dp_encode_params *p, int32_t max_windows, void *workspace,
                      int64_t workspace_bytes, void *stream) {
    DP_REQUIRE(reads && p && reads->n_reads >= 0 && reads->n_cigar >= 0 && reads->n_bases >= 0 && max_windows >= 0, "bad argument");
    const dp::WsLayout lay = dp::ws_layout(reads->n_reads, reads->n_cigar, reads->n_bases, max_windows);
    DP_REQUIRE(workspace && workspace_bytes >= lay.total, "workspace too small (dp_encode_workspace_bytes)");
    return dp::encode_prepare(reads, p, (uint8_t *)workspace, lay, (cudaStream_t)stream, nullptr);
}

int dp_encode_prepared(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs, const int32_t *win_lo,
                       const int32_t *win_hi, const uint8_t *ref, int64_t ref_base, int64_t ref_len,
                       const dp_encode_params *p, uint8_t *out_img, int32_t *out_rowpos, int32_t *out_rowins,
                       int32_t *out_status, void *workspace, int64_t workspace_bytes, int32_t max_windows,
                       int32_t bases_per_window, void *stream) {
    return dp_encode_impl(reads, n_windows, win_qs, win_lo, win_hi, ref, ref_base, ref_len, p, out_img, out_rowpos, out_rowins,
                          out_status, workspace, workspace_bytes, max_windows, bases_per_window, 1, stream);
}

int dp_window_layout(const dp_reads *reads, int32_t n_windows, const int32_t *win_qs, const int32_t *win_lo,
                     const int32_t *win_hi, const uint8_t *ref, int64_t ref_base, int64_t ref_len, const dp_encode_params *p,
                     int32_t *out_colrows, int32_t *out_status, void *workspace, int64_t workspace_bytes, void *stream) {
    DP_REQUIRE(reads && p && out_colrows, "null argument");
    DP_REQUIRE(n_windows >= 0 && reads->n_reads >= 0 && reads->n_cigar >= 0 && reads->n_bases >= 0, "negative extent");
    if (n_windows == 0) return DP_OK;
    DP_REQUIRE(win_qs && win_lo && win_hi && ref, "null argument");
    const int L = 2 * p->flank + 1;
    DP_REQUIRE(p->flank >= 0 && L <= dp::ENC_THREADS, "flankLength must satisfy 2*flank+1 <= 256");
    cudaStream_t st = (cudaStream_t)stream;
    const dp::WsLayout lay = dp::ws_layout(reads->n_reads, reads->n_cigar, reads->n_bases, n_windows);
    DP_REQUIRE(workspace && workspace_bytes >= lay.total, "workspace too small (dp_encode_workspace_bytes)");
    uint8_t *wsb = (uint8_t *)workspace;
    int rc = dp::encode_prepare(reads, p, wsb, lay, st, nullptr);
    if (rc) return rc;
    dp::EncArgs a;
    memset(&a, 0, sizeof(a));
    a.R = *reads;
    a.aux = (dp::ReadAux *)(wsb + lay.aux); a.RP = (int *)(wsb + lay.RP); a.QP = (int *)(wsb + lay.QP);
    a.info = (dp::ReadInfo *)(wsb + lay.info); a.ckpt = (int *)(wsb + lay.ckpt); a.endk = (int *)(wsb + lay.endk);
    a.code0 = wsb + lay.code0;
    a.win_qs = win_qs; a.win_lo = win_lo; a.win_hi = win_hi;
    a.ref = ref; a.ref_base = ref_base; a.ref_len = ref_len;
    a.L = L; a.W = p->coverage_cutoff; a.H = p->window_cutoff; a.mapq_cut = p->map_quality_cutoff;
    a.out_status = out_status; a.out_colrows = out_colrows; a.layout_only = 1;
    const size_t smem = dp::encoder_smem_bytes(p);
    DP_REQUIRE(smem <= 227 * 1024, "coverageCutoff*windowCutoff too large for shared-memory staging");
    DP_CUDA_CHECK(cudaFuncSetAttribute(dp::encode_windows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dp::encode_windows_kernel<<<n_windows, dp::ENC_THREADS, smem, st>>>(a, nullptr, nullptr);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

int dp_window_ranges(const dp_reads *reads, const void *workspace, int64_t workspace_bytes, int32_t max_windows,
                     int32_t n_windows, const int32_t *win_qs, int32_t flank, int32_t *win_lo, int32_t *win_hi,
                     int32_t *unsorted, void *stream) {
    DP_REQUIRE(reads && reads->n_reads >= 0 && n_windows >= 0 && flank >= 0 && max_windows >= 0, "bad argument");
    if (n_windows == 0) return DP_OK;
    DP_REQUIRE(win_qs && win_lo && win_hi && unsorted, "null argument");
    const dp::WsLayout lay = dp::ws_layout(reads->n_reads, reads->n_cigar, reads->n_bases, max_windows);
    DP_REQUIRE(workspace && workspace_bytes >= lay.total, "workspace too small (dp_encode_workspace_bytes)");
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t *wsb = (uint8_t *)workspace;
    const dp::ReadAux *aux = (const dp::ReadAux *)(wsb + lay.aux);
    int *pmax = (int *)(wsb + lay.pmax), *tmax = (int *)(wsb + lay.tmax);
    const int n = reads->n_reads;
    DP_CUDA_CHECK(cudaMemsetAsync(unsorted, 0, sizeof(int32_t), st));
    if (n > 0) {
        const int tiles = (n + dp::PM_TILE - 1) / dp::PM_TILE;
        dp::prefmax_tiles_kernel<<<tiles, dp::PM_THREADS, 0, st>>>(aux, reads->pos, n, pmax, tmax, unsorted);
        dp::prefmax_spine_kernel<<<1, 1024, 0, st>>>(tmax, tiles);
        DP_CUDA_CHECK(cudaGetLastError());
    }
    dp::window_ranges_kernel<<<(n_windows + 255) / 256, 256, 0, st>>>(reads->pos, pmax, tmax, n, win_qs, n_windows, 2 * flank + 1,
                                                                       win_lo, win_hi);
    DP_CUDA_CHECK(cudaGetLastError());
    return DP_OK;
}

}  // extern "C"
