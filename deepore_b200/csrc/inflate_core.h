/* DEFLATE (RFC 1951) decoder core of the device-side BGZF stage.
 *
 * Replaces, per BGZF block, what the reference's host stage does through zlib: dep/htslib/bgzf.c:456-499
 * (inflate_block: inflateInit2(-15) / inflate(Z_FINISH) on the block payload) -- zlib itself is a third-party
 * dependency that is not under /root/reference; the algorithm restated here is the published RFC 1951 one
 * (stored / fixed / dynamic blocks, canonical Huffman codes decoded by a 9-bit lookup table with the count/first-code walk behind it).
 *
 * One decoder instance = one warp (device) or one thread (host) decoding one BGZF payload front to back; the caller
 * provides the table scratch (shared memory on the device).  Everything is bounds-checked: malformed input ends with a negative
 * code, never with a read or write outside [in, in+in_len) / [out, out+out_cap).
 *
 * Plain C so that the same source compiles for the device (nvcc) and, for the CPU test suite only, for the host
 * (oracle/inflate_host.c).  The product path only ever runs it inside bgzf_inflate_kernel. */
#ifndef DP_INFLATE_CORE_H
#define DP_INFLATE_CORE_H
#include <stdint.h>

#ifdef __CUDACC__
#define DPI_FN __device__ __forceinline__
#define DPI_CONST __constant__
/* On the device the 32 lanes of a warp run ONE decoder in lock step: every lane holds the same bit reservoir and takes the
 * same branches (same-address loads broadcast, no divergence), and only the output is split -- lane 0 stores literals,
 * the lanes stride over the bytes of a match or of a stored block.  DPI_SYNC orders the lanes' earlier stores before a
 * match reads them back. */
#define DPI_LANE  ((uint32_t)(threadIdx.x & 31))
#define DPI_LANES 32u
#define DPI_SYNC() __syncwarp()
#else
#define DPI_FN static inline
#define DPI_CONST static const
#define DPI_LANE  0u
#define DPI_LANES 1u
#define DPI_SYNC() ((void)0)
#endif

#define DPI_OK            0
#define DPI_E_INPUT      -1   /* ran out of input                                  */
#define DPI_E_OUTPUT     -2   /* output would exceed out_cap                       */
#define DPI_E_BTYPE      -3   /* reserved block type 3                             */
#define DPI_E_STORED     -4   /* stored block: LEN != ~NLEN                        */
#define DPI_E_TABLE      -5   /* over-subscribed / incomplete code, bad repeat     */
#define DPI_E_SYMBOL     -6   /* no code matches, or symbol out of range           */
#define DPI_E_DISTANCE   -7   /* distance reaches before the start of the output   */
#define DPI_E_LENGTH     -8   /* inflated size differs from ISIZE                  */

#define DPI_MAXBITS 15
#define DPI_NLEN    288
#define DPI_NDIST   30
/* first-level lookup of a code: DPI_FAST_BITS peeked bits -> 32-bit entry
 *     bits 0-3 code length | bits 4-7 extra bits | bits 8-23 base value | bits 24-25 kind
 * with everything the symbol means already resolved (a literal's byte, a length's or distance's base and extra-bit count:
 * no second table, no symbol-range compares on the hot path); 0 = the code is longer than DPI_FAST_BITS, or the symbol is
 * not a legal one (decoded by the count/first-code walk instead, which also reports the error) */
#define DPI_FAST_BITS 9
#define DPI_FAST_SIZE (1 << DPI_FAST_BITS)
#define DPI_K_LITERAL 1u
#define DPI_K_BASE    2u      /* length or distance: base + extra bits */
#define DPI_K_END     3u
/* scratch per decoder, in uint16 units: counts + symbols of both codes, the code-length list, the two lookup tables
 * (literal/length and distance) */
#define DPI_SLOW_U16 (16 + DPI_NLEN + 16 + DPI_NDIST + (DPI_NLEN + DPI_NDIST + 2) / 2 + 1)
#define DPI_SCRATCH_U16 (((DPI_SLOW_U16 + 1) & ~1) + 4 * DPI_FAST_SIZE)      /* the two tables hold 32-bit entries, 4-byte aligned */

DPI_CONST uint16_t dpi_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83,
                                       99, 115, 131, 163, 195, 227, 258};
DPI_CONST uint8_t dpi_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
DPI_CONST uint16_t dpi_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769,
                                        1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
DPI_CONST uint8_t dpi_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11,
                                        12, 12, 13, 13};
DPI_CONST uint8_t dpi_clen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

typedef struct dpi_stream {
    const uint8_t *in;
    uint32_t in_len, in_pos;
    uint64_t bitbuf;          /* LSB-first bit reservoir */
    int32_t bitcnt;
    uint8_t *out;
    uint32_t out_cap, out_pos;
    int32_t err;
} dpi_stream;

/* Top the reservoir up: one aligned 32-bit load where the input pointer allows it (every refill after the first few bytes
 * of a stream; a byte load per 8 bits was a dependent load per output byte), single bytes at unaligned positions and at
 * the tail.  Never reads outside [in, in + in_len). */
DPI_FN void dpi_refill(dpi_stream *s)
{
    if (s->bitcnt > 32) return;
    if (s->in_pos + 4 <= s->in_len && ((uintptr_t)(s->in + s->in_pos) & 3u) == 0) {
        const uint32_t w = *(const uint32_t *)(s->in + s->in_pos);           /* little-endian hosts and devices */
        s->bitbuf |= (uint64_t)w << s->bitcnt;
        s->bitcnt += 32;
        s->in_pos += 4;
        return;
    }
    while (s->bitcnt <= 56 && s->in_pos < s->in_len && (s->bitcnt <= 24 || ((uintptr_t)(s->in + s->in_pos) & 3u) != 0)) {
        s->bitbuf |= (uint64_t)s->in[s->in_pos++] << s->bitcnt;
        s->bitcnt += 8;
    }
}

/* Take `need` (<= 16) bits; past the end of the input the stream goes into error and zeros are returned. */
DPI_FN uint32_t dpi_bits(dpi_stream *s, int need)
{
    if (s->bitcnt < need) {
        dpi_refill(s);
        if (s->bitcnt < need) {
            s->err = DPI_E_INPUT;
            return 0;
        }
    }
    uint32_t v = (uint32_t)(s->bitbuf & ((1u << need) - 1u));
    s->bitbuf >>= need;
    s->bitcnt -= need;
    return v;
}

/* Canonical code described by count[len] (codes of each length) and symbol[] (symbols ordered by code). */
DPI_FN int dpi_decode(dpi_stream *s, const uint16_t *count, const uint16_t *symbol)
{
    int32_t code = 0, first = 0, index = 0;
    for (int len = 1; len <= DPI_MAXBITS; ++len) {
        code |= (int32_t)dpi_bits(s, 1);
        if (s->err) return s->err;
        int32_t c = count[len];
        if (code - c < first) return symbol[index + (code - first)];
        index += c;
        first = (first + c) << 1;
        code <<= 1;
    }
    return DPI_E_SYMBOL;
}

/* Build count[]/symbol[] from code lengths.  Returns 0 for a complete code, > 0 for an incomplete one (allowed only
 * where RFC 1951 allows a single code), < 0 for an over-subscribed one. */
DPI_FN int dpi_build(uint16_t *count, uint16_t *symbol, const uint8_t *length, int n)
{
    /* the tables are shared by the lanes of a warp: lane 0 writes them (read-modify-write on shared memory must not be
     * repeated by 32 lanes), every lane reads them back behind a DPI_SYNC */
    DPI_SYNC();                                                  /* nobody still decodes with the previous tables */
    if (DPI_LANE == 0) {
        for (int len = 0; len <= DPI_MAXBITS; ++len) count[len] = 0;
        for (int i = 0; i < n; ++i) count[length[i]]++;
    }
    DPI_SYNC();
    if (count[0] == n) return 0;
    int left = 1;
    for (int len = 1; len <= DPI_MAXBITS; ++len) {
        left = (left << 1) - count[len];
        if (left < 0) return left;
    }
    if (DPI_LANE == 0) {
        uint16_t offs[DPI_MAXBITS + 1];
        offs[1] = 0;
        for (int len = 1; len < DPI_MAXBITS; ++len) offs[len + 1] = offs[len] + count[len];
        for (int i = 0; i < n; ++i)
            if (length[i]) symbol[offs[length[i]]++] = (uint16_t)i;
    }
    DPI_SYNC();
    return left;
}

/* Lookup table of the literal/length code from its code lengths (RFC 1951 3.2.2: codes of one length are consecutive,
 * shorter codes first; the stream carries them most-significant bit first, so the peeked index is the reversed code). */
DPI_FN uint32_t dpi_entry(int sym, int len, int is_dist)
{
    if (is_dist) {
        if (sym >= DPI_NDIST) return 0;
        return (uint32_t)len | ((uint32_t)dpi_dist_extra[sym] << 4) | ((uint32_t)dpi_dist_base[sym] << 8) | (DPI_K_BASE << 24);
    }
    if (sym < 256) return (uint32_t)len | ((uint32_t)sym << 8) | (DPI_K_LITERAL << 24);
    if (sym == 256) return (uint32_t)len | (DPI_K_END << 24);
    if (sym - 257 >= 29) return 0;
    return (uint32_t)len | ((uint32_t)dpi_len_extra[sym - 257] << 4) | ((uint32_t)dpi_len_base[sym - 257] << 8) | (DPI_K_BASE << 24);
}

DPI_FN void dpi_build_fast(uint32_t *fast, const uint16_t *count, const uint8_t *length, int n, int is_dist)
{
    DPI_SYNC();
    for (uint32_t i = DPI_LANE; i < DPI_FAST_SIZE; i += DPI_LANES) fast[i] = 0;
    DPI_SYNC();
    if (DPI_LANE == 0) {
        uint16_t next[DPI_MAXBITS + 1];
        uint32_t code = 0;
        next[0] = 0;
        for (int len = 1; len <= DPI_MAXBITS; ++len) {
            code = (code + (len > 1 ? count[len - 1] : 0)) << 1;
            next[len] = (uint16_t)code;
        }
        for (int sym = 0; sym < n; ++sym) {
            const int len = length[sym];
            if (len == 0) continue;
            const uint32_t c = next[len]++;
            if (len > DPI_FAST_BITS) continue;
            uint32_t rev = 0;
            for (int b = 0; b < len; ++b) rev |= ((c >> b) & 1u) << (len - 1 - b);
            const uint32_t e = dpi_entry(sym, len, is_dist);
            for (uint32_t idx = rev; idx < DPI_FAST_SIZE; idx += 1u << len) fast[idx] = e;
        }
    }
    DPI_SYNC();
}

/* One symbol as a table entry (format above): a lookup when the code is short and enough input is buffered, else the walk. */
DPI_FN uint32_t dpi_symbol(dpi_stream *s, const uint32_t *fast, const uint16_t *count, const uint16_t *symbol, int is_dist)
{
    const uint32_t e = fast[(uint32_t)s->bitbuf & (DPI_FAST_SIZE - 1)];
    const int len = (int)(e & 15u);
    if (e && len <= s->bitcnt) {
        s->bitbuf >>= len;
        s->bitcnt -= len;
        return e;
    }
    const int sym = dpi_decode(s, count, symbol);
    if (sym < 0) { s->err = sym; return 0; }
    const uint32_t w = dpi_entry(sym, 1, is_dist);
    if (!w) s->err = DPI_E_SYMBOL;
    return w;
}

/* Output goes through a batch of up to DPI_LANES PENDING bytes kept in registers -- lane i holds the i-th pending byte --
 * and leaves with ONE store instruction per batch (consecutive bytes, a 32-byte run per warp instruction instead of
 * single-byte stores by lane 0).  Literals append one byte; a match of distance >= DPI_LANES appends its bytes too: the
 * lanes that receive them load them from the already written output (consecutive addresses; such a source never lies in
 * the pending batch).  Only matches closer than DPI_LANES -- whose source may be pending or may be the match itself -- flush
 * the batch and copy directly.  `dirty` remembers that this warp stored since the lanes last synchronised, so that a
 * source byte written by another lane is ordered before its load. */
#define DPI_FLUSH_LITERALS()                                                        \
    do {                                                                            \
        if (DPI_LANE < run) s->out[s->out_pos - run + DPI_LANE] = mylit;            \
        if (run) dirty = 1;                                                         \
        run = 0;                                                                    \
        cap = DPI_LANES;                                                            \
    } while (0)
/* The first batch of a stream is cut so that it ENDS on a DPI_LANES-byte boundary: every later batch then covers one aligned
 * 32-byte sector and the L2 never has to fill a partly written sector from DRAM. */
#define DPI_FIRST_CAP(p) (DPI_LANES - (uint32_t)((uintptr_t)(p) & (DPI_LANES - 1u)))

DPI_FN int dpi_codes(dpi_stream *s, const uint32_t *fast, const uint32_t *dfast, const uint16_t *lcnt, const uint16_t *lsym,
                     const uint16_t *dcnt, const uint16_t *dsym)
{
    uint32_t run = 0;                                            /* bytes decoded but not yet stored */
    uint32_t cap = DPI_FIRST_CAP(s->out + s->out_pos);           /* bytes the current batch takes */
    uint32_t dirty = 0;
    uint8_t mylit = 0;
    for (;;) {
        if (s->bitcnt < 20) dpi_refill(s);                       /* a literal/length code (<= 15 bits) + its extra bits (<= 5) */
        const uint32_t e = dpi_symbol(s, fast, lcnt, lsym, 0);
        const uint32_t kind = e >> 24;
        if (kind == DPI_K_LITERAL) {
            if (s->out_pos >= s->out_cap) { DPI_FLUSH_LITERALS(); return DPI_E_OUTPUT; }
            if (DPI_LANE == run) mylit = (uint8_t)(e >> 8);
            run++;
            s->out_pos++;
            if (run == cap) DPI_FLUSH_LITERALS();
        } else if (kind == DPI_K_BASE) {
            uint32_t len = ((e >> 8) & 0xffffu) + dpi_bits(s, (int)((e >> 4) & 15u));
            if (s->bitcnt < 28) dpi_refill(s);                   /* a distance code (<= 15 bits) + its extra bits (<= 13) */
            const uint32_t de = dpi_symbol(s, dfast, dcnt, dsym, 1);
            if (!de) { DPI_FLUSH_LITERALS(); return s->err ? s->err : DPI_E_SYMBOL; }
            const uint32_t dist = ((de >> 8) & 0xffffu) + dpi_bits(s, (int)((de >> 4) & 15u));
            if (s->err) { DPI_FLUSH_LITERALS(); return s->err; }
            if (dist > s->out_pos) { DPI_FLUSH_LITERALS(); return DPI_E_DISTANCE; }
            if (s->out_pos + len > s->out_cap) { DPI_FLUSH_LITERALS(); return DPI_E_OUTPUT; }
            if (dist >= DPI_LANES) {
                /* every source byte is already in memory: out_pos + k - dist < out_pos - run for k < DPI_LANES - run */
                if (dirty) { DPI_SYNC(); dirty = 0; }
                while (len) {
                    const uint32_t n = len < cap - run ? len : cap - run;
                    if (DPI_LANE >= run && DPI_LANE < run + n) mylit = s->out[s->out_pos + (DPI_LANE - run) - dist];
                    run += n;
                    s->out_pos += n;
                    len -= n;
                    if (run == cap) {
                        DPI_FLUSH_LITERALS();
                        if (len && dist < 2 * DPI_LANES) { DPI_SYNC(); dirty = 0; }   /* the next piece may read this batch */
                    }
                }
            } else {
                DPI_FLUSH_LITERALS();
                uint8_t *d = s->out + s->out_pos;
                const uint8_t *f = d - dist;
                DPI_SYNC();
                if (dist >= len) {
                    for (uint32_t i = DPI_LANE; i < len; i += DPI_LANES) d[i] = f[i];
                } else {                                         /* the match overlaps itself: its first `dist` bytes repeat */
                    for (uint32_t i = DPI_LANE; i < len; i += DPI_LANES) d[i] = f[i % dist];
                }
                dirty = 1;
                s->out_pos += len;
                cap = DPI_FIRST_CAP(s->out + s->out_pos);        /* re-align the batches behind the direct copy */
            }
        } else if (kind == DPI_K_END) {
            DPI_FLUSH_LITERALS();
            return DPI_OK;
        } else {                                                 /* no code matches / illegal symbol / input exhausted */
            DPI_FLUSH_LITERALS();
            return s->err ? s->err : DPI_E_SYMBOL;
        }
    }
}

/* Inflate one raw DEFLATE stream.  scratch: DPI_SCRATCH_U16 uint16.  Returns DPI_OK or a DPI_E_* code;
 * *out_len = bytes written. */
DPI_FN int dpi_inflate(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t out_cap, uint16_t *scratch,
                       uint32_t *out_len)
{
    dpi_stream st;
    st.in = in; st.in_len = in_len; st.in_pos = 0; st.bitbuf = 0; st.bitcnt = 0;
    st.out = out; st.out_cap = out_cap; st.out_pos = 0; st.err = 0;
    uint16_t *lcnt = scratch, *lsym = lcnt + 16, *dcnt = lsym + DPI_NLEN, *dsym = dcnt + 16;
    uint8_t *lengths = (uint8_t *)(dsym + DPI_NDIST);
    uint32_t *fast = (uint32_t *)(scratch + ((DPI_SLOW_U16 + 1) & ~1)), *dfast = fast + DPI_FAST_SIZE;
    int last, rc = DPI_OK;
    do {
        last = (int)dpi_bits(&st, 1);
        int type = (int)dpi_bits(&st, 2);
        if (st.err) { rc = st.err; break; }
        if (type == 0) {
            st.in_pos -= (uint32_t)st.bitcnt >> 3;               /* whole bytes the table lookup read ahead go back */
            st.bitbuf = 0; st.bitcnt = 0;                         /* skip to the byte boundary */
            if (st.in_pos + 4 > st.in_len) { rc = DPI_E_INPUT; break; }
            uint32_t len = st.in[st.in_pos] | ((uint32_t)st.in[st.in_pos + 1] << 8);
            uint32_t nlen = st.in[st.in_pos + 2] | ((uint32_t)st.in[st.in_pos + 3] << 8);
            st.in_pos += 4;
            if ((len ^ 0xffffu) != nlen) { rc = DPI_E_STORED; break; }
            if (st.in_pos + len > st.in_len) { rc = DPI_E_INPUT; break; }
            if (st.out_pos + len > st.out_cap) { rc = DPI_E_OUTPUT; break; }
            for (uint32_t i = DPI_LANE; i < len; i += DPI_LANES) st.out[st.out_pos + i] = st.in[st.in_pos + i];
            st.in_pos += len; st.out_pos += len;
        } else if (type == 1) {
            int i = 0;
            for (; i < 144; ++i) lengths[i] = 8;
            for (; i < 256; ++i) lengths[i] = 9;
            for (; i < 280; ++i) lengths[i] = 7;
            for (; i < DPI_NLEN; ++i) lengths[i] = 8;
            dpi_build(lcnt, lsym, lengths, DPI_NLEN);
            dpi_build_fast(fast, lcnt, lengths, DPI_NLEN, 0);
            for (i = 0; i < DPI_NDIST; ++i) lengths[i] = 5;
            dpi_build(dcnt, dsym, lengths, DPI_NDIST);
            dpi_build_fast(dfast, dcnt, lengths, DPI_NDIST, 1);
            rc = dpi_codes(&st, fast, dfast, lcnt, lsym, dcnt, dsym);
        } else if (type == 2) {
            int nlen = (int)dpi_bits(&st, 5) + 257, ndist = (int)dpi_bits(&st, 5) + 1, ncode = (int)dpi_bits(&st, 4) + 4;
            if (st.err) { rc = st.err; break; }
            if (nlen > 286 || ndist > DPI_NDIST) { rc = DPI_E_TABLE; break; }
            int i;
            for (i = 0; i < 19; ++i) lengths[i] = 0;
            for (i = 0; i < ncode; ++i) lengths[dpi_clen_order[i]] = (uint8_t)dpi_bits(&st, 3);
            if (st.err) { rc = st.err; break; }
            if (dpi_build(lcnt, lsym, lengths, 19) != 0) { rc = DPI_E_TABLE; break; }   /* code-length code must be complete */
            i = 0;
            while (i < nlen + ndist) {
                int sym = dpi_decode(&st, lcnt, lsym);
                if (sym < 0) { rc = sym; break; }
                if (sym < 16) {
                    lengths[i++] = (uint8_t)sym;
                } else {
                    int prev = 0, rep;
                    if (sym == 16) {
                        if (i == 0) { rc = DPI_E_TABLE; break; }
                        prev = lengths[i - 1];
                        rep = 3 + (int)dpi_bits(&st, 2);
                    } else if (sym == 17) {
                        rep = 3 + (int)dpi_bits(&st, 3);
                    } else {
                        rep = 11 + (int)dpi_bits(&st, 7);
                    }
                    if (st.err) { rc = st.err; break; }
                    if (i + rep > nlen + ndist) { rc = DPI_E_TABLE; break; }
                    while (rep--) lengths[i++] = (uint8_t)prev;
                }
            }
            if (rc != DPI_OK) break;
            if (lengths[256] == 0) { rc = DPI_E_TABLE; break; }                         /* no end-of-block code */
            /* the distance lengths must be read before lsym is overwritten: build the distance code first, from a copy
             * position that dpi_build of the literal code does not touch (lengths[] is separate from both tables) */
            int e = dpi_build(dcnt, dsym, lengths + nlen, ndist);
            if (e < 0 || (e > 0 && ndist - dcnt[0] != 1)) { rc = DPI_E_TABLE; break; }
            dpi_build_fast(dfast, dcnt, lengths + nlen, ndist, 1);
            e = dpi_build(lcnt, lsym, lengths, nlen);
            if (e < 0 || (e > 0 && nlen - lcnt[0] != 1)) { rc = DPI_E_TABLE; break; }
            dpi_build_fast(fast, lcnt, lengths, nlen, 0);
            rc = dpi_codes(&st, fast, dfast, lcnt, lsym, dcnt, dsym);
        } else {
            rc = DPI_E_BTYPE;
        }
    } while (rc == DPI_OK && !last);
    *out_len = st.out_pos;
    return rc;
}

#endif /* DP_INFLATE_CORE_H */
