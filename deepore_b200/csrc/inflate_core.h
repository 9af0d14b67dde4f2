/* DEFLATE (RFC 1951) decoder core of the device-side BGZF stage.
 *
 * Replaces, per BGZF block, what the reference's host stage does through zlib: dep/htslib/bgzf.c:456-499
 * (inflate_block: inflateInit2(-15) / inflate(Z_FINISH) on the block payload) -- zlib itself is a third-party
 * dependency that is not under /root/reference; the algorithm restated here is the published RFC 1951 one
 * (stored / fixed / dynamic blocks, canonical Huffman codes decoded by a first-level lookup table with the count/first-code walk behind it).
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
 * same branches (same-address loads broadcast, no divergence), and only the output is split -- lane i holds the i-th
 * pending output byte, the lanes stride over the bytes of a close match or of a stored block.  DPI_SYNC orders the lanes'
 * earlier stores before a match reads them back. */
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

#if defined(__GNUC__) || defined(__CUDACC__)
#define DPI_LIKELY(x)   __builtin_expect(!!(x), 1)
#define DPI_UNLIKELY(x) __builtin_expect(!!(x), 0)
#else
#define DPI_LIKELY(x)   (x)
#define DPI_UNLIKELY(x) (x)
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
/* First-level lookup of a code: the peeked low bits of the reservoir index a table of 32-bit entries in which everything
 * the symbol means is already resolved (no second table, no symbol-range compares on the hot path):
 *     bits 0-7 code length | bits 8-15 number of extra bits | bits 16-30 literal byte / length base / distance base
 * literal/length table (DPI_LBITS index bits): bit 31 marks a literal (one sign test), bit 30 marks "special": the
 *     end-of-block code when the code length is non-zero, else a miss (code longer than DPI_LBITS, or an illegal symbol);
 * distance table (DPI_DBITS index bits): 0 is a miss.
 * A miss consumes nothing and goes to the count/first-code walk, which also reports the errors. */
#ifndef DPI_LBITS
#define DPI_LBITS 10
#endif
#ifndef DPI_DBITS
#define DPI_DBITS 8
#endif
#define DPI_LSIZE (1 << DPI_LBITS)
#define DPI_DSIZE (1 << DPI_DBITS)
#define DPI_F_LITERAL 0x80000000u
#define DPI_F_SPECIAL 0x40000000u
/* scratch per decoder, in uint16 units: counts + symbols of both codes, the code-length list, the two lookup tables
 * (literal/length and distance; 32-bit entries, 4-byte aligned) */
#define DPI_SLOW_U16 (16 + DPI_NLEN + 16 + DPI_NDIST + (DPI_NLEN + DPI_NDIST + 2) / 2 + 1)
#define DPI_SCRATCH_U16 (((DPI_SLOW_U16 + 1) & ~1) + 2 * (DPI_LSIZE + DPI_DSIZE))

DPI_CONST uint16_t dpi_len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83,
                                       99, 115, 131, 163, 195, 227, 258};
DPI_CONST uint8_t dpi_len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
DPI_CONST uint16_t dpi_dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769,
                                        1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
DPI_CONST uint8_t dpi_dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11,
                                        12, 12, 13, 13};
DPI_CONST uint8_t dpi_clen_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

/* The bit reader works on the aligned 32-bit words that cover [in, in + in_len): `head` = in & 3 bytes of the first word
 * precede the stream and are skipped at the start.  It keeps a WINDOW of two words (lo, hi) and the bit position `pos`
 * inside it: nothing is shifted per symbol -- a peek is one funnel shift of the window by pos, consuming a code is an
 * addition to pos -- and when pos passes 32 the window slides by one word (two moves).  The word behind the window is
 * loaded one slide early (`ahead`), so the load's latency is off the decoding chain.  Past the last word that holds a
 * stream byte that word repeats, so the hot loop never tests for the end of the input: a stream that runs out keeps decoding
 * those bits until it ends, errs or fills the output (all bounded), and dpi_exhausted() -- consumed bits > stream bits --
 * turns that into DPI_E_INPUT afterwards.  Nothing outside the words that hold at least one stream byte is ever read. */
typedef struct dpi_stream {
    const uint8_t *in;        /* the stream itself (not aligned) */
    uint32_t in_len;
    uint32_t head;            /* in & 3 */
    uint32_t in_last;         /* last aligned word that holds a stream byte */
    uint32_t in_pos;          /* index of the word in `ahead`; the window holds words in_pos - 2 and in_pos - 1 */
    uint32_t lo, hi, ahead;
    uint32_t pos;             /* next unread bit of the window; < 32 after dpi_slide */
    uint8_t *out;
    uint32_t out_cap, out_pos;
} dpi_stream;

#ifdef __CUDACC__
#define DPI_LOAD_WORD(s, i) (((const uint32_t *)((s)->in - (s)->head))[i])
/* wrap-mode shifts and masks: only the low 5 bits of the count operand are used, so a count may carry other fields above */
#define DPI_PEEK(lo, hi, n)  __funnelshift_r((lo), (hi), (n))          /* 32 bits of hi:lo from bit (n & 31) */
#define DPI_SHR(v, n)        __funnelshift_r((v), 0u, (n))             /* v >> (n & 31) */
DPI_FN uint32_t dpi_mask(uint32_t n)                                   /* (1 << (n & 31)) - 1 */
{
    uint32_t m;
    asm("bmsk.wrap.b32 %0, 0, %1;" : "=r"(m) : "r"(n));
    return m;
}
#else
DPI_FN uint32_t dpi_load_word_host(const dpi_stream *s, uint32_t i)   /* same bytes, assembled without touching a byte outside the stream */
{
    uint32_t w = 0;
    for (uint32_t k = 0; k < 4; ++k) {
        const uint64_t at = (uint64_t)i * 4u + k;
        if (at >= s->head && at - s->head < s->in_len) w |= (uint32_t)s->in[at - s->head] << (8 * k);
    }
    return w;
}
#define DPI_LOAD_WORD(s, i) dpi_load_word_host(s, i)
#define DPI_PEEK(lo, hi, n)  ((uint32_t)(((((uint64_t)(hi)) << 32) | (lo)) >> ((n) & 31u)))
#define DPI_SHR(v, n)        ((uint32_t)(v) >> ((n) & 31u))
DPI_FN uint32_t dpi_mask(uint32_t n) { return (1u << (n & 31u)) - 1u; }
#endif
#define DPI_WORD_AT(s, i) DPI_LOAD_WORD(s, (i) < (s)->in_last ? (i) : (s)->in_last)

/* Slide the window word by word until the position is back in its first word. */
DPI_FN void dpi_slide(dpi_stream *s)
{
    while (s->pos >= 32u) {
        s->lo = s->hi;
        s->hi = s->ahead;
        s->in_pos++;
        s->ahead = DPI_WORD_AT(s, s->in_pos);
        s->pos -= 32u;
    }
}

/* Take `need` (<= 16) bits. */
DPI_FN uint32_t dpi_bits(dpi_stream *s, uint32_t need)
{
    dpi_slide(s);
    const uint32_t v = DPI_PEEK(s->lo, s->hi, s->pos) & dpi_mask(need);
    s->pos += need;
    return v;
}

/* Bits consumed so far, counted from the first aligned word; the position of the next unread bit as a byte offset from
 * `in` (rounded down); and whether more bits were consumed than the stream holds. */
DPI_FN uint64_t dpi_consumed(const dpi_stream *s) { return ((uint64_t)s->in_pos - 2u) * 32u + s->pos; }
DPI_FN uint32_t dpi_byte_pos(const dpi_stream *s) { return (uint32_t)(dpi_consumed(s) / 8u) - s->head; }
DPI_FN int dpi_exhausted(const dpi_stream *s) { return dpi_consumed(s) > ((uint64_t)s->head + s->in_len) * 8u; }

/* Continue reading at byte `off` of the stream. */
DPI_FN void dpi_seek(dpi_stream *s, uint32_t off)
{
    const uint32_t at = s->head + off, w = at >> 2;
    s->lo = DPI_WORD_AT(s, w);
    s->hi = DPI_WORD_AT(s, w + 1u);
    s->ahead = DPI_WORD_AT(s, w + 2u);
    s->in_pos = w + 2u;
    s->pos = (at & 3u) * 8u;
}

/* Canonical code described by count[len] (codes of each length) and symbol[] (symbols ordered by code): the walk behind
 * the lookup tables, and the decoder of the code-length code.  `v` = the peeked bits (>= 15 valid); *used = the length of
 * the code found.  DPI_E_SYMBOL if no code matches. */
DPI_FN int dpi_walk(uint32_t v, const uint16_t *count, const uint16_t *symbol, uint32_t *used)
{
    int32_t code = 0, first = 0, index = 0;
    for (uint32_t len = 1; len <= DPI_MAXBITS; ++len) {
        code |= (int32_t)(v & 1u);
        v >>= 1;
        const int32_t c = count[len];
        if (code - c < first) {
            *used = len;
            return symbol[index + (code - first)];
        }
        index += c;
        first = (first + c) << 1;
        code <<= 1;
    }
    *used = 0;
    return DPI_E_SYMBOL;
}

DPI_FN int dpi_decode(dpi_stream *s, const uint16_t *count, const uint16_t *symbol)
{
    dpi_slide(s);
    uint32_t used;
    const int sym = dpi_walk(DPI_PEEK(s->lo, s->hi, s->pos), count, symbol, &used);
    s->pos += used;
    return sym;
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

/* Table entry of a symbol (format above) with `len` in the code-length field; 0 for a symbol that must not occur. */
DPI_FN uint32_t dpi_entry(int sym, uint32_t len, int is_dist)
{
    if (is_dist) {
        if (sym >= DPI_NDIST) return 0;
        return len | ((uint32_t)dpi_dist_extra[sym] << 8) | ((uint32_t)dpi_dist_base[sym] << 16);
    }
    if (sym < 256) return len | ((uint32_t)sym << 16) | DPI_F_LITERAL;
    if (sym == 256) return len | DPI_F_SPECIAL;
    if (sym - 257 >= 29) return 0;
    return len | ((uint32_t)dpi_len_extra[sym - 257] << 8) | ((uint32_t)dpi_len_base[sym - 257] << 16);
}

DPI_FN uint32_t dpi_reverse(uint32_t code, uint32_t len)     /* the stream carries codes most-significant bit first */
{
#ifdef __CUDACC__
    return __brev(code) >> (32u - len);
#else
    uint32_t r = 0;
    for (uint32_t b = 0; b < len; ++b) r |= ((code >> b) & 1u) << (len - 1u - b);
    return r;
#endif
}

/* Lookup table of a code from count[]/symbol[] (RFC 1951 3.2.2: codes of one length are consecutive, shorter codes first, so
 * the i-th symbol in code order has the code first[len] + (i - number of shorter codes)).  The lanes take every DPI_LANES-th
 * symbol of the code order and write all the replicas of its entry. */
DPI_FN void dpi_build_fast(uint32_t *fast, uint32_t bits, const uint16_t *count, const uint16_t *symbol, int n, int is_dist)
{
    const uint32_t size = 1u << bits;
    DPI_SYNC();
    for (uint32_t i = DPI_LANE; i < size; i += DPI_LANES) fast[i] = is_dist ? 0u : DPI_F_SPECIAL;
    DPI_SYNC();
    const uint32_t used = (uint32_t)n - count[0];
    for (uint32_t i = DPI_LANE; i < used; i += DPI_LANES) {
        uint32_t first = 0, below = 0, len = 1;
        for (; len <= DPI_MAXBITS; ++len) {
            const uint32_t c = count[len];
            if (i < below + c) break;
            below += c;
            first = (first + c) << 1;
        }
        if (len > bits) break;                                   /* code order is length order: the rest is longer still */
        const uint32_t e = dpi_entry(symbol[i], len, is_dist);
        if (!e) continue;
        for (uint32_t idx = dpi_reverse(first + (i - below), len); idx < size; idx += 1u << len) fast[idx] = e;
    }
    DPI_SYNC();
}

/* Output goes through a batch of DPI_LANES PENDING bytes kept in registers -- lane i holds slot i -- and leaves with ONE
 * store instruction per batch (a 32-byte run per warp instruction instead of single-byte stores).  Batches cover aligned
 * DPI_LANES-byte pieces of the output (the L2 never has to fill a partly written sector from DRAM): where the output
 * position is not aligned, the first `skip` slots of the batch are dead.
 *   - a literal puts its table entry into the slot (bit 31 set, the byte in bits 16-23: one move, unpacked when the
 *     batch is closed);
 *   - a match of distance >= DPI_LANES only NOTES its distance in its slots (bit 31 clear).  Its source bytes lie
 *     behind the batch (slot j reads out[j - dist]), so they are fetched when the batch is closed: one load instruction
 *     for all the matches of a batch, issued after the previous batch has been stored;
 *   - the loaded bytes are stored one batch LATER (`p*` below is the closed batch in flight): the load's latency -- an L2
 *     round trip -- overlaps the decoding of the next 32 bytes instead of stalling every match;
 *   - only matches closer than DPI_LANES -- whose source may be pending or may be the match itself -- drain both batches and
 *     copy directly.
 * The output bound is checked when a batch is closed.
 *     base  output position of slot 0 (may be negative by `skip` at the very start)
 *     run   slots in use, dead ones included;  room = out_cap - base;  optr = out + base + lane */
#define DPI_REALIGN(pos)                                                                   \
    do {                                                                                   \
        skip = (uint32_t)((uintptr_t)(s->out + (pos)) & (DPI_LANES - 1u));                 \
        base = (int32_t)(pos) - (int32_t)skip;                                             \
        run = skip;                                                                        \
        room = s->out_cap - (uint32_t)base;                                                \
        optr = s->out + base + (int32_t)DPI_LANE;                                          \
    } while (0)
/* store the batch in flight (a no-op range when there is none) and order the stores before the loads that follow */
#define DPI_RETIRE()                                                                       \
    do {                                                                                   \
        if (DPI_LANE - pskip < prun - pskip) *pptr = (uint8_t)pval;                        \
        prun = pskip = 0;                                                                  \
        DPI_SYNC();                                                                        \
    } while (0)
/* close slots [skip, upto) of the current batch: it becomes the batch in flight, its match bytes being loaded */
#define DPI_CLOSE(upto)                                                                    \
    do {                                                                                   \
        DPI_RETIRE();                                                                      \
        pval = (myval >> 16) & 0xffu; pptr = optr; pskip = skip; prun = (upto);            \
        if ((int32_t)myval >= 0 && DPI_LANE - skip < (upto) - skip) pval = *(optr - myval); \
    } while (0)
#define DPI_LEAVE(code)                                                                    \
    do {                                                                                   \
        uint32_t ok_ = run < room ? run : room;                                            \
        if (ok_ < skip) ok_ = skip;                                                        \
        DPI_CLOSE(ok_);                                                                    \
        DPI_RETIRE();                                                                      \
        s->out_pos = (uint32_t)(base + (int32_t)ok_);                                      \
        s->lo = lo; s->hi = hi; s->ahead = ahead; s->in_pos = ipos; s->pos = pacc & 127u; \
        return (run > room && (code) == DPI_OK) ? DPI_E_OUTPUT : (code);                   \
    } while (0)
#define DPI_FLUSH()                                                                        \
    do {                                                                                   \
        if (run > room) DPI_LEAVE(DPI_E_OUTPUT);                                           \
        DPI_CLOSE(run);                                                                    \
        optr += run; base += (int32_t)run; room -= run;                                    \
        run = 0; skip = 0;                                                                 \
    } while (0)

/* the hot loop keeps the reader in locals; `pacc` is pos with other fields of the added table entries above bit 7 (only
 * its low bits are ever looked at: bit 5 = "pos >= 32", and the wrap-mode shifts take the low 5) */
#define DPI_HOT_SLIDE()                                                                    \
    do {                                                                                   \
        lo = hi; hi = ahead; ipos++;                                                       \
        ahead = DPI_LOAD_WORD(s, ipos < s->in_last ? ipos : s->in_last);                   \
        pacc -= 32u;                                                                       \
    } while (0)

DPI_FN int dpi_codes(dpi_stream *s, const uint32_t *fast, const uint32_t *dfast, const uint16_t *lcnt, const uint16_t *lsym,
                     const uint16_t *dcnt, const uint16_t *dsym)
{
    uint32_t skip, run, room, myval = 0x80000000u;
    uint32_t pskip = 0, prun = 0, pval = 0;
    int32_t base;
    uint8_t *optr, *pptr = s->out;
    dpi_slide(s);
    uint32_t lo = s->lo, hi = s->hi, ahead = s->ahead, ipos = s->in_pos, pacc = s->pos;
    DPI_REALIGN(s->out_pos);
    for (;;) {
        if (DPI_UNLIKELY(pacc & 32u)) DPI_HOT_SLIDE();
        uint32_t p = DPI_PEEK(lo, hi, pacc);                     /* 32 valid bits: a code (<= 15) and its extra bits (<= 13) */
        uint32_t e = fast[p & (DPI_LSIZE - 1)];
        pacc += e;                                               /* the code length sits in the low byte */
    dispatch:
        if (DPI_LIKELY((int32_t)e < 0)) {                        /* a literal */
            if (DPI_LANE == run) myval = e;
            if (DPI_UNLIKELY(++run == DPI_LANES)) DPI_FLUSH();
            continue;
        }
        if (DPI_UNLIKELY(e & DPI_F_SPECIAL)) {
            if (e & 0xffu) DPI_LEAVE(DPI_OK);                    /* end of block */
            uint32_t used;
            const int sym = dpi_walk(p, lcnt, lsym, &used);      /* a miss: the walk */
            if (sym < 0) DPI_LEAVE(sym);
            pacc += used;
            if (sym == 256) DPI_LEAVE(DPI_OK);
            e = dpi_entry(sym, 0, 0);
            if (!e) DPI_LEAVE(DPI_E_SYMBOL);
            p >>= used;
            goto dispatch;
        }
        uint32_t t = e >> 8;                                     /* a match: length = base + extra bits, then the distance */
        uint32_t len = (e >> 16) + (DPI_SHR(p, e) & dpi_mask(t));
        pacc += t;
        if (DPI_UNLIKELY(pacc & 32u)) DPI_HOT_SLIDE();
        p = DPI_PEEK(lo, hi, pacc);
        uint32_t de = dfast[p & (DPI_DSIZE - 1)];
        if (DPI_UNLIKELY(!de)) {
            uint32_t used;
            const int sym = dpi_walk(p, dcnt, dsym, &used);
            if (sym < 0) DPI_LEAVE(sym);
            pacc += used;
            de = dpi_entry(sym, 0, 1);
            if (!de) DPI_LEAVE(DPI_E_SYMBOL);
            p >>= used;
        }
        pacc += de;
        t = de >> 8;
        const uint32_t dist = (de >> 16) + (DPI_SHR(p, de) & dpi_mask(t));
        pacc += t;
        if (DPI_UNLIKELY(dist > (uint32_t)(base + (int32_t)run))) DPI_LEAVE(DPI_E_DISTANCE);
        if (DPI_LIKELY(dist >= DPI_LANES)) {
            if (DPI_LIKELY(run + len < DPI_LANES)) {             /* the usual case: the match ends inside the batch */
                if (DPI_LANE - run < len) myval = dist;
                run += len;
                continue;
            }
            for (;;) {
                const uint32_t n = len < DPI_LANES - run ? len : DPI_LANES - run;
                if (DPI_LANE - run < n) myval = dist;
                run += n;
                len -= n;
                if (run == DPI_LANES) DPI_FLUSH();
                if (!len) break;
            }
        } else {
            DPI_FLUSH();                                         /* run = 0: base is the output position */
            DPI_RETIRE();
            if (len > room) DPI_LEAVE(DPI_E_OUTPUT);
            uint8_t *d = s->out + base;
            const uint8_t *f = d - dist;
            if (dist >= len) {
                for (uint32_t i = DPI_LANE; i < len; i += DPI_LANES) d[i] = f[i];
            } else {                                             /* the match overlaps itself: its first `dist` bytes repeat */
                for (uint32_t i = DPI_LANE; i < len; i += DPI_LANES) d[i] = f[i % dist];
            }
            DPI_SYNC();
            DPI_REALIGN((uint32_t)base + len);
        }
    }
}

/* Inflate one raw DEFLATE stream.  scratch: DPI_SCRATCH_U16 uint16, 4-byte aligned.  Returns DPI_OK or a DPI_E_* code;
 * *out_len = bytes written. */
DPI_FN int dpi_inflate(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t out_cap, uint16_t *scratch,
                       uint32_t *out_len)
{
    dpi_stream st;
    st.in = in; st.in_len = in_len;
    st.head = (uint32_t)((uintptr_t)in & 3u);
    if (in_len == 0) { *out_len = 0; return DPI_E_INPUT; }
    st.in_last = (st.head + in_len - 1u) >> 2;
    st.out = out; st.out_cap = out_cap; st.out_pos = 0;
    dpi_seek(&st, 0);
    uint16_t *lcnt = scratch, *lsym = lcnt + 16, *dcnt = lsym + DPI_NLEN, *dsym = dcnt + 16;
    uint8_t *lengths = (uint8_t *)(dsym + DPI_NDIST);
    uint32_t *fast = (uint32_t *)(scratch + ((DPI_SLOW_U16 + 1) & ~1)), *dfast = fast + DPI_LSIZE;
    int last, rc = DPI_OK;
    do {
        last = (int)dpi_bits(&st, 1);
        const int type = (int)dpi_bits(&st, 2);
        if (type == 0) {
            st.pos = (st.pos + 7u) & ~7u;                        /* skip to the byte boundary */
            const uint32_t len = dpi_bits(&st, 16), nlen = dpi_bits(&st, 16);
            if (dpi_exhausted(&st)) { rc = DPI_E_INPUT; break; }
            if ((len ^ 0xffffu) != nlen) { rc = DPI_E_STORED; break; }
            const uint32_t at = dpi_byte_pos(&st);
            if (at + len > in_len) { rc = DPI_E_INPUT; break; }
            if (st.out_pos + len > out_cap) { rc = DPI_E_OUTPUT; break; }
            for (uint32_t i = DPI_LANE; i < len; i += DPI_LANES) out[st.out_pos + i] = in[at + i];
            DPI_SYNC();
            st.out_pos += len;
            dpi_seek(&st, at + len);
        } else if (type == 1) {
            int i = 0;
            DPI_SYNC();
            if (DPI_LANE == 0) {
                for (; i < 144; ++i) lengths[i] = 8;
                for (; i < 256; ++i) lengths[i] = 9;
                for (; i < 280; ++i) lengths[i] = 7;
                for (; i < DPI_NLEN; ++i) lengths[i] = 8;
            }
            dpi_build(lcnt, lsym, lengths, DPI_NLEN);
            dpi_build_fast(fast, DPI_LBITS, lcnt, lsym, DPI_NLEN, 0);
            if (DPI_LANE == 0)
                for (i = 0; i < DPI_NDIST; ++i) lengths[i] = 5;
            dpi_build(dcnt, dsym, lengths, DPI_NDIST);
            dpi_build_fast(dfast, DPI_DBITS, dcnt, dsym, DPI_NDIST, 1);
            rc = dpi_codes(&st, fast, dfast, lcnt, lsym, dcnt, dsym);
        } else if (type == 2) {
            const int nlen = (int)dpi_bits(&st, 5) + 257, ndist = (int)dpi_bits(&st, 5) + 1, ncode = (int)dpi_bits(&st, 4) + 4;
            if (nlen > 286 || ndist > DPI_NDIST) { rc = dpi_exhausted(&st) ? DPI_E_INPUT : DPI_E_TABLE; break; }
            int i;
            DPI_SYNC();
            for (i = 0; i < 19; ++i) lengths[i] = 0;
            for (i = 0; i < ncode; ++i) lengths[dpi_clen_order[i]] = (uint8_t)dpi_bits(&st, 3);
            if (dpi_build(lcnt, lsym, lengths, 19) != 0) { rc = DPI_E_TABLE; break; }   /* code-length code must be complete */
            i = 0;
            while (i < nlen + ndist) {
                const int sym = dpi_decode(&st, lcnt, lsym);
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
                    if (i + rep > nlen + ndist) { rc = DPI_E_TABLE; break; }
                    while (rep--) lengths[i++] = (uint8_t)prev;
                }
            }
            if (rc != DPI_OK) break;
            if (dpi_exhausted(&st)) { rc = DPI_E_INPUT; break; }
            if (lengths[256] == 0) { rc = DPI_E_TABLE; break; }                         /* no end-of-block code */
            /* the distance lengths must be read before lsym is overwritten: build the distance code first, from a copy
             * position that dpi_build of the literal code does not touch (lengths[] is separate from both tables) */
            int e = dpi_build(dcnt, dsym, lengths + nlen, ndist);
            if (e < 0 || (e > 0 && ndist - dcnt[0] != 1)) { rc = DPI_E_TABLE; break; }
            dpi_build_fast(dfast, DPI_DBITS, dcnt, dsym, ndist, 1);
            e = dpi_build(lcnt, lsym, lengths, nlen);
            if (e < 0 || (e > 0 && nlen - lcnt[0] != 1)) { rc = DPI_E_TABLE; break; }
            dpi_build_fast(fast, DPI_LBITS, lcnt, lsym, nlen, 0);
            rc = dpi_codes(&st, fast, dfast, lcnt, lsym, dcnt, dsym);
        } else {
            rc = DPI_E_BTYPE;
        }
    } while (rc == DPI_OK && !last);
    if (dpi_exhausted(&st)) rc = DPI_E_INPUT;
    *out_len = st.out_pos;
    return rc;
}

#endif /* DP_INFLATE_CORE_H */
