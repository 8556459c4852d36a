// zsb_device.cuh — device side of the batched inflate engine: bit input, Huffman table build in
// shared memory, the symbol decode loop, LZ77 copies.  Hand-written for sm_100a.
//
// Behavioural contract: per member, the observable result of the reference's
//   inflateInit2(wbits); inflate(Z_FINISH); inflateEnd()
// (reference src/inflate.c:186-224, 597-1275; src/inffast.c:50-304; src/inftrees.c:33-321).
// Nothing here is translated from those files: tables are 16-bit entries with our own layout, the
// distance code uses an 8-bit root plus a canonical-code fallback, a group of G lanes (G = 1..32)
// owns one member, and input is consumed through a 64-bit bit buffer.  What is reproduced exactly is
// the order of the checks, which input bytes count as consumed at every stop, and the accept/reject
// decisions of the table builder (including the Deflate64-only ones, src/inftrees.c:131-132,211-228).
#pragma once
#include <stdint.h>
#include "../../include/zsb.h"

namespace zsb {

#define ZSB_DEV __device__ __forceinline__

// ---- events produced by the decoder leader lane ----
enum : int { EV_NONE = 0, EV_MATCH = 1, EV_EOB = 2, EV_ERR = 3, EV_STARVE = 4, EV_OUTFULL = 5,
             EV_DONE = 6, EV_NEEDDICT = 7, EV_INITERR = 8, EV_LIT = 9,
             EV_BLOCKSTOP = 10 };    // z_stream surface only: inflate(Z_BLOCK / Z_TREES) left at a block boundary / header end

// ---- message ids (index into zsb_message(); same order as the reference's strm->msg strings) ----
enum : int { M_HDRCHECK = 1, M_METHOD = 2, M_WINDOW = 3, M_GZFLAGS = 4, M_HCRC = 5, M_BTYPE = 6,
             M_STORED = 7, M_TOOMANY = 8, M_CLENS = 9, M_REPEAT = 10, M_NOEOB = 11, M_LITLENS = 12,
             M_DISTS = 13, M_LITCODE = 14, M_DISTCODE = 15, M_TOOFAR = 16, M_DATACHECK = 17,
             M_LENCHECK = 18 };

// ---- per-member shared-memory layout, in uint16 units ----
constexpr int LROOT = 9;                 // literal/length root bits (same root as the reference, so
                                         // the table-space accounting of inftrees.c:211-228 falls out)
constexpr int DROOT = 8;                 // distance root bits (ours)
constexpr int T_L = 0;                   // [0,852)    literal/length root + sub-tables
constexpr int L_CAP = 852;               //            (bound for 286 symbols, root 9, 15-bit codes)
constexpr int T_D = 852;                 // [852,1108) distance root; doubles as lens8[320] in headers
constexpr int T_DLIM = 1108;             // [1108,1115) left-aligned code limits for lengths 9..15
constexpr int T_DBASE = 1115;            // [1115,1122) sorted-symbol index bias for lengths 9..15
constexpr int T_DSYM = 1124;             // [1124,1140) 32 bytes: distance symbols sorted by (len,sym)
constexpr int T_DL = 1140;               // [1140,1156) 32 bytes: copy of the distance code lengths
constexpr int T_CNT = 1156;              // [1156,1172) count per code length
constexpr int T_NXT = 1172;              // [1172,1188) next canonical code / sort offset per length
constexpr int STREAM_U16 = 1192;         // 2384 bytes per member
constexpr int STREAM_BYTES = STREAM_U16 * 2;

// ---- per-CTA lookup words (shared memory) ----
constexpr int LUT_LEN = 0;               // [0,64)   length symbol -> base | extra<<16   ([32..64) = Deflate64)
constexpr int LUT_DIST = 64;             // [64,128) distance symbol -> base | extra<<16 | invalid<<31
constexpr int LUT_ORDER = 128;           // [128,147) code-length-code permutation
constexpr int LUT_WORDS = 152;
constexpr int LUT_BYTES = LUT_WORDS * 4;

// table entry: bits[3:0] code bits at this level, bits[5:4] kind, bits[15:6] payload
constexpr uint32_t K_LIT = 0u << 4, K_LEN = 1u << 4, K_SPECIAL = 2u << 4, K_LINK = 3u << 4;
constexpr uint32_t E_EOB = K_SPECIAL | (0u << 6);
constexpr uint32_t E_INVALID = K_SPECIAL | (1u << 6);

// what the decode kernel hands to the checksum/trailer kernel
struct Aux {
    uint32_t kind;         // 0 raw, 1 zlib (Adler-32), 2 gzip (CRC-32)
    uint32_t trailer_pos;  // byte offset of the trailer when need_trailer
    uint32_t need_trailer; // 1: all blocks decoded, trailer still to be checked
    uint32_t init_check;   // strm->adler if no output byte is produced
    uint32_t nseq;         // lane path: LZ sequences k_lz has to execute for this member
    uint32_t pad[3];
};

// ------------------------------------------------------------------------------------------------
// Bit input: 64-bit buffer, refilled by aligned 32-bit loads while at least one whole word is left,
// byte-wise for the last bytes.  bit position = 8*next - nbits.
// ------------------------------------------------------------------------------------------------
struct BitIn {
    const uint8_t *p;
    uint32_t len;
    uint32_t next;
    uint32_t nbits;
    uint64_t hold;
};

ZSB_DEV void bi_seek(BitIn &b, uint32_t bytepos) {
    b.next = bytepos;
    b.hold = 0;
    b.nbits = 0;
    while (b.next < b.len && ((uintptr_t)(b.p + b.next) & 3)) {
        b.hold |= (uint64_t)b.p[b.next++] << b.nbits;
        b.nbits += 8;
    }
}
ZSB_DEV void bi_loadword(BitIn &b) {
    uint32_t w = __ldg((const uint32_t *)(b.p + b.next));
    b.hold |= (uint64_t)w << b.nbits;
    b.nbits += 32;
    b.next += 4;
}
// at least 32 bits while whole words are left, else every byte that is left (up to 64 bits)
ZSB_DEV void bi_fill(BitIn &b) {
    if (b.next + 4 <= b.len) {           // whole words left: keep `next` word aligned
        if (b.nbits < 32) bi_loadword(b);
        return;
    }
    while (b.nbits <= 56 && b.next < b.len) {
        b.hold |= (uint64_t)b.p[b.next++] << b.nbits;
        b.nbits += 8;
    }
}
// make n (<= 32) bits available; false = input exhausted first (everything left is then loaded)
ZSB_DEV bool bi_need(BitIn &b, uint32_t n) {
    if (b.nbits >= n) return true;
    bi_fill(b);
    return b.nbits >= n;
}
ZSB_DEV uint32_t bi_peek(const BitIn &b, uint32_t n) { return (uint32_t)b.hold & ((1u << n) - 1u); }
ZSB_DEV void bi_drop(BitIn &b, uint32_t n) {
    b.hold >>= n;
    b.nbits -= n;
}
ZSB_DEV uint64_t bi_bitpos(const BitIn &b) { return 8ull * b.next - b.nbits; }

// ------------------------------------------------------------------------------------------------
// Leader-lane decoder state
// ------------------------------------------------------------------------------------------------
struct Dec {
    BitIn b;
    uint8_t *out;
    uint32_t pos, cap;
    uint32_t d64;        // 0 / 1
    uint32_t last;       // BFINAL of the current block
    uint32_t at_type;    // stopped exactly at a block boundary reached through mode TYPE (+128 in data_type)
    int msg;
    uint32_t extent;     // bits looked at beyond the bit position when an error fired between need and drop
    uint32_t mlen, mdist, mtrunc;
};

// ---- group helpers -------------------------------------------------------------------------------
template <int G> ZSB_DEV uint32_t grp_bcast(uint32_t v, unsigned gmask) {
    if (G == 1) return v;
    return __shfl_sync(gmask, v, 0, G);
}
template <int G> ZSB_DEV void grp_sync(unsigned gmask) {
    if (G > 1) __syncwarp(gmask);
}

// ------------------------------------------------------------------------------------------------
// Table construction (leader lane, serial; per block)
// ------------------------------------------------------------------------------------------------
ZSB_DEV uint32_t rev_bits(uint32_t code, uint32_t len) { return __brev(code) >> (32 - len); }

// Count code lengths, check the set (over-subscribed / incomplete), leave first canonical codes in
// nxt[].  Returns: 0 complete, 1 no codes at all, 2 single 1-bit code (allowed incomplete),
// -1 invalid.  maxl receives the longest length.
ZSB_DEV int code_census(const uint8_t *lens, uint32_t n, uint16_t *cnt, uint16_t *nxt, bool allow_single,
                        uint32_t &maxl) {
#pragma unroll
    for (int i = 0; i < 16; i++) cnt[i] = 0;
    for (uint32_t i = 0; i < n; i++) cnt[lens[i]]++;
    maxl = 15;
    while (maxl >= 1 && cnt[maxl] == 0) maxl--;
    if (maxl == 0) return 1;
    int left = 1;
    for (uint32_t l = 1; l <= 15; l++) {
        left = left * 2 - (int)cnt[l];
        if (left < 0) return -1;
    }
    uint32_t code = 0;
    for (uint32_t l = 1; l <= 15; l++) {
        nxt[l] = (uint16_t)code;
        code = (code + cnt[l]) << 1;
    }
    if (left > 0) return (allow_single && maxl == 1) ? 2 : -1;
    return 0;
}

// code-length code: 19 symbols, 7-bit root, entry = bits | sym<<4.  Returns false -> "invalid code lengths set".
ZSB_DEV bool build_clen(uint16_t *T, bool d64) {
    const uint8_t *lens = (const uint8_t *)(T + T_D);
    uint16_t *cnt = T + T_CNT, *nxt = T + T_NXT, *C = T + T_L;
    uint32_t maxl;
    int k = code_census(lens, 19, cnt, nxt, false, maxl);
    if (k < 0) return false;
    if (k == 1) {                    // no codes: every lookup yields "length 0, one bit" (the caller
        if (d64) return false;       // never looks at the invalid marker; rejected outright in Deflate64)
        for (int i = 0; i < 128; i++) C[i] = 1;
        return true;
    }
    for (uint32_t s = 0; s < 19; s++) {
        uint32_t l = lens[s];
        if (!l) continue;
        uint32_t c = nxt[l]++;
        uint32_t e = l | (s << 4);
        for (uint32_t i = rev_bits(c, l); i < 128; i += 1u << l) C[i] = (uint16_t)e;
    }
    return true;
}

ZSB_DEV uint32_t litlen_entry(uint32_t sym) {
    if (sym < 256) return K_LIT | (sym << 6);
    if (sym == 256) return E_EOB;
    if (sym < 286) return K_LEN | ((sym - 257) << 6);
    return E_INVALID;
}

// literal/length code from lens8[0..n) -> T[T_L..].  false -> "invalid literal/lengths set".
ZSB_DEV bool build_litlen(uint16_t *T, uint32_t n, bool d64) {
    const uint8_t *lens = (const uint8_t *)(T + T_D);
    uint16_t *cnt = T + T_CNT, *nxt = T + T_NXT, *L = T + T_L;
    uint32_t maxl;
    int k = code_census(lens, n, cnt, nxt, true, maxl);
    if (k < 0 || k == 1) return false;
    if (k == 2)
        for (int i = 0; i < 512; i++) L[i] = (uint16_t)(E_INVALID | 1);
    // sub-tables: one per 9-bit prefix that carries codes longer than 9 bits, sized by the longest
    // code under that prefix; prefixes are met in canonical (increasing) order.
    uint32_t used = 512, plo = 512;
    for (uint32_t l = LROOT + 1; l <= maxl; l++) {
        if (!cnt[l]) continue;
        uint32_t lo = nxt[l] >> (l - LROOT), hi = (nxt[l] + cnt[l] - 1u) >> (l - LROOT);
        if (plo == 512) plo = lo;
        for (uint32_t p = lo; p <= hi; p++) L[rev_bits(p, LROOT)] = (uint16_t)(K_LINK | (l - LROOT));
    }
    for (uint32_t p = plo; p < 512; p++) {
        uint32_t ri = rev_bits(p, LROOT);
        uint32_t curr = L[ri] & 15u;
        L[ri] = (uint16_t)(K_LINK | curr | (used << 6));
        used += 1u << curr;
    }
    if (used > (d64 ? 851u : 852u)) return false;      // inftrees.c:211-212,226-228
    for (uint32_t s = 0; s < n; s++) {
        uint32_t l = lens[s];
        if (!l) continue;
        uint32_t r = rev_bits(nxt[l]++, l);
        uint32_t e = litlen_entry(s);
        if (l <= LROOT) {
            e |= l;
            for (uint32_t i = r; i < 512; i += 1u << l) L[i] = (uint16_t)e;
        } else {
            uint32_t lk = L[r & 511u];
            uint32_t off = lk >> 6, size = 1u << (lk & 15u);
            e |= (l - LROOT);
            for (uint32_t i = r >> LROOT; i < size; i += 1u << (l - LROOT)) L[off + i] = (uint16_t)e;
        }
    }
    return true;
}

// distance code from dl8[0..n) -> T[T_D..], fallback arrays for codes longer than 8 bits.
// false -> "invalid distances set".
ZSB_DEV bool build_dist(uint16_t *T, uint32_t n, bool d64) {
    const uint8_t *lens = (const uint8_t *)(T + T_DL);
    uint16_t *cnt = T + T_CNT, *nxt = T + T_NXT, *D = T + T_D;
    uint8_t *dsym = (uint8_t *)(T + T_DSYM);
    uint32_t maxl;
    int k = code_census(lens, n, cnt, nxt, true, maxl);
    if (k < 0) return false;
    if (k == 1) {                                       // inftrees.c:131-139
        if (d64) return false;
        for (int i = 0; i < 256; i++) D[i] = (uint16_t)(E_INVALID | 1);
        return true;
    }
    if (d64 && maxl > 6) {
        // table space the reference's builder (root 6) would need; Deflate64 rejects > 593
        uint32_t used = 64, last_p = 0xffffffffu, last_curr = 0;
        for (uint32_t l = 7; l <= maxl; l++) {
            if (!cnt[l]) continue;
            uint32_t lo = nxt[l] >> (l - 6), hi = (nxt[l] + cnt[l] - 1u) >> (l - 6);
            if (lo == last_p) {
                used += (1u << (l - 6)) - (1u << last_curr);
                lo++;
            }
            if (hi + 1 > lo) used += (hi + 1 - lo) << (l - 6);
            last_p = hi;
            last_curr = l - 6;
        }
        if (used > 593u) return false;
    }
    if (k == 2)
        for (int i = 0; i < 256; i++) D[i] = (uint16_t)(E_INVALID | 1);
    for (uint32_t l = DROOT + 1; l <= 15; l++)
        T[T_DLIM + l - (DROOT + 1)] = (uint16_t)((uint32_t)(nxt[l] + cnt[l]) << (15 - l));
    // cnt[] -> sort offsets
    {
        uint32_t o = 0;
        for (uint32_t l = 1; l <= 15; l++) {
            uint32_t c = cnt[l];
            cnt[l] = (uint16_t)o;
            o += c;
        }
    }
    for (uint32_t l = DROOT + 1; l <= 15; l++) T[T_DBASE + l - (DROOT + 1)] = (uint16_t)(cnt[l] - nxt[l]);
    for (uint32_t s = 0; s < n; s++) {
        uint32_t l = lens[s];
        if (!l) continue;
        uint32_t r = rev_bits(nxt[l]++, l);
        dsym[cnt[l]++] = (uint8_t)s;
        if (l <= DROOT) {
            uint32_t e = l | (s << 6);
            for (uint32_t i = r; i < 256; i += 1u << l) D[i] = (uint16_t)e;
        } else {
            D[r & 255u] = (uint16_t)K_LINK;
        }
    }
    return true;
}

// lengths of the fixed code (reference src/inflate.c:271-289) -> tables
ZSB_DEV void build_fixed(uint16_t *T) {
    uint8_t *lens = (uint8_t *)(T + T_D);
    for (int i = 0; i < 288; i++) lens[i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
    build_litlen(T, 288, false);
    uint8_t *dl = (uint8_t *)(T + T_DL);
    for (int i = 0; i < 32; i++) dl[i] = 5;
    build_dist(T, 32, false);
}

// ------------------------------------------------------------------------------------------------
// Code lookups.  Both take the (zero padded) bit buffer and return the entry; nb = total code bits.
// ------------------------------------------------------------------------------------------------
ZSB_DEV uint32_t lookup_litlen(const uint16_t *L, uint64_t hold, uint32_t &nb) {
    uint32_t e = L[(uint32_t)hold & 511u];
    if ((e & 0x30u) == K_LINK) {
        uint32_t curr = e & 15u;
        e = L[(e >> 6) + (((uint32_t)(hold >> LROOT)) & ((1u << curr) - 1u))];
        nb = LROOT + (e & 15u);
    } else {
        nb = e & 15u;
    }
    return e;
}
ZSB_DEV uint32_t lookup_dist(const uint16_t *T, uint64_t hold, uint32_t &nb) {
    uint32_t e = T[T_D + ((uint32_t)hold & 255u)];
    if ((e & 0x30u) == K_LINK) {
        uint32_t c15 = __brev((uint32_t)hold) >> 17;
        uint32_t l = DROOT + 1;
        while (l < 15 && c15 >= T[T_DLIM + l - (DROOT + 1)]) l++;
        uint32_t idx = (T[T_DBASE + l - (DROOT + 1)] + (c15 >> (15 - l))) & 31u;
        e = ((uint32_t)((const uint8_t *)(T + T_DSYM))[idx]) << 6;
        nb = l;
    } else {
        nb = e & 15u;
    }
    return e;
}

// ------------------------------------------------------------------------------------------------
// Dynamic block header (reference src/inflate.c:905-1033).  Leader lane.  Returns EV_NONE when the
// tables are ready, else EV_ERR (d.msg set) or EV_STARVE.
// ------------------------------------------------------------------------------------------------
ZSB_DEV int dynamic_header(Dec &d, uint16_t *T, const uint32_t *lut) {
    BitIn &b = d.b;
    uint8_t *lens = (uint8_t *)(T + T_D);
    if (!bi_need(b, 14)) return EV_STARVE;
    uint32_t nlen = bi_peek(b, 5) + 257;
    bi_drop(b, 5);
    uint32_t ndist = bi_peek(b, 5) + 1;
    bi_drop(b, 5);
    uint32_t ncode = bi_peek(b, 4) + 4;
    bi_drop(b, 4);
    if (nlen > 286 || (!d.d64 && ndist > 30)) {
        d.msg = M_TOOMANY;
        return EV_ERR;
    }
    const uint8_t *order = (const uint8_t *)(lut + LUT_ORDER);
    for (uint32_t i = 0; i < 19; i++) {
        uint32_t v = 0;
        if (i < ncode) {
            if (!bi_need(b, 3)) return EV_STARVE;
            v = bi_peek(b, 3);
            bi_drop(b, 3);
        }
        lens[order[i]] = (uint8_t)v;
    }
    if (!build_clen(T, d.d64)) {
        d.msg = M_CLENS;
        return EV_ERR;
    }
    const uint16_t *C = T + T_L;
    uint32_t have = 0, total = nlen + ndist;
    while (have < total) {
        if (b.nbits < 32) bi_fill(b);
        uint32_t e = C[(uint32_t)b.hold & 127u];
        uint32_t nb = e & 15u, sym = e >> 4;
        if (nb > b.nbits) return EV_STARVE;
        if (sym < 16) {
            bi_drop(b, nb);
            lens[have++] = (uint8_t)sym;
            continue;
        }
        uint32_t xb = sym == 16 ? 2u : sym == 17 ? 3u : 7u;
        if (nb + xb > b.nbits) return EV_STARVE;
        bi_drop(b, nb);
        uint32_t val = 0;
        if (sym == 16) {
            if (have == 0) {
                d.msg = M_REPEAT;
                d.extent = 2;         // the two repeat bits were pulled but not dropped (inflate.c:959-966)
                return EV_ERR;
            }
            val = lens[have - 1];
        }
        uint32_t rep = (sym == 18 ? 11u : 3u) + bi_peek(b, xb);
        bi_drop(b, xb);
        if (have + rep > total) {
            d.msg = M_REPEAT;
            return EV_ERR;
        }
        while (rep--) lens[have++] = (uint8_t)val;
    }
    if (lens[256] == 0) {
        d.msg = M_NOEOB;
        return EV_ERR;
    }
    // keep the distance lengths aside: the distance table is built over the lens8 area
    uint8_t *dl = (uint8_t *)(T + T_DL);
    for (uint32_t i = 0; i < ndist; i++) dl[i] = lens[nlen + i];
    if (!build_litlen(T, nlen, d.d64)) {
        d.msg = M_LITLENS;
        return EV_ERR;
    }
    if (!build_dist(T, ndist, d.d64)) {
        d.msg = M_DISTS;
        return EV_ERR;
    }
    return EV_NONE;
}

// ------------------------------------------------------------------------------------------------
// Symbol decoding (leader lane).  decode_burst() turns symbols into TOKENS in a small shared-memory
// queue without touching the output: the group then executes the queued tokens together
// (run_tokens), so the latency of reading earlier output is paid once per burst instead of once per
// match.  A token is  literal: TOK_LIT | byte   or   match: (dist-1) | len<<16  (len <= TOK_MAXLEN).
// The leader keeps the output position, so capacity and "distance too far back" are decided here,
// at the symbol where the reference decides them (src/inflate.c:1143-1153, src/inffast.c:166-175).
// Symbols that start with at least two whole input words still unread take the unchecked path; the
// last bytes of a member go through tail_symbol(), which checks availability before every step
// exactly where the reference would wait for input (src/inflate.c:1046-1142).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t TOK_LIT = 0x80000000u;
constexpr uint32_t TOK_MAXLEN = 511;
constexpr int QCAP = 24;                  // tokens per burst; the queue overlays T_DL..T_NXT (idle then)
constexpr int T_Q = T_DL;                 // 25 words of the member slot

ZSB_DEV int tail_symbol(Dec &d, const uint16_t *T, const uint32_t *lut) {
    BitIn &b = d.b;
    bi_fill(b);
    uint32_t nb;
    uint32_t e = lookup_litlen(T + T_L, b.hold, nb);
    if (nb > b.nbits) return EV_STARVE;
    bi_drop(b, nb);
    uint32_t kind = e & 0x30u, v = e >> 6;
    if (kind == K_LIT) {
        d.mdist = v;
        return EV_LIT;
    }
    if (kind == K_SPECIAL) {
        if (v == 0) return EV_EOB;
        d.msg = M_LITCODE;
        return EV_ERR;
    }
    uint32_t lw = lut[LUT_LEN + d.d64 * 32 + v];
    uint32_t xb = lw >> 16;
    if (!bi_need(b, xb)) return EV_STARVE;
    d.mlen = (lw & 0xffffu) + (xb ? bi_peek(b, xb) : 0u);
    bi_drop(b, xb);
    bi_fill(b);
    e = lookup_dist(T, b.hold, nb);
    if (nb > b.nbits) return EV_STARVE;
    bi_drop(b, nb);
    if ((e & 0x30u) == K_SPECIAL) {
        d.msg = M_DISTCODE;
        return EV_ERR;
    }
    uint32_t dw = lut[LUT_DIST + d.d64 * 32 + (e >> 6)];
    if (dw >> 31) {
        d.msg = M_DISTCODE;
        return EV_ERR;
    }
    xb = (dw >> 16) & 15u;
    if (!bi_need(b, xb)) return EV_STARVE;
    d.mdist = (dw & 0xffffu) + (xb ? bi_peek(b, xb) : 0u);
    bi_drop(b, xb);
    return EV_MATCH;
}

// One symbol -> one token.  Returns EV_NONE after queueing a token, EV_MATCH for a match too long for
// a token (d.mlen bytes to copy now, d.mtrunc set when the capacity cut it), else the stop event.
// Called in lock step by all leader lanes of a warp (the caller reconverges after every step).
ZSB_DEV int decode_step(Dec &d, const uint16_t *T, const uint32_t *lut, uint32_t *q, uint32_t &nt) {
    BitIn &b = d.b;
    const uint16_t *L = T + T_L;
    const uint32_t lutoff = d.d64 * 32;
    {
        if (b.next + 8 > b.len) {
            int ev = tail_symbol(d, T, lut);
            if (ev == EV_LIT) {
                if (d.pos >= d.cap) return EV_OUTFULL;
                q[nt++] = TOK_LIT | d.mdist;
                d.pos++;
                return EV_NONE;
            }
            if (ev != EV_MATCH) return ev;
        } else {
            if (b.nbits < 32) bi_loadword(b);
            uint32_t e = L[(uint32_t)b.hold & 511u];
            if ((e & 0x30u) == K_LINK) {
                uint32_t curr = e & 15u;
                e = L[(e >> 6) + (((uint32_t)(b.hold >> LROOT)) & ((1u << curr) - 1u))];
                bi_drop(b, LROOT);
            }
            bi_drop(b, e & 15u);
            uint32_t kind = e & 0x30u, v = e >> 6;
            if (kind == K_LIT) {
                if (d.pos >= d.cap) return EV_OUTFULL;
                q[nt++] = TOK_LIT | v;
                d.pos++;
                return EV_NONE;
            }
            if (kind == K_SPECIAL) {
                if (v == 0) return EV_EOB;
                d.msg = M_LITCODE;
                return EV_ERR;
            }
            uint32_t lw = lut[LUT_LEN + lutoff + v];
            uint32_t xb = lw >> 16;
            d.mlen = (lw & 0xffffu) + ((uint32_t)b.hold & ((1u << xb) - 1u));
            bi_drop(b, xb);
            if (b.nbits < 32) bi_loadword(b);
            uint32_t nb;
            e = lookup_dist(T, b.hold, nb);
            bi_drop(b, nb);
            if ((e & 0x30u) == K_SPECIAL) {
                d.msg = M_DISTCODE;
                return EV_ERR;
            }
            uint32_t dw = lut[LUT_DIST + lutoff + (e >> 6)];
            if (dw >> 31) {
                d.msg = M_DISTCODE;
                return EV_ERR;
            }
            xb = (dw >> 16) & 15u;
            d.mdist = (dw & 0xffffu) + ((uint32_t)b.hold & ((1u << xb) - 1u));
            bi_drop(b, xb);
        }
        // a match: capacity first, then the distance (src/inflate.c:1144-1153)
        if (d.pos >= d.cap) return EV_OUTFULL;
        if (d.mdist > d.pos) {
            d.msg = M_TOOFAR;
            return EV_ERR;
        }
        uint32_t n = d.cap - d.pos;
        d.mtrunc = n < d.mlen;
        if (n > d.mlen) n = d.mlen;
        if (n > TOK_MAXLEN) {
            d.mlen = n;
            return EV_MATCH;
        }
        q[nt++] = (d.mdist - 1u) | (n << 16);
        d.pos += n;
        if (d.mtrunc) return EV_OUTFULL;
    }
    return EV_NONE;
}

// ------------------------------------------------------------------------------------------------
// LZ77 copy: out[pos+i] = out[pos-dist + (i mod dist)], i < n.  Every source byte lies before pos, so
// lanes never depend on each other's stores of this copy.
// ------------------------------------------------------------------------------------------------
// one lane, serial: loads of a step are issued before its stores so they overlap
ZSB_DEV void lane_copy(uint8_t *out, uint32_t pos, uint32_t dist, uint32_t n) {
    uint8_t *dst = out + pos;
    const uint8_t *src = dst - dist;
    if (dist >= 4) {
        while (n >= 4) {
            uint8_t a = src[0], bb = src[1], c = src[2], e = src[3];
            dst[0] = a; dst[1] = bb; dst[2] = c; dst[3] = e;
            src += 4; dst += 4; n -= 4;
        }
        while (n--) *dst++ = *src++;
    } else {
        uint8_t p0 = src[0], p1 = dist > 1 ? src[1] : p0, p2 = dist > 2 ? src[2] : p0;
        uint32_t ph = 0;                      // pattern of period dist (1..3)
        for (uint32_t i = 0; i < n; i++) {
            dst[i] = ph == 0 ? p0 : ph == 1 ? p1 : p2;
            if (++ph == dist) ph = 0;
        }
    }
}

// all G lanes on one (long) match
template <int G>
ZSB_DEV void grp_copy_match(uint8_t *out, uint32_t pos, uint32_t dist, uint32_t n, int sub) {
    if (G == 1) {
        lane_copy(out, pos, dist, n);
        return;
    }
    const uint8_t *src = out + pos - dist;
    uint8_t *dst = out + pos;
    if (dist >= n) {
        for (uint32_t i = sub; i < n; i += G) dst[i] = src[i];
    } else {
        uint32_t off = (uint32_t)sub % dist, step = (uint32_t)G % dist;
        for (uint32_t i = sub; i < n; i += G) {
            dst[i] = src[off];
            off += step;
            if (off >= dist) off -= dist;
        }
    }
}

// Execute a burst of nt queued tokens starting at output position pos0.  G tokens at a time, one per
// lane: positions by a prefix sum over the group, literals stored at once, a match copied as soon as
// every token before its source range's end is finished (the first unfinished token always is), so
// independent matches of a chunk overlap their memory latency.
template <int G>
ZSB_DEV void run_tokens(uint8_t *out, uint32_t pos0, const uint32_t *q, uint32_t nt, int sub, unsigned gmask) {
    if (G == 1) {
        uint32_t pos = pos0;
        for (uint32_t t = 0; t < nt; t++) {
            uint32_t tok = q[t];
            if (tok & TOK_LIT) {
                out[pos++] = (uint8_t)tok;
            } else {
                uint32_t n = tok >> 16;
                lane_copy(out, pos, (tok & 0xffffu) + 1u, n);
                pos += n;
            }
        }
        return;
    }
    uint32_t base = pos0;
    for (uint32_t c = 0; c < nt; c += G) {
        const uint32_t t = c + sub;
        const bool have = t < nt;
        const uint32_t tok = have ? q[t] : TOK_LIT;
        const bool lit = (tok & TOK_LIT) != 0;
        const uint32_t n = have ? (lit ? 1u : tok >> 16) : 0u;
        const uint32_t dist = (tok & 0xffffu) + 1u;
        uint32_t incl = n;
#pragma unroll
        for (int o = 1; o < G; o <<= 1) {
            uint32_t v = __shfl_up_sync(gmask, incl, o, G);
            if (sub >= o) incl += v;
        }
        const uint32_t mypos = base + incl - n;
        base += __shfl_sync(gmask, incl, G - 1, G);
        if (have && lit) out[mypos] = (uint8_t)tok;
        bool done = lit;
        const uint32_t src_hi = dist >= n ? mypos - dist + n : mypos;   // end of the bytes this match reads
        __syncwarp(gmask);
        for (;;) {
            const unsigned pending = __ballot_sync(gmask, !done) & gmask;
            if (!pending) break;
            const uint32_t frontier = __shfl_sync(gmask, mypos, __ffs(pending) - 1);
            if (!done && src_hi <= frontier) {
                lane_copy(out, mypos, dist, n);
                done = true;
            }
            __syncwarp(gmask);
        }
    }
}

template <int G>
ZSB_DEV void grp_copy_stored(uint8_t *dst, const uint8_t *src, uint32_t n, int sub) {
    for (uint32_t i = sub; i < n; i += G) dst[i] = __ldg(src + i);
}

// ------------------------------------------------------------------------------------------------
// Wrapper headers (reference src/inflate.c:629-830).  Leader lane, byte cursor.  On success the bit
// reader is positioned on the first block.  Result fields for the early exits are filled directly.
// ------------------------------------------------------------------------------------------------
ZSB_DEV uint32_t crc32_bytes(uint32_t crc, const uint8_t *p, uint32_t n) {
    crc = ~crc;
    for (uint32_t i = 0; i < n; i++) {
        crc ^= p[i];
        for (int k = 0; k < 8; k++) crc = (crc >> 1) ^ (0xEDB88320u & (0u - (crc & 1u)));
    }
    return ~crc;
}

struct HeadOut {
    int ev;              // EV_NONE: data follows at `pos`
    uint32_t pos;        // byte offset of the first block
    uint32_t pulled;     // total_in for EV_ERR
    uint32_t bits;       // bits left in the reference's bit buffer at the stop (data_type low bits)
    uint32_t kind;       // 0 raw, 1 zlib, 2 gzip
    uint32_t adler;      // strm->adler so far
    int msg;
};

ZSB_DEV HeadOut parse_wrapper(const uint8_t *in, uint32_t len, int wbits_arg) {
    HeadOut h;
    h.ev = EV_NONE; h.pos = 0; h.pulled = 0; h.bits = 0; h.kind = 0; h.adler = 0; h.msg = 0;
    if (wbits_arg < 0) return h;
    uint32_t wrap = ((uint32_t)wbits_arg >> 4) + 5;
    uint32_t wbits = (uint32_t)wbits_arg & 15u;
    h.adler = wrap & 1u;
    h.kind = 1;
    if (len < 2) { h.ev = EV_STARVE; h.bits = 8 * len; return h; }
    uint32_t b0 = in[0], b1 = in[1];
    if ((wrap & 2) && b0 == 0x1f && b1 == 0x8b) {
        h.kind = 2;
        uint32_t crc = crc32_bytes(0, in, 2);
        uint32_t pos = 2;
        if (len - pos < 2) { h.ev = EV_STARVE; h.bits = 8 * (len - pos); return h; }
        uint32_t flags = in[2] | ((uint32_t)in[3] << 8);
        if ((flags & 0xff) != 8) { h.ev = EV_ERR; h.msg = M_METHOD; h.pulled = 4; h.bits = 16; return h; }
        if (flags & 0xe000) { h.ev = EV_ERR; h.msg = M_GZFLAGS; h.pulled = 4; h.bits = 16; return h; }
        bool hc = (flags & 0x0200) != 0;
        if (hc) crc = crc32_bytes(crc, in + pos, 2);
        pos = 4;
        if (len - pos < 4) { h.ev = EV_STARVE; h.bits = 8 * (len - pos); return h; }
        if (hc) crc = crc32_bytes(crc, in + pos, 4);
        pos = 8;
        if (len - pos < 2) { h.ev = EV_STARVE; h.bits = 8 * (len - pos); return h; }
        if (hc) crc = crc32_bytes(crc, in + pos, 2);
        pos = 10;
        if (flags & 0x0400) {
            if (len - pos < 2) { h.ev = EV_STARVE; h.bits = 8 * (len - pos); return h; }
            uint32_t xlen = in[pos] | ((uint32_t)in[pos + 1] << 8);
            if (hc) crc = crc32_bytes(crc, in + pos, 2);
            pos += 2;
            uint32_t copy = xlen < len - pos ? xlen : len - pos;
            if (hc) crc = crc32_bytes(crc, in + pos, copy);
            pos += copy;
            if (copy < xlen) { h.ev = EV_STARVE; return h; }
        }
        for (uint32_t bit = 0x0800; bit <= 0x1000; bit <<= 1) {
            if (!(flags & bit)) continue;
            if (pos >= len) { h.ev = EV_STARVE; return h; }
            uint32_t k = 0, c;
            do { c = in[pos + k++]; } while (c && pos + k < len);
            if (hc) crc = crc32_bytes(crc, in + pos, k);
            pos += k;
            if (c) { h.ev = EV_STARVE; return h; }
        }
        if (hc) {
            if (len - pos < 2) { h.ev = EV_STARVE; h.bits = 8 * (len - pos); return h; }
            uint32_t v = in[pos] | ((uint32_t)in[pos + 1] << 8);
            if (v != (crc & 0xffffu)) { h.ev = EV_ERR; h.msg = M_HCRC; h.pulled = pos + 2; h.bits = 16; return h; }
            pos += 2;
        }
        h.adler = 0;
        h.pos = pos;
        return h;
    }
    h.pulled = 2;
    h.bits = 16;
    if (!(wrap & 1) || ((b0 << 8) + b1) % 31) { h.ev = EV_ERR; h.msg = M_HDRCHECK; return h; }
    if ((b0 & 15) != 8) { h.ev = EV_ERR; h.msg = M_METHOD; return h; }
    h.bits = 12;                                           // the method nibble was dropped first
    uint32_t wl = (b0 >> 4) + 8;
    if (wbits == 0) wbits = wl;
    if (wl > 15 || wl > wbits) { h.ev = EV_ERR; h.msg = M_WINDOW; return h; }
    h.adler = 1;
    h.bits = 0;
    if (b1 & 0x20) {
        if (len - 2 < 4) { h.ev = EV_STARVE; h.bits = 8 * (len - 2); return h; }
        h.adler = ((uint32_t)in[2] << 24) | ((uint32_t)in[3] << 16) | ((uint32_t)in[4] << 8) | in[5];
        h.ev = EV_NEEDDICT;
        return h;
    }
    h.pos = 2;
    return h;
}

// ------------------------------------------------------------------------------------------------
// Trailer (reference src/inflate.c:1194-1232): `check` is the Adler-32 / CRC-32 of the E output bytes
// (or strm->adler's initial value when E == 0).  Completes the result the decode kernel left pending.
// ------------------------------------------------------------------------------------------------
ZSB_DEV uint32_t le32(const uint8_t *p) {
    return p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}
ZSB_DEV void finish_member(zsb_result &r, const Aux &a, const uint8_t *in, uint32_t len, uint32_t check,
                           uint32_t E) {
    r.check = check;
    if (!a.need_trailer) return;
    uint32_t pos = a.trailer_pos;
    const int dt_last = 64;
    if (len - pos < 4) {
        r.ret = -5; r.total_in = len; r.data_type = (int)(8 * (len - pos)) + dt_last;
        return;
    }
    uint32_t v = le32(in + pos);
    if (a.kind == 1) v = (v >> 24) | ((v >> 8) & 0xff00u) | ((v & 0xff00u) << 8) | (v << 24);
    pos += 4;
    if (v != check) {
        r.ret = -3; r.msg_id = M_DATACHECK; r.total_in = pos; r.data_type = 32 + dt_last;
        return;
    }
    if (a.kind == 2) {
        if (len - pos < 4) {
            r.ret = -5; r.total_in = len; r.data_type = (int)(8 * (len - pos)) + dt_last;
            return;
        }
        uint32_t isz = le32(in + pos);
        pos += 4;
        if (isz != E) {
            r.ret = -3; r.msg_id = M_LENCHECK; r.total_in = pos; r.data_type = 32 + dt_last;
            return;
        }
    }
    r.ret = 1; r.total_in = pos; r.data_type = dt_last;
}

// ------------------------------------------------------------------------------------------------
// One member, start to finish (all G lanes of the group call this; the leader lane holds the state)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void init_lut(uint32_t *lut, int tid, int nthreads) {
    for (int i = tid; i < LUT_WORDS; i += nthreads) {
        uint32_t v = 0;
        if (i < 64) {                                  // length symbols 257..285 (RFC 1951 3.2.5)
            uint32_t s = i & 31, d64 = i >> 5;
            if (s < 8) v = 3 + s;
            else if (s < 28) {
                uint32_t xb = (s >> 2) - 1;
                v = (3 + ((4 + (s & 3)) << xb)) | (xb << 16);
            } else if (s == 28) v = d64 ? (3u | (16u << 16)) : 258u;   // inftrees.c:59-79
        } else if (i < 128) {                          // distance symbols
            uint32_t s = i & 31, d64 = (i - 64) >> 5;
            if (s < 4) v = 1 + s;
            else {
                uint32_t xb = (s >> 1) - 1;
                v = (1 + ((2 + (s & 1)) << xb)) | (xb << 16);
            }
            if (s >= 30 && !d64) v = 0x80000000u;       // inftrees.c:65-72: codes 30,31 invalid in deflate
        } else if (i < 128 + 5) {
            // permutation of the code-length code lengths (RFC 1951 3.2.7), 4 per word
            const unsigned char ord[20] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15, 0};
            int k = (i - 128) * 4;
            v = ord[k] | (ord[k + 1] << 8) | (ord[k + 2] << 16) | ((uint32_t)ord[k + 3] << 24);
        }
        lut[i] = v;
    }
}

// inflateInit2's windowBits rules (reference src/inflate.c:147-184): false -> Z_STREAM_ERROR
__host__ __device__ inline bool wbits_valid(int wb) {
    int d64 = 0;
    if (wb < 0) {
        if (wb < -16) return false;
        d64 = wb == -16;
        wb = -wb;
    } else if (wb < 48) {
        wb &= 15;
    }
    return !(wb && (wb < 8 || wb > (d64 ? 16 : 15)));
}

// what the leader keeps about its member besides the decoder state
struct Mem {
    uint32_t idx, kind, adler0;
    uint32_t hdr_tin, hdr_bits;
    bool hdr_stop;
};

// member ended (for this kernel): write the result and what k_check needs
ZSB_DEV void write_result(const Dec &d, const Mem &mm, int ev, zsb_result *res, Aux *aux) {
    zsb_result r;
    Aux a;
    a.kind = mm.kind; a.trailer_pos = 0; a.need_trailer = 0; a.init_check = mm.adler0;
    a.nseq = 0; a.pad[0] = a.pad[1] = a.pad[2] = 0;
    r.msg_id = d.msg;
    r.total_out = d.pos;
    r.check = mm.adler0;
    const BitIn &b = d.b;
    uint32_t tin = b.next - (b.nbits >> 3), bits = b.nbits & 7u;
    int extra128 = 0;
    switch (ev) {
    case EV_INITERR:
        r.ret = -2; tin = 0; bits = 0; r.total_out = 0; r.msg_id = 0; r.check = 0; a.kind = 0;
        break;
    case EV_NEEDDICT:                          // src/inflate.c:817-827: returns before the exit bookkeeping
        r.ret = 2; tin = 0; bits = 0; a.kind = 0;
        break;
    case EV_DONE:
        if (mm.kind == 0) r.ret = 1;
        else { r.ret = 0; a.need_trailer = 1; a.trailer_pos = tin; }
        bits = 0;
        break;
    case EV_ERR:
        r.ret = -3;
        if (mm.hdr_stop) { tin = mm.hdr_tin; bits = mm.hdr_bits; }
        else {
            uint64_t bp = bi_bitpos(b);
            tin = (uint32_t)((bp + d.extent + 7) >> 3);
            bits = (uint32_t)(8ull * tin - bp);
        }
        break;
    case EV_STARVE:
        r.ret = -5;
        if (mm.hdr_stop) { tin = mm.hdr_tin; bits = mm.hdr_bits; }
        else { tin = b.len; bits = b.nbits; extra128 = d.at_type ? 128 : 0; }
        break;
    case EV_BLOCKSTOP: {                       // src/inflate.c:832,854-857,883,1032: Z_OK with the bits held back
        r.ret = 0;
        const uint64_t bp = bi_bitpos(b);
        tin = (uint32_t)((bp + 7) >> 3);
        bits = (uint32_t)(8ull * tin - bp);
        extra128 = d.at_type ? 128 : 256;      // mode TYPE, or LEN_ / COPY_ (src/inflate.c:1269-1271)
        break;
    }
    default:                                   // EV_OUTFULL
        r.ret = -5;
        break;
    }
    r.total_in = tin;
    r.data_type = (ev == EV_INITERR || ev == EV_NEEDDICT) ? 0 : (int)(bits + (d.last ? 64 : 0) + extra128);
    *res = r;
    *aux = a;
}

#ifndef ZSB_FETCH_SLOT
#define ZSB_FETCH_SLOT(counter) atomicAdd((counter), 1u)
#endif

enum : int { ST_FETCH = 0, ST_BLOCK = 1, ST_SYM = 2, ST_IDLE = 3 };

// ------------------------------------------------------------------------------------------------
// The decoder as a warp-synchronous state machine.  A warp holds 32/G groups of G lanes; every group
// owns one member at a time and carries a state (fetch a member / read a block header / decode
// symbols / no more work).  The warp walks through the phases together so that the lanes stay in
// lock step where it pays: all leaders decode one symbol per step into their token queues
// (reconverging after every step), then all groups execute their queued tokens.  Member start-up
// and block headers are done per group between bursts.
// ------------------------------------------------------------------------------------------------
template <int G>
__device__ __forceinline__ void warp_inflate(const zsb_member *members, zsb_result *results, Aux *aux,
                                             const uint32_t *order, uint32_t n, uint32_t *counter, uint16_t *T,
                                             const uint32_t *lut, const uint16_t *fixed_tab, int sub,
                                             unsigned gmask) {
    Dec d;
    Mem mm;
    d.pos = 0; d.mlen = 0; d.mdist = 0; d.mtrunc = 0; d.out = nullptr; d.b.p = nullptr;
    mm.idx = 0;
    int st = ST_FETCH;
    const uint8_t *in = nullptr;
    uint8_t *out = nullptr;
    uint32_t *q = (uint32_t *)(T + T_Q);
    const bool leader = sub == 0;

    for (;;) {
        int ev = EV_NONE;
        // ---------------- a new member: descriptor, wrapper header ----------------
        if (st == ST_FETCH) {
            uint32_t slot = 0;
            if (leader) slot = ZSB_FETCH_SLOT(counter);
            slot = grp_bcast<G>(slot, gmask);
            if (slot >= n) {
                st = ST_IDLE;
            } else {
                const uint32_t idx = order ? order[slot] : slot;
                const zsb_member m = members[idx];
                in = (const uint8_t *)m.in;
                out = (uint8_t *)m.out;
                if (leader) {
                    mm.idx = idx; mm.kind = 0; mm.adler0 = 0; mm.hdr_stop = false; mm.hdr_tin = 0; mm.hdr_bits = 0;
                    d.b.p = in;
                    d.b.len = (uint32_t)m.in_len;
                    d.b.next = 0; d.b.nbits = 0; d.b.hold = 0;
                    d.out = out;
                    d.pos = 0;
                    d.cap = (uint32_t)m.out_cap;
                    d.d64 = m.window_bits == -16;
                    d.last = 0; d.at_type = 0; d.msg = 0; d.extent = 0; d.mlen = 0; d.mdist = 0; d.mtrunc = 0;
                    if (m.flags & ZSB_M_STORED) {      // copied (and checked) by k_check; nothing to decode
                        Aux a;
                        a.kind = 0; a.trailer_pos = 0; a.need_trailer = 0; a.init_check = 0; a.nseq = 0;
                        a.pad[0] = a.pad[1] = a.pad[2] = 0;
                        aux[idx] = a;
                        ev = EV_INITERR;               // any event: the group fetches its next member
                    } else if (!wbits_valid(m.window_bits) || (in == nullptr && m.in_len != 0) ||
                        (out == nullptr && m.out_cap != 0)) {
                        ev = EV_INITERR;
                    } else {
                        HeadOut h = parse_wrapper(in, d.b.len, m.window_bits);
                        mm.kind = h.kind;
                        mm.adler0 = h.adler;
                        ev = h.ev;
                        if (ev == EV_NONE) {
                            bi_seek(d.b, h.pos);
                            d.at_type = mm.kind != 0;
                        } else {
                            mm.hdr_stop = true;
                            d.msg = h.msg;
                            mm.hdr_tin = ev == EV_STARVE ? d.b.len : h.pulled;
                            mm.hdr_bits = h.bits;
                        }
                    }
                    if (ev != EV_NONE && !(m.flags & ZSB_M_STORED)) write_result(d, mm, ev, results + idx, aux + idx);
                }
                ev = (int)grp_bcast<G>((uint32_t)ev, gmask);
                st = ev == EV_NONE ? ST_BLOCK : ST_FETCH;
                ev = EV_NONE;
            }
        }
        // ---------------- block header (one block per turn) ----------------
        if (st == ST_BLOCK) {
            uint32_t btype = 0, slen = 0, spos = 0;
            if (leader) {
                BitIn &b = d.b;
                if (d.last) {
                    ev = EV_DONE;
                } else if (!bi_need(b, 3)) {
                    ev = EV_STARVE;
                } else {
                    d.last = bi_peek(b, 1);
                    bi_drop(b, 1);
                    btype = bi_peek(b, 2);
                    bi_drop(b, 2);
                    d.at_type = 0;
                    if (btype == 3) {
                        d.msg = M_BTYPE;
                        ev = EV_ERR;
                    } else if (btype == 2) {
                        ev = dynamic_header(d, T, lut);
                    } else if (btype == 0) {
                        bi_drop(b, b.nbits & 7);
                        uint32_t bytepos = b.next - (b.nbits >> 3);
                        if (b.len - bytepos < 4) {
                            b.nbits = 8 * (b.len - bytepos);
                            b.next = b.len;
                            ev = EV_STARVE;
                        } else {
                            uint32_t v = le32(in + bytepos);
                            if ((v & 0xffffu) != ((v >> 16) ^ 0xffffu)) {
                                d.msg = M_STORED;
                                d.extent = 32;
                                ev = EV_ERR;
                            } else {
                                spos = bytepos + 4;
                                slen = v & 0xffffu;
                                uint32_t have = b.len - spos, left = d.cap - d.pos;
                                if (slen > have) slen = have, btype |= 4;
                                if (slen > left) slen = left, btype |= 4;
                            }
                        }
                    }
                }
            }
            uint32_t w0 = grp_bcast<G>((uint32_t)ev | (btype << 8), gmask);
            ev = (int)(w0 & 0xffu);
            btype = w0 >> 8;
            if (ev != EV_NONE) {
                if (leader) write_result(d, mm, ev, results + mm.idx, aux + mm.idx);
                st = ST_FETCH;
            } else if ((btype & 3) == 0) {            // stored block (reference src/inflate.c:870-904)
                slen = grp_bcast<G>(slen, gmask);
                spos = grp_bcast<G>(spos, gmask);
                uint32_t pos = grp_bcast<G>(d.pos, gmask);
                grp_copy_stored<G>(out + pos, in + spos, slen, sub);
                grp_sync<G>(gmask);
                if (leader) {
                    d.pos += slen;
                    bi_seek(d.b, spos + slen);
                    d.at_type = 1;
                    if (btype & 4) {                  // input or output ran out inside the block
                        d.b.next = spos + slen;
                        d.b.nbits = 0;
                        d.at_type = 0;
                        write_result(d, mm, EV_OUTFULL, results + mm.idx, aux + mm.idx);
                    }
                }
                if (btype & 4) st = ST_FETCH;
            } else {
                if (btype == 1) {                     // fixed code: copy the context's prebuilt tables
                    uint32_t *t32 = (uint32_t *)T;
                    const uint32_t *f32 = (const uint32_t *)fixed_tab;
                    for (int i = sub; i < 256; i += G) t32[i] = f32[i];
                    for (int i = T_D / 2 + sub; i < T_DL / 2; i += G) t32[i] = f32[i];
                    grp_sync<G>(gmask);
                }
                st = ST_SYM;
            }
            ev = EV_NONE;
        }
        __syncwarp();
        // ---------------- decode burst: every leader, one symbol per step, in lock step ----------------
        uint32_t nt = 0, pos0 = d.pos;
        {
            const bool dec = leader && st == ST_SYM;
#pragma unroll 1
            for (int it = 0; it < QCAP; it++) {
                const bool go = dec && ev == EV_NONE;
                if (!__any_sync(0xffffffffu, go)) break;
                if (go) ev = decode_step(d, T, lut, q, nt);
                __syncwarp();
            }
        }
        // ---------------- execute the queued tokens ----------------
        if (G > 1) {
            uint32_t w = grp_bcast<G>((uint32_t)ev | (nt << 8), gmask);
            ev = (int)(w & 0xffu);
            nt = w >> 8;
            pos0 = grp_bcast<G>(pos0, gmask);
        }
        grp_sync<G>(gmask);                   // queue and earlier output visible to the group
        if (nt) run_tokens<G>(out, pos0, q, nt, sub, gmask);
        if (ev == EV_MATCH) {                 // match longer than a token holds (Deflate64): all lanes copy
            grp_sync<G>(gmask);
            uint32_t nn = grp_bcast<G>(d.mlen, gmask);
            uint32_t dist = grp_bcast<G>(d.mdist, gmask);
            uint32_t pos = grp_bcast<G>(d.pos, gmask);
            uint32_t trunc = grp_bcast<G>(d.mtrunc, gmask);
            grp_copy_match<G>(out, pos, dist, nn, sub);
            if (leader) d.pos += nn;
            ev = trunc ? EV_OUTFULL : EV_NONE;
        }
        grp_sync<G>(gmask);
        // ---------------- events ----------------
        if (st == ST_SYM && ev != EV_NONE) {
            if (ev == EV_EOB) {
                st = ST_BLOCK;
                if (leader) d.at_type = 1;
            } else {
                if (leader) write_result(d, mm, ev, results + mm.idx, aux + mm.idx);
                st = ST_FETCH;
            }
        }
        if (__all_sync(0xffffffffu, st == ST_IDLE)) break;
    }
}

}  // namespace zsb
