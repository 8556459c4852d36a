// zsb_lane.cuh — the lane-per-member decoder ("phase A") of the batched inflate engine, sm_100a.
//
// Why: the decode loop is bound by instruction issue, not by HBM.  With a group of lanes per member
// only the leader lane decodes, so a warp instruction yields at most 32/G symbols.  Here every lane of
// a warp owns one member and decodes one symbol per loop trip, so a warp instruction works on 32
// symbols.  To keep the 32 lanes in lock step the loop never touches the output: symbols are turned
// into 8-byte LZ SEQUENCES (up to three literals + one match) in a per-member scratch array, and
// k_lz (zsb_lz.cuh, one warp per member, coalesced) turns the sequences into bytes afterwards.
//
// Shared memory is what limits the number of resident members, so the per-member tables are compact
// (976 bytes): an 8-bit literal/length root of 16-bit entries, a 6-bit distance root of 8-bit entries,
// and canonical-code side arrays (left-aligned limits + symbols in code order) for longer codes.
// Per-lane work that is rare and long (block headers, the last bytes of a member, wrapper headers) is
// done in a service section between runs of the lock-step symbol loop; Huffman tables are built by
// the whole warp together, one member after the other.
//
// Behavioural contract: identical to zsb_device.cuh (reference src/inflate.c:597-1275,
// src/inffast.c:50-304, src/inftrees.c:33-321) — same order of checks, same consumed-input
// accounting at every stop, same accept/reject decisions of the table builder.
#pragma once
#include "zsb_device.cuh"

namespace zsb {
namespace lane {

// ---- per-member table slot (bytes) ----
// Canonical-code tables, the same three arrays for both codes: for every code length l = 1..15 the
// left-aligned (15 bit) limit lim[l] = (first code of length l + number of codes of length l) <<
// (15 - l) — non-decreasing in l, so the length of the code at the head of the bit buffer is the
// first l whose limit exceeds the next 15 bits — a bias that turns those bits into an index into the
// symbols in code order, and those symbols (one byte each; for the literal/length code the ninth bit
// comes from lt256[l], the index of the first symbol >= 256 of that length).
// 470 bytes per member: 14 warps x 32 members per SM, enough to hold a whole 65,536-member batch in
// flight on 148 SMs.
constexpr int S_LLIM = 0;                 // u16[15]  literal/length limits, lengths 1..15
constexpr int S_LBASE = 30;               // u16[15]  index bias per length
constexpr int S_LT256 = 60;               // u16[15]  first index holding a symbol >= 256, per length
constexpr int S_LS = 90;                  // u8[288]  low byte of the symbols in code order (the fixed code has 288)
constexpr int S_D = 378;                  // distance code: same shape without lt256
constexpr int S_DLIM = S_D;               // u16[15]
constexpr int S_DBASE = S_D + 30;         // u16[15]
constexpr int S_DS = S_D + 60;            // u8[32]
constexpr int SLOT = 470;                 // bytes per member
// while a block header is read the slot holds (the tables are built over it afterwards):
constexpr int H_LENS = 0;                 // u8[320]   code lengths of the litlen + distance codes
constexpr int H_CL = 320;                 // u8[128]   code-length code root (7 bits): sym<<3 | len
constexpr int H_CLL = 448;                // u8[19]    code-length code lengths
constexpr int H_CNT = 0;                  // u16[16]   census of the code-length code (before H_LENS is filled)
constexpr int H_NXT = 32;                 // u16[16]

// ---- per-warp scratch (bytes) for the cooperative table build ----
constexpr int W_SORT = 0;                 // u16[288]  symbols in code order
constexpr int W_CNT = 576;                // u32[16]   codes per length
constexpr int W_LIT = 640;                // u32[16]   codes per length whose symbol is < 256
constexpr int W_OFF = 704;                // u32[16]   running sort offset per length
constexpr int W_DL = 768;                 // u8[32]    distance code lengths
constexpr int WSCR = 992;

constexpr int WARPS = 14;                 // warps per CTA, one CTA per SM: 448 members in flight per SM
constexpr size_t smem_for(size_t warps) { return LUT_BYTES + warps * WSCR + warps * 32 * SLOT + warps * 32 * 16; }
constexpr size_t SMEM = smem_for(WARPS);
constexpr int PIPE_WARPS = 2;             // CTA size of the chunks of a pipelined host batch: 6 CTAs share an SM

// ---- LZ sequence (8 bytes): lo = lit0 | lit1<<8 | lit2<<16 | flags, hi = payload ----
constexpr uint32_t Q_NLIT_SHIFT = 24;     // bits 24,25: number of literals (0..3)
constexpr uint32_t Q_MATCH = 1u << 26;    // hi = (len-3)<<16 | (dist-1)
constexpr uint32_t Q_RAW = 1u << 27;      // stored bytes: hi = offset in the member's input, lo[15:0] = count
constexpr uint32_t Q_ABSLEN = 1u << 28;   // with Q_MATCH: hi>>16 is the length itself (capacity-cut 1 or 2)

struct seq_t {
    uint32_t lo, hi;
};

}  // namespace lane
namespace lz {           // zsb_lz.cuh; k_big resolves its queue with it
template <int W>
ZSB_DEV uint32_t resolve_range(uint8_t *out, const uint8_t *in, const lane::seq_t *seq, uint32_t nseq, int lane, uint32_t base);
}
namespace lane {

// ---- warp primitives, W = 32 on the device, 1 in the host emulation of the test tier ----
template <int W> struct Wp {
    static ZSB_DEV unsigned ballot(bool p) { return __ballot_sync(0xffffffffu, p); }
    static ZSB_DEV bool any(bool p) { return __any_sync(0xffffffffu, p); }
    static ZSB_DEV uint32_t shfl(uint32_t v, int src) { return __shfl_sync(0xffffffffu, v, src); }
    static ZSB_DEV uint32_t shfl_up(uint32_t v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
    static ZSB_DEV unsigned match(uint32_t v) { return __match_any_sync(0xffffffffu, v); }
    static ZSB_DEV unsigned reduce_or(unsigned v) { return __reduce_or_sync(0xffffffffu, v); }
    static ZSB_DEV uint32_t shfl_down(uint32_t v, int d) { return __shfl_down_sync(0xffffffffu, v, d); }
    static ZSB_DEV void sync() { __syncwarp(); }
};
template <> struct Wp<1> {
    static ZSB_DEV unsigned ballot(bool p) { return p ? 1u : 0u; }
    static ZSB_DEV bool any(bool p) { return p; }
    static ZSB_DEV uint32_t shfl(uint32_t v, int) { return v; }
    static ZSB_DEV uint32_t shfl_up(uint32_t v, int) { return v; }
    static ZSB_DEV unsigned match(uint32_t) { return 1u; }
    static ZSB_DEV unsigned reduce_or(unsigned v) { return v; }
    static ZSB_DEV uint32_t shfl_down(uint32_t v, int) { return v; }
    static ZSB_DEV void sync() {}
};

// per-lane decoder state
struct Lane {
    Dec d;               // bit input, output accounting, stop bookkeeping (zsb_device.cuh)
    Mem mm;
    uint32_t *ring;      // 4 words of shared memory (RING_STRIDE apart): the input words at d.b.next, +4, +8 arrive here
                         // through cp.async, so a refill never waits on a load scoreboard
    seq_t *seq;          // where this member's next sequence goes
    seq_t *seq0;         // start of its sequence array
    seq_t *seq_lim;      // k_lane_seg: end of the arena range the segment may fill
    uint32_t lw, nl;     // literals waiting for their sequence
    uint32_t hist;       // valid bytes of history before out[0] (0 for one-shot members)
    uint32_t nlen, ndist;
    uint32_t mode;       // fast_step: 0 = the next code is a literal/length code, 1 = a distance code
    uint32_t pl;         // fast_step: length of the match whose distance code comes next (mode 1)
    uint32_t ring_sa;    // shared-window address of `ring` (device only)
    uint32_t pre;        // ZSB_V_PREFETCH: the input word at d.b.next, loaded one refill ahead
    seq_t *q0, *qend;    // k_big: the warp's sequence queue in shared memory (NULL in k_lane); L.seq runs inside it
    uint32_t qbase;      // k_big: output position of the queue's first sequence
};

ZSB_DEV void emit(Lane &L, uint32_t lo, uint32_t hi) {
    seq_t s;
    s.lo = lo;
    s.hi = hi;
    *L.seq++ = s;
}
ZSB_DEV void flush_literals(Lane &L) {
    if (L.nl) emit(L, L.lw | (L.nl << Q_NLIT_SHIFT), 0);
    L.lw = 0;
    L.nl = 0;
}
ZSB_DEV void push_literal(Lane &L, uint32_t v) {
    if (L.nl == 3) flush_literals(L);
    L.lw |= v << (8 * L.nl);
    L.nl++;
    L.d.pos++;
}
// a match of n bytes (already cut to the capacity) at the current position
ZSB_DEV void push_match(Lane &L, uint32_t n, uint32_t dist) {
    uint32_t lo = L.lw | (L.nl << Q_NLIT_SHIFT) | Q_MATCH;
    uint32_t hi = dist - 1u;
    if (n < 3) {
        lo |= Q_ABSLEN;
        hi |= n << 16;
    } else {
        hi |= (n - 3u) << 16;
    }
    emit(L, lo, hi);
    L.lw = 0;
    L.nl = 0;
    L.d.pos += n;
}
ZSB_DEV void push_raw(Lane &L, uint32_t in_off, uint32_t n) {
    flush_literals(L);
    if (n) emit(L, Q_RAW | n, in_off);
    L.d.pos += n;
}

// ---- asynchronous input prefetch: global -> the lane's ring in shared memory ----
#ifndef ZSB_CP4
#define ZSB_CP4(sdst, gsrc)                                                                             \
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(sdst)), "l"(gsrc))
#define ZSB_CP_COMMIT() asm volatile("cp.async.commit_group;")
#define ZSB_CP_WAIT(n) asm volatile("cp.async.wait_group %0;" ::"n"(n) : "memory")
#endif
constexpr int RING_WORDS = 4;
// word w of a lane's ring is ring[w * RING_STRIDE]: on the device the 32 rings of a warp are interleaved, so the
// lanes hit 32 different banks whatever word each one is at (4 consecutive words per lane were a 4-way conflict)
#if defined(__CUDA_ARCH__)
constexpr int RING_STRIDE = 32;
#else
constexpr int RING_STRIDE = 1;
#endif

// start the copy of the input word at byte offset `off` (one group per word, empty past the end)
ZSB_DEV void fetch_word(const Lane &L, uint32_t off) {
    const BitIn &b = L.d.b;
    if (off < b.len) ZSB_CP4(&L.ring[((off >> 2) & 3u) * RING_STRIDE], b.p + off);
    ZSB_CP_COMMIT();
}
ZSB_DEV void prime(Lane &L) {
    const BitIn &b = L.d.b;
#if ZSB_V_PREFETCH
    L.pre = 0;
    if ((((uintptr_t)(b.p + b.next)) & 3) == 0 && b.next + 4 <= b.len) L.pre = __ldg((const uint32_t *)(b.p + b.next));
    return;
#endif
    ZSB_CP_WAIT(0);
    if ((((uintptr_t)(b.p + b.next)) & 3) != 0) return;       // only at the very end of a member
    fetch_word(L, b.next);
    fetch_word(L, b.next + 4);
    fetch_word(L, b.next + 8);
}

// ------------------------------------------------------------------------------------------------
// Code lookups on the canonical tables.  `h32` holds the next input bits (zero padded past the end of
// the input); nb receives the code length.  A bit pattern that is no code (possible only with the
// incomplete sets the reference accepts: a single one-bit code, or no distance code at all,
// src/inftrees.c:131-153) reads as the reference's filler entry: invalid, one bit (`:301-317`).
// ------------------------------------------------------------------------------------------------
ZSB_DEV uint32_t sym_entry(uint32_t sym) {        // symbol -> entry without the length field
    if (sym < 256) return K_LIT | (sym << 6);
    if (sym == 256) return E_EOB;
    if (sym < 286) return K_LEN | ((sym - 257) << 6);
    return E_INVALID;
}
// number of lengths whose limit is at or below c15 (0..15) = code length - 1; 15 = no code
#ifndef ZSB_V_SPLIT
#define ZSB_V_SPLIT 0
#endif
#ifndef ZSB_V_PREFETCH
#define ZSB_V_PREFETCH 0
#endif
ZSB_DEV uint32_t canon_len(const uint16_t *lim, uint32_t c15) {
#if ZSB_V_SPLIT
    // the three splitters of the first two levels are independent loads (their addresses do not depend on the
    // code): two dependent shared-memory round trips instead of four
    const uint32_t l7 = lim[7], l3 = lim[3], l11 = lim[11];
    uint32_t k = c15 >= l7 ? 8u : 0u;
    k += c15 >= (k ? l11 : l3) ? 4u : 0u;
#else
    uint32_t k = c15 >= lim[7] ? 8u : 0u;
    k += c15 >= lim[k + 3] ? 4u : 0u;
#endif
    k += c15 >= lim[k + 1] ? 2u : 0u;
    k += c15 >= lim[k] ? 1u : 0u;
    return k;
}
ZSB_DEV uint32_t lookup_litlen(const uint8_t *T, uint32_t h32, uint32_t &nb) {
    const uint32_t c15 = __brev(h32) >> 17;
    const uint32_t k = canon_len((const uint16_t *)(T + S_LLIM), c15);
    if (k == 15) {
        nb = 1;
        return E_INVALID;
    }
    const uint32_t idx = (((const uint16_t *)(T + S_LBASE))[k] + (c15 >> (14 - k))) & 0x1ffu;
    uint32_t sym = T[S_LS + idx];
    if (idx >= ((const uint16_t *)(T + S_LT256))[k]) sym += 256;
    nb = k + 1;
    return sym_entry(sym);
}
// distance code: returns the symbol, 32 for an invalid code
ZSB_DEV uint32_t lookup_dist(const uint8_t *T, uint32_t h32, uint32_t &nb) {
    const uint32_t c15 = __brev(h32) >> 17;
    const uint32_t k = canon_len((const uint16_t *)(T + S_DLIM), c15);
    if (k == 15) {
        nb = 1;
        return 32;
    }
    const uint32_t idx = (((const uint16_t *)(T + S_DBASE))[k] + (c15 >> (14 - k))) & 31u;
    nb = k + 1;
    return T[S_DS + idx];
}

// ------------------------------------------------------------------------------------------------
// One symbol with every input step checked (the last bytes of a member): stops exactly where the
// reference would wait for input (src/inflate.c:1046-1142).  Returns EV_LIT (byte in d.mdist),
// EV_MATCH (d.mlen, d.mdist), EV_EOB, EV_ERR or EV_STARVE.
// ------------------------------------------------------------------------------------------------
// have_len: the length code and its extra bits have been consumed already (d.mlen is set): resume at
// the distance code, as the reference resumes in mode DIST (src/inflate.c:1097).
ZSB_DEV int careful_symbol(Dec &d, const uint8_t *T, const uint32_t *lut, bool have_len = false) {
    BitIn &b = d.b;
    uint32_t nb, xb;
    if (!have_len) {
        bi_fill(b);
        uint32_t e = lookup_litlen(T, (uint32_t)b.hold, nb);
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
        xb = lw >> 16;
        if (!bi_need(b, xb)) return EV_STARVE;
        d.mlen = (lw & 0xffffu) + (xb ? bi_peek(b, xb) : 0u);
        bi_drop(b, xb);
    }
    bi_fill(b);
    uint32_t s = lookup_dist(T, (uint32_t)b.hold, nb);
    if (nb > b.nbits) return EV_STARVE;
    bi_drop(b, nb);
    if (s >= 32) {
        d.msg = M_DISTCODE;
        return EV_ERR;
    }
    uint32_t dw = lut[LUT_DIST + d.d64 * 32 + s];
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

// capacity / distance checks of a decoded match, then the sequence (src/inflate.c:1143-1153)
ZSB_DEV int accept_match(Lane &L) {
    Dec &d = L.d;
    if (d.pos >= d.cap) return EV_OUTFULL;
    if ((uint64_t)d.mdist > (uint64_t)d.pos + L.hist) {        // 64-bit: pos + hist may pass 2^32 in a member near the 4 GiB limit
        d.msg = M_TOOFAR;
        return EV_ERR;
    }
    uint32_t n = d.cap - d.pos;
    const bool cut = n < d.mlen;
    if (!cut) n = d.mlen;
    push_match(L, n, d.mdist);
    return cut ? EV_OUTFULL : EV_NONE;
}

// ------------------------------------------------------------------------------------------------
// The lock-step loop of k_lane: ONE Huffman code per trip and lane.  A match takes two trips (the
// length code, then the distance code), so that a trip's dependent chain is one root lookup (+ one
// side-array lookup) + one base/extra-bits lookup for every lane, whatever it decodes, and literal
// and match lanes run the same instructions.  Everything is straight-line and predicated; the only
// branch is the search for codes the roots do not resolve (rare).
// Returns false without consuming the code when the lane has to leave the loop (end of block, invalid
// code, distance too far back, capacity reached): the checked path takes over at that code, with
// L.mode telling it whether the length has been consumed already.
// ------------------------------------------------------------------------------------------------
ZSB_DEV uint32_t ring_load(const Lane &L, uint32_t byteoff) {          // byteoff = 0, 4, 8 or 12
#if defined(__CUDA_ARCH__)
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(L.ring_sa + byteoff * RING_STRIDE) : "memory");
    return v;
#else
    return L.ring[byteoff >> 2];
#endif
}
// if p: start the copy of the input word at byte offset off (nothing past the end) and close a group
ZSB_DEV void fetch_word_if(const Lane &L, bool p, uint32_t off) {
    const BitIn &b = L.d.b;
#if defined(__CUDA_ARCH__)
    const uint32_t go = (p && off < b.len) ? 1u : 0u, grp = p ? 1u : 0u;
    asm volatile(
        "{\n\t.reg .pred q, r;\n\tsetp.ne.u32 q, %0, 0;\n\tsetp.ne.u32 r, %1, 0;\n\t"
        "@q cp.async.ca.shared.global [%2], [%3], 4;\n\t@r cp.async.commit_group;\n\t}" ::"r"(go),
        "r"(grp), "r"(L.ring_sa + (off & 12u) * RING_STRIDE), "l"(b.p + off)
        : "memory");
#else
    if (p && off < b.len) L.ring[(off >> 2) & 3u] = *(const uint32_t *)(b.p + off);
#endif
}
ZSB_DEV void store_seq_if(seq_t *at, bool p, uint32_t lo, uint32_t hi) {
#if defined(__CUDA_ARCH__)
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %0, 0;\n\t@q st.global.v2.u32 [%1], {%2, %3};\n\t}" ::"r"(
                     (uint32_t)p),
                 "l"(at), "r"(lo), "r"(hi)
                 : "memory");
#else
    if (p) {
        at->lo = lo;
        at->hi = hi;
    }
#endif
}
ZSB_DEV uint32_t sel(uint32_t m, uint32_t a, uint32_t b) { return (a & m) | (b & ~m); }   // m ? a : b, m = 0 / ~0

ZSB_DEV bool fast_step(Lane &L, const uint8_t *T, const uint32_t *lut) {
    Dec &d = L.d;
    BitIn &b = d.b;
    // ---- refill: 32 more bits when fewer than 32 are held; the word three ahead is requested ----
    const bool need = b.nbits < 32;
#if ZSB_V_PREFETCH
    // the word at b.next was requested at the previous refill; the loop's entry condition (8 more input bytes)
    // makes the next word readable
    const uint32_t rw = L.pre;
    b.hold |= need ? (uint64_t)rw << b.nbits : 0ull;
    b.nbits += need ? 32u : 0u;
    b.next += need ? 4u : 0u;
    if (need) L.pre = __ldg((const uint32_t *)(b.p + b.next));
#else
    ZSB_CP_WAIT(2);
    const uint32_t rw = ring_load(L, b.next & 12u);
    fetch_word_if(L, need, b.next + 12);
    b.hold |= need ? (uint64_t)rw << b.nbits : 0ull;
    b.nbits += need ? 32u : 0u;
    b.next += need ? 4u : 0u;
#endif
    // ---- the code at the head of the buffer, in the lane's current code (mask arithmetic, no branch) ----
    const uint32_t w = (uint32_t)b.hold;
    const uint32_t m = 0u - L.mode;                                // ~0 in distance mode
    const uint32_t c16 = __brev(w) >> 16, c15 = c16 >> 1;
    const uint8_t *tb = T + (m & (uint32_t)S_D);
    const uint32_t k = canon_len((const uint16_t *)tb, c15);       // code length - 1; 15 = no such code
    const uint32_t bias = ((const uint16_t *)(tb + S_LBASE))[k];
    const uint32_t lt = ((const uint16_t *)(T + S_LT256))[k];
    const uint32_t idx = (bias + (c16 >> (15 - k))) & 0x1ffu;
    const uint32_t sb = tb[sel(m, (uint32_t)(S_DS - S_D), (uint32_t)S_LS) + idx];
    const uint32_t sym = sb + ((~m & (idx >= lt ? 256u : 0u)));
    const uint32_t nb = k + 1;
    const bool islit = !m && sym < 256;
    const uint32_t v = sel(m, sym, sym - 257u);                    // index of the base / extra-bits entry
    const bool hasx = k < 15 && (m || v < 29u);                    // a length or a distance code
    const uint32_t lw = lut[sel(m, LUT_DIST, LUT_LEN) + d.d64 * 32 + (v & 31u)];
    const uint32_t xb = hasx ? (lw >> 16) & 31u : 0u;
    const uint32_t val = (lw & 0xffffu) + ((w >> nb) & ((1u << xb) - 1u));
    const uint32_t used = nb + xb;
    // ---- may this code be committed here? ----
    const uint32_t okD = (hasx && !(lw >> 31) && (uint64_t)val <= (uint64_t)d.pos + L.hist && L.pl <= d.cap - d.pos) ? 1u : 0u;
    const uint32_t okL = (hasx || (k < 15 && islit && d.pos < d.cap)) ? 1u : 0u;
    if (!sel(m, okD, okL)) return false;
    // ---- commit ----
    b.hold >>= used;
    b.nbits -= used;
    const bool put = m != 0 || (islit && L.nl == 3);
    store_seq_if(L.seq, put, L.lw | (L.nl << Q_NLIT_SHIFT) | (m & Q_MATCH), m & ((val - 1u) | ((L.pl - 3u) << 16)));
    L.seq += put ? 1 : 0;
    L.lw = put ? 0u : L.lw;
    L.nl = put ? 0u : L.nl;
    L.lw |= islit ? sym << (8 * L.nl) : 0u;
    L.nl += islit ? 1u : 0u;
    d.pos += sel(m, L.pl, islit ? 1u : 0u);
    L.pl = val;                   // meaningful after a length code
    L.mode = (~m & (hasx ? 1u : 0u));
    return true;
}

// ------------------------------------------------------------------------------------------------
// The large-member decoder (k_big): ONE member per warp, decoded by lane 0 with direct-indexed tables.
// A lane of k_lane decodes its member at a fixed pace (about 200 ns per byte of input + output at full
// occupancy), so a member far larger than the rest of its batch would hold the batch up.  Such members
// get a warp of their own and 3 KB of lookup tables beside the canonical ones: 9 index bits for the
// literal/length code, 8 for the distance code, one 32-bit entry = code length | extra-bit count << 4 |
// kind << 9 | literal or base << 16.  Codes longer than the index and everything that ends a run of
// plain symbols (end of block, invalid code, distance too far, capacity, the last bytes of the input) are
// left to the same checked path as in k_lane, so results are identical by construction.
// ------------------------------------------------------------------------------------------------
constexpr int BIG_LBITS = 9, BIG_DBITS = 8;
constexpr int BIG_LUT_BYTES = ((1 << BIG_LBITS) + (1 << BIG_DBITS)) * 4;
// The warp's sequence queue: lane 0 decodes into it, the whole warp resolves it (lz::resolve_range) whenever
// it is nearly full or a block / the member ends: no sequence arena and no k_lz pass for these members.
constexpr int BIG_QCAP = 512, BIG_QMARGIN = 128;    // margin: what the checked path may emit before the next flush
constexpr uint32_t BK_LIT = 0, BK_BASE = 1, BK_EOB = 2, BK_INVALID = 3;

// entry of the literal/length code at the head of h32 (nb = 1..15 in bits 0-3)
ZSB_DEV uint32_t big_ll_entry(const uint8_t *T, const uint32_t *lut, uint32_t d64, uint32_t h32) {
    uint32_t nb;
    const uint32_t e = lookup_litlen(T, h32, nb);
    const uint32_t kind = e & 0x30u, v = e >> 6;
    if (kind == K_LIT) return nb | (BK_LIT << 9) | (v << 16);
    if (kind == K_SPECIAL) return nb | ((v == 0 ? BK_EOB : BK_INVALID) << 9);
    const uint32_t lw = lut[LUT_LEN + d64 * 32 + v];
    return nb | ((lw >> 16) << 4) | (BK_BASE << 9) | (lw << 16);
}
ZSB_DEV uint32_t big_dd_entry(const uint8_t *T, const uint32_t *lut, uint32_t d64, uint32_t h32) {
    uint32_t nb;
    const uint32_t sym = lookup_dist(T, h32, nb);
    if (sym >= 32) return nb | (BK_INVALID << 9);
    const uint32_t dw = lut[LUT_DIST + d64 * 32 + sym];
    if (dw >> 31) return nb | (BK_INVALID << 9);
    return nb | (((dw >> 16) & 15u) << 4) | (BK_BASE << 9) | (dw << 16);
}
// the direct tables of the member whose canonical tables are at T, by the whole warp; codes longer than
// the index leave 0 (the loop then asks big_*_entry)
template <int W>
ZSB_DEV void big_build(uint32_t *LL, const uint8_t *T, const uint32_t *lut, uint32_t d64, int lane) {
    uint32_t *DD = LL + (1 << BIG_LBITS);
    for (uint32_t i = lane; i < (1u << BIG_LBITS); i += W) {
        const uint32_t e = big_ll_entry(T, lut, d64, i);
        LL[i] = (e & 15u) <= (uint32_t)BIG_LBITS ? e : 0u;
    }
    for (uint32_t i = lane; i < (1u << BIG_DBITS); i += W) {
        const uint32_t e = big_dd_entry(T, lut, d64, i);
        DD[i] = (e & 15u) <= (uint32_t)BIG_DBITS ? e : 0u;
    }
    Wp<W>::sync();
}

// plain literals and whole matches until something else comes up; leaves the bit input at the start of the
// first code it did not take
// Returns true when it stopped only because the queue has to be flushed (the lane stays in the loop).
ZSB_DEV bool big_run(Lane &L, const uint8_t *T, const uint32_t *lut, const uint32_t *LL) {
    Dec &d = L.d;
    BitIn &b = d.b;
    const uint32_t *DD = LL + (1 << BIG_LBITS);
    if (b.next + 16 > b.len) return false;
    // (b.next is word aligned while whole words are left: the invariant bi_loadword relies on)
    uint32_t pre = __ldg((const uint32_t *)(b.p + b.next));     // the next input word, requested one refill ahead
    for (;;) {
        // Codes that can be taken without looking at the queue, the end of the input or (for literals) the
        // capacity again: a code queues at most one sequence, reads at most 8 input bytes and yields at least
        // one output byte.
        if (b.next + 16 > b.len) return false;
        const long qfree = (long)(L.qend - L.seq) - BIG_QMARGIN;
        if (qfree <= 0) return true;
        uint32_t budget = (b.len - 16 - b.next) / 8 + 1;
        if ((long)budget > qfree) budget = (uint32_t)qfree;
        if (budget > d.cap - d.pos) budget = d.cap - d.pos;
        if (budget == 0) return false;                   // no room for a literal: the checked path decides
        while (budget) {
            budget--;
            if (b.nbits < 32) {
                b.hold |= (uint64_t)pre << b.nbits;
                b.nbits += 32;
                b.next += 4;
                pre = __ldg((const uint32_t *)(b.p + b.next));
            }
            uint32_t e = LL[(uint32_t)b.hold & ((1u << BIG_LBITS) - 1u)];
            if (!(e & 15u)) e = big_ll_entry(T, lut, d.d64, (uint32_t)b.hold);
            const uint32_t nb = e & 15u, kind = (e >> 9) & 3u;
            if (kind == BK_LIT) {
                b.hold >>= nb;
                b.nbits -= nb;
                push_literal(L, e >> 16);
                continue;
            }
            if (kind != BK_BASE) return false;           // end of block / invalid code: the checked path's
            const uint64_t h0 = b.hold;                  // a match is taken whole or not at all
            const uint32_t n0 = b.nbits, x0 = b.next;
            const uint32_t xb = (e >> 4) & 31u;
            const uint32_t mlen = (e >> 16) + ((uint32_t)(b.hold >> nb) & ((1u << xb) - 1u));
            b.hold >>= nb + xb;
            b.nbits -= nb + xb;
            if (b.nbits < 32) {
                b.hold |= (uint64_t)pre << b.nbits;
                b.nbits += 32;
                b.next += 4;
                pre = __ldg((const uint32_t *)(b.p + b.next));
            }
            uint32_t f = DD[(uint32_t)b.hold & ((1u << BIG_DBITS) - 1u)];
            if (!(f & 15u)) f = big_dd_entry(T, lut, d.d64, (uint32_t)b.hold);
            const uint32_t nb2 = f & 15u, xb2 = (f >> 4) & 31u;
            const uint32_t dist = (f >> 16) + ((uint32_t)(b.hold >> nb2) & ((1u << xb2) - 1u));
            if (((f >> 9) & 3u) != BK_BASE || (uint64_t)dist > (uint64_t)d.pos + L.hist || mlen > d.cap - d.pos) {
                b.hold = h0;
                b.nbits = n0;
                b.next = x0;
                return false;
            }
            b.hold >>= nb2 + xb2;
            b.nbits -= nb2 + xb2;
            push_match(L, mlen, dist);
            if (budget > d.cap - d.pos) budget = d.cap - d.pos;
        }
    }
}

// the warp executes what lane 0 has queued (all lanes call)
template <int W>
ZSB_DEV void big_flush(Lane &L, const seq_t *q, int lane) {
    const uint32_t nq = Wp<W>::shfl(lane == 0 ? (uint32_t)(L.seq - L.q0) : 0u, 0);
    if (!nq) return;
    const uint64_t o = (uint64_t)(uintptr_t)L.d.out, i = (uint64_t)(uintptr_t)L.d.b.p;
    const uint64_t ob = (uint64_t)Wp<W>::shfl((uint32_t)o, 0) | ((uint64_t)Wp<W>::shfl((uint32_t)(o >> 32), 0) << 32);
    const uint64_t ib = (uint64_t)Wp<W>::shfl((uint32_t)i, 0) | ((uint64_t)Wp<W>::shfl((uint32_t)(i >> 32), 0) << 32);
    uint32_t base = Wp<W>::shfl(L.qbase, 0);
    Wp<W>::sync();                                   // lane 0's queue stores are visible to the warp
    base = lz::resolve_range<W>((uint8_t *)(uintptr_t)ob, (const uint8_t *)(uintptr_t)ib, q, nq, lane, base);
    Wp<W>::sync();
    if (lane == 0) {
        L.qbase = base;
        L.seq = L.q0;
    }
}

// ------------------------------------------------------------------------------------------------
// Dynamic block header, per lane (reference src/inflate.c:905-1004): the counts, the code-length
// code and the code lengths of both codes, left as bytes at T[H_LENS..].  The two decoding tables
// are built afterwards by the whole warp (build_tables).  Returns EV_NONE, EV_ERR or EV_STARVE.
// ------------------------------------------------------------------------------------------------

ZSB_DEV bool build_clen8(uint8_t *T, bool d64) {
    const uint8_t *lens = T + H_CLL;
    uint16_t *cnt = (uint16_t *)(T + H_CNT), *nxt = (uint16_t *)(T + H_NXT);
    uint8_t *C = T + H_CL;
    uint32_t maxl;
    int k = code_census(lens, 19, cnt, nxt, false, maxl);
    if (k < 0) return false;
    if (k == 1) {                    // no codes: every lookup reads "length 0, one bit" (src/inftrees.c:131-139)
        if (d64) return false;
        for (int i = 0; i < 128; i++) C[i] = 1;
        return true;
    }
    for (uint32_t s = 0; s < 19; s++) {
        uint32_t l = lens[s];
        if (!l) continue;
        uint32_t c = nxt[l]++;
        uint32_t e = l | (s << 3);
        for (uint32_t i = rev_bits(c, l); i < 128; i += 1u << l) C[i] = (uint8_t)e;
    }
    return true;
}

// EARLY (the block-start search of zsb_seg.cuh, which tries positions that are mostly no headers): give up as soon as
// the lengths read so far over-subscribe the literal/length or the distance code — sets build_tables would refuse.
template <bool EARLY>
ZSB_DEV int lane_dynamic_header_t(Lane &L, uint8_t *T, const uint32_t *lut) {
    Dec &d = L.d;
    BitIn &b = d.b;
    uint8_t *lens = T + H_LENS;
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
        T[H_CLL + order[i]] = (uint8_t)v;
    }
    if (!build_clen8(T, d.d64)) {
        d.msg = M_CLENS;
        return EV_ERR;
    }
    {   // entries of the reference's table for this code (root 7, src/inftrees.c:126-142): the z_stream surface reports it
        uint32_t mn = 16, mx = 0;
        for (uint32_t i = 0; i < 19; i++) {
            const uint32_t l = T[H_CLL + i];
            if (l) {
                mn = l < mn ? l : mn;
                mx = l > mx ? l : mx;
            }
        }
        const uint32_t root = mx == 0 ? 1u : (7u < mx ? (7u > mn ? 7u : mn) : mx);
        L.ndist = 0x80000000u | (1u << root);
    }
    const uint8_t *C = T + H_CL;
    uint32_t have = 0, total = nlen + ndist;
    uint32_t kl = 0, kd = 0;                      // EARLY: sums of 2^(15 - length) over the two codes
    while (have < total) {
        if (b.nbits < 32) bi_fill(b);
        uint32_t e = C[(uint32_t)b.hold & 127u];
        uint32_t nb = e & 7u, sym = e >> 3;
        if (nb > b.nbits) return EV_STARVE;
        if (sym < 16) {
            bi_drop(b, nb);
            if (EARLY && sym) {
                uint32_t &k = have < nlen ? kl : kd;
                k += 32768u >> sym;
                if (k > 32768u) return EV_ERR;
            }
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
                d.extent = 2;         // the two repeat bits were pulled but not dropped (src/inflate.c:959-966)
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
        while (rep--) {
            if (EARLY && val) {
                uint32_t &k = have < nlen ? kl : kd;
                k += 32768u >> val;
            }
            lens[have++] = (uint8_t)val;
        }
        if (EARLY && (kl > 32768u || kd > 32768u)) return EV_ERR;
    }
    if (lens[256] == 0) {
        d.msg = M_NOEOB;
        return EV_ERR;
    }
    L.nlen = nlen;
    L.ndist = ndist;
    return EV_NONE;
}
ZSB_DEV int lane_dynamic_header(Lane &L, uint8_t *T, const uint32_t *lut) { return lane_dynamic_header_t<false>(L, T, lut); }

// ------------------------------------------------------------------------------------------------
// Table construction by the whole warp for ONE member (reference src/inftrees.c:33-321 decides what
// is accepted; the tables themselves are ours).  `lens` = n code lengths in shared memory, `scr` =
// the warp's scratch.  All lanes call with the same arguments and get the same return value.
// ------------------------------------------------------------------------------------------------
#ifndef ZSB_FETCH_ADD
#define ZSB_FETCH_ADD(p, v) atomicAdd((p), (v))
#endif
#ifndef ZSB_SMEM_ADD
#define ZSB_SMEM_ADD(p, v) atomicAdd((p), (v))
#endif
constexpr int W_FIRST = 800;              // u32[16] first canonical code per length
constexpr int W_LIM = 864;                // u32[16] (first+count) left-aligned to 15 bits
constexpr int W_OFF0 = 928;               // u32[16] index of the first symbol of each length in code order
// (entry 0 of W_LIM / W_OFF0 is no length's: after build_tables they hold the table entries the reference's
//  inflate_table() would have allocated for the literal/length and the distance code: inflateCodesUsed())

// Table space the reference's two-level builder would use with `root` index bits: the root plus one
// sub-table per root prefix that carries longer codes, sized by the longest code under the prefix
// (src/inftrees.c:222-228, 275-297).  cnt/first are per-length arrays.
ZSB_DEV uint32_t ref_table_space(const uint32_t *cnt, const uint32_t *first, uint32_t root, uint32_t maxl) {
    uint32_t used = 1u << root, last_p = 0xffffffffu, last_curr = 0;
    for (uint32_t l = root + 1; l <= maxl; l++) {
        if (!cnt[l]) continue;
        uint32_t lo = first[l] >> (l - root), hi = (first[l] + cnt[l] - 1u) >> (l - root);
        if (lo == last_p) {
            used += (1u << (l - root)) - (1u << last_curr);
            lo++;
        }
        if (hi + 1 > lo) used += (hi + 1 - lo) << (l - root);
        last_p = hi;
        last_curr = l - root;
    }
    return used;
}

// Census + code-order sort.  Returns 0 complete, 1 no codes, 2 a single 1-bit code, -1 invalid set.
template <int W>
ZSB_DEV int census_sort(const uint8_t *lens, uint32_t n, uint8_t *scr, int lane, uint32_t &maxl) {
    uint32_t *cnt = (uint32_t *)(scr + W_CNT), *lit = (uint32_t *)(scr + W_LIT), *off = (uint32_t *)(scr + W_OFF);
    uint32_t *first = (uint32_t *)(scr + W_FIRST), *lim = (uint32_t *)(scr + W_LIM), *off0 = (uint32_t *)(scr + W_OFF0);
    uint16_t *sorted = (uint16_t *)(scr + W_SORT);
    for (int i = lane; i < 48; i += W) cnt[i] = 0;          // cnt, lit, off are contiguous
    Wp<W>::sync();
    for (uint32_t s = lane; s < n; s += W) {
        uint32_t l = lens[s];
        if (l) {
            ZSB_SMEM_ADD(&cnt[l], 1u);
            if (s < 256) ZSB_SMEM_ADD(&lit[l], 1u);
        }
    }
    Wp<W>::sync();
    // every lane walks the 15 lengths (same values everywhere); lane 0 records the per-length arrays
    int left = 1, bad = 0;
    uint32_t code = 0, o = 0;
    maxl = 0;
    for (uint32_t l = 1; l <= 15; l++) {
        const uint32_t c = cnt[l];
        left = left * 2 - (int)c;            // may already be negative (no shift of a negative value)
        if (left < 0) bad = 1;
        if (c) maxl = l;
        if (lane == 0) {
            first[l] = code;
            lim[l] = (code + c) << (15 - l);
            off0[l] = o;
            off[l] = o;
        }
        code = (code + c) << 1;
        o += c;
    }
    Wp<W>::sync();
    if (bad) return -1;
    if (maxl == 0) return 1;
    int k = 0;
    if (left > 0) {
        if (maxl != 1) return -1;
        k = 2;
    }
    // symbols in code order: by length, then by symbol value
    for (uint32_t s0 = 0; s0 < n; s0 += W) {
        const uint32_t s = s0 + lane;
        const uint32_t l = s < n ? lens[s] : 0u;
        const unsigned m = Wp<W>::match(l);
        const uint32_t rank = __popc(m & ((1u << lane) - 1u));
        uint32_t at = 0;
        if (l) at = off[l] + rank;
        Wp<W>::sync();
        if (l) {
            sorted[at] = (uint16_t)s;
            if (rank == 0) off[l] += __popc(m);
        }
        Wp<W>::sync();
    }
    return k;
}

// literal/length code: lens[0..nlen).  false -> "invalid literal/lengths set"
template <int W>
ZSB_DEV bool build_litlen_tab(uint8_t *T, uint8_t *scr, const uint8_t *lens, uint32_t nlen, bool d64, int lane) {
    uint32_t maxl;
    const int k = census_sort<W>(lens, nlen, scr, lane, maxl);
    if (k < 0 || k == 1) return false;
    const uint32_t *cnt = (const uint32_t *)(scr + W_CNT), *lit = (const uint32_t *)(scr + W_LIT);
    const uint32_t *first = (const uint32_t *)(scr + W_FIRST), *lim = (const uint32_t *)(scr + W_LIM);
    const uint32_t *off0 = (const uint32_t *)(scr + W_OFF0);
    const uint16_t *sorted = (const uint16_t *)(scr + W_SORT);
    uint32_t minl = 1;
    while (minl < maxl && !cnt[minl]) minl++;
    const uint32_t root = 9u < maxl ? (9u > minl ? 9u : minl) : maxl;      // src/inftrees.c:126-142
    const uint32_t used = maxl > root ? ref_table_space(cnt, first, root, maxl) : (1u << root);
    if (used > (d64 ? 851u : 852u)) return false;
    if (lane == 0) ((uint32_t *)(scr + W_LIM))[0] = used;       // what inflateCodesUsed() would count for this code
    for (int j = lane; j < 15; j += W) {
        const uint32_t l = j + 1;
        ((uint16_t *)(T + S_LLIM))[j] = (uint16_t)lim[l];
        ((uint16_t *)(T + S_LBASE))[j] = (uint16_t)(off0[l] - first[l]);
        ((uint16_t *)(T + S_LT256))[j] = (uint16_t)(off0[l] + lit[l]);
    }
    const uint32_t total = off0[15] + cnt[15];
    for (uint32_t j = lane; j < total; j += W) T[S_LS + j] = (uint8_t)sorted[j];
    Wp<W>::sync();
    return true;
}

// distance code: lens[0..ndist).  false -> "invalid distances set"
template <int W>
ZSB_DEV bool build_dist_tab(uint8_t *T, uint8_t *scr, const uint8_t *lens, uint32_t ndist, bool d64, int lane) {
    uint32_t maxl;
    const int k = census_sort<W>(lens, ndist, scr, lane, maxl);
    if (k < 0) return false;
    if (k == 1) {                                           // no codes at all: src/inftrees.c:131-139
        if (d64) return false;
        for (int j = lane; j < 15; j += W) ((uint16_t *)(T + S_DLIM))[j] = 0;      // every pattern: invalid, one bit
        if (lane == 0) ((uint32_t *)(scr + W_OFF0))[0] = 2;
        Wp<W>::sync();
        return true;
    }
    const uint32_t *cnt = (const uint32_t *)(scr + W_CNT);
    const uint32_t *first = (const uint32_t *)(scr + W_FIRST), *lim = (const uint32_t *)(scr + W_LIM);
    const uint32_t *off0 = (const uint32_t *)(scr + W_OFF0);
    const uint16_t *sorted = (const uint16_t *)(scr + W_SORT);
    uint32_t minl = 1;
    while (minl < maxl && !cnt[minl]) minl++;
    const uint32_t root = 6u < maxl ? (6u > minl ? 6u : minl) : maxl;
    const uint32_t used = maxl > root ? ref_table_space(cnt, first, root, maxl) : (1u << root);
    if (d64 && used > 593u) return false;
    if (lane == 0) ((uint32_t *)(scr + W_OFF0))[0] = used;
    for (int j = lane; j < 15; j += W) {
        const uint32_t l = j + 1;
        ((uint16_t *)(T + S_DLIM))[j] = (uint16_t)lim[l];
        ((uint16_t *)(T + S_DBASE))[j] = (uint16_t)(off0[l] - first[l]);
    }
    const uint32_t total = off0[15] + cnt[15];
    for (uint32_t j = lane; j < total; j += W) T[S_DS + j] = (uint8_t)sorted[j];
    Wp<W>::sync();
    return true;
}

// both tables of a dynamic block from T[H_LENS..]: returns 0 or the message id
template <int W>
ZSB_DEV int build_tables(uint8_t *T, uint8_t *scr, uint32_t nlen, uint32_t ndist, bool d64, int lane) {
    uint8_t *dl = scr + W_DL;
    for (uint32_t i = lane; i < ndist; i += W) dl[i] = T[H_LENS + nlen + i];
    Wp<W>::sync();
    if (!build_litlen_tab<W>(T, scr, T + H_LENS, nlen, d64, lane)) return M_LITLENS;
    if (!build_dist_tab<W>(T, scr, dl, ndist, d64, lane)) return M_DISTS;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// The warp: 32 members in flight, one per lane.
// ------------------------------------------------------------------------------------------------
enum : int { LS_FETCH = 0, LS_BLOCK = 1, LS_SYM = 2, LS_TAIL = 3, LS_IDLE = 4, LS_BUILD = 5, LS_FIXED = 6 };

// A SEGMENT of a large member (zsb_seg.cuh): the blocks between two block starts that k_seg_find found near
// multiples of `segb` compressed bytes.  k_lane_seg decodes a segment like a member of its own (one lane, its own
// range of the sequence arena); k_seg_chain then checks that every segment ended exactly where the next one began.
constexpr uint64_t SEG_NONE = ~0ull;
enum : int { EV_SEGEND = 11, EV_SEGFULL = 12, EV_ABSENT = 13 };
struct Seg {
    uint32_t member, k, first, nseg;    // member index; ordinal; slot of the member's segment 0; its segment count
    uint64_t start_bit;                 // bit position of the block header it starts at (k = 0: 0); SEG_NONE: no start found
    uint64_t target;                    // where it has to stop: the next found start (SEG_NONE: the member's end)
    int32_t ev, msg;                    // how it ended (EV_*), message id
    uint32_t pos, next, nbits, extent, at_type, last;   // Dec / BitIn at the end (pos = bytes the segment produced)
    uint32_t nseq;                      // sequences written
    uint32_t kind, adler0, hdr_tin, hdr_bits, hdr_stop;  // k = 0: the wrapper (Mem)
    uint32_t base;                      // k_seg_chain: output position of the segment's first byte
    uint32_t live;                      // k_seg_chain: 1 = part of the verified chain (resolve it)
};

struct Batch {
    Seg *segs = nullptr;         // k_lane_seg: slots are segments (order is not used); NULL everywhere else
    uint32_t seg_q = 0;          // sequences of arena per segment
    const zsb_member *members;
    zsb_result *results;
    Aux *aux;
    const uint32_t *order;       // slot -> member index
    const uint64_t *seq_off;     // per slot: first sequence of the member's array in seq_base
    seq_t *seq_base;
    const uint32_t *hist;        // optional: valid history bytes before each member's out
    uint32_t n;
    uint32_t *counter;
    const uint8_t *fixed_tab;    // SLOT bytes: tables of the fixed code
};

ZSB_DEV void finish(Lane &L, int ev, const Batch &B) {
    flush_literals(L);
    if (B.segs) {                      // a segment: the state goes to its record, k_seg_chain writes the member's result
        Seg &g = B.segs[L.mm.idx];
        g.ev = ev; g.msg = L.d.msg; g.pos = L.d.pos; g.next = L.d.b.next; g.nbits = L.d.b.nbits; g.extent = L.d.extent;
        g.at_type = L.d.at_type; g.last = L.d.last; g.nseq = (uint32_t)(L.seq - L.seq0);
        g.kind = L.mm.kind; g.adler0 = L.mm.adler0; g.hdr_tin = L.mm.hdr_tin; g.hdr_bits = L.mm.hdr_bits; g.hdr_stop = L.mm.hdr_stop;
        return;
    }
    write_result(L.d, L.mm, ev, B.results + L.mm.idx, B.aux + L.mm.idx);
    B.aux[L.mm.idx].nseq = L.q0 ? 0u : (uint32_t)(L.seq - L.seq0);      // k_big resolves its own sequences
}

// descriptor + wrapper header of a new member; returns the lane's next state
ZSB_DEV int start_member(Lane &L, uint32_t slot, const Batch &B) {
    const uint32_t idx = B.segs ? B.segs[slot].member : B.order[slot];
    if (idx == 0xffffffffu) return LS_FETCH;        // an empty slot (the fallback list of the segmented members)
    const zsb_member m = B.members[idx];
    Dec &d = L.d;
    Mem &mm = L.mm;
    const uint8_t *in = (const uint8_t *)m.in;
    mm.idx = idx; mm.kind = 0; mm.adler0 = 0; mm.hdr_stop = false; mm.hdr_tin = 0; mm.hdr_bits = 0;
    d.b.p = in;
    d.b.len = (uint32_t)m.in_len;
    d.b.next = 0; d.b.nbits = 0; d.b.hold = 0;
    d.out = (uint8_t *)m.out;
    d.pos = 0;
    d.cap = (uint32_t)m.out_cap;
    d.d64 = m.window_bits == -16;
    d.last = 0; d.at_type = 0; d.msg = 0; d.extent = 0; d.mlen = 0; d.mdist = 0; d.mtrunc = 0;
    L.seq = L.seq0 = B.seq_base + B.seq_off[slot];
    L.lw = 0; L.nl = 0; L.mode = 0; L.pl = 0;
    L.hist = B.hist ? B.hist[idx] : 0u;
    int ev = EV_NONE;
    if (B.segs) {
        Seg &g = B.segs[slot];
        mm.idx = slot;
        // it stops at the next start that was found, and may write sequences up to that segment's range
        uint32_t j = g.k + 1;
        while (j < g.nseg && B.segs[g.first + j].start_bit == SEG_NONE) j++;
        g.target = j < g.nseg ? B.segs[g.first + j].start_bit : SEG_NONE;
        L.seq_lim = L.seq0 + (size_t)(j - g.k) * B.seg_q;
        if (!g.k && g.start_bit == SEG_NONE) {       // the member is not decoded in segments after all (seg::plan_member)
            g.ev = EV_SEGFULL;
            g.nseq = 0;
            return LS_FETCH;
        }
        if (g.k) {
            if (g.start_bit == SEG_NONE) {
                g.ev = EV_ABSENT;
                g.nseq = 0;
                return LS_FETCH;
            }
            // a block header in the middle of the member: no wrapper, unknown output position (at least a window's
            // worth, k_seg_chain checks), no capacity of its own
            bi_seek(d.b, (uint32_t)(g.start_bit >> 3));
            bi_need(d.b, 8);
            bi_drop(d.b, (uint32_t)(g.start_bit & 7u));
            d.cap = 0xfffffff0u;
            d.at_type = 1;
            L.hist = 65536u;
            return LS_BLOCK;
        }
    }
    if (m.flags & ZSB_M_STORED) {          // copied (and checked) by k_check; nothing to decode
        Aux a;
        a.kind = 0; a.trailer_pos = 0; a.need_trailer = 0; a.init_check = 0; a.nseq = 0;
        a.pad[0] = a.pad[1] = a.pad[2] = 0;
        B.aux[idx] = a;
        return LS_FETCH;
    }
    if (!wbits_valid(m.window_bits) || (in == nullptr && m.in_len != 0) || (m.out == nullptr && m.out_cap != 0)) {
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
    if (ev != EV_NONE) {
        finish(L, ev, B);
        return LS_FETCH;
    }
    return LS_BLOCK;
}

// block headers of one lane until a block with symbols starts (reference src/inflate.c:831-1004);
// stored blocks become raw sequences on the way.  Returns the lane's next state.
ZSB_DEV int lane_block(Lane &L, uint8_t *T, const uint32_t *lut, const Batch &B) {
    Dec &d = L.d;
    BitIn &b = d.b;
    for (;;) {
        if (d.last) {
            finish(L, EV_DONE, B);
            return LS_FETCH;
        }
        if (B.segs) {                              // a segment ends at the first block boundary at or behind its target
            if (bi_bitpos(b) >= B.segs[L.mm.idx].target) {
                finish(L, EV_SEGEND, B);
                return LS_FETCH;
            }
            if (L.seq + 16 > L.seq_lim) {
                finish(L, EV_SEGFULL, B);
                return LS_FETCH;
            }
        }
        if (!bi_need(b, 3)) {
            finish(L, EV_STARVE, B);
            return LS_FETCH;
        }
        d.last = bi_peek(b, 1);
        bi_drop(b, 1);
        const uint32_t btype = bi_peek(b, 2);
        bi_drop(b, 2);
        d.at_type = 0;
        if (btype == 3) {
            d.msg = M_BTYPE;
            finish(L, EV_ERR, B);
            return LS_FETCH;
        }
        if (btype == 2) {
            const int ev = lane_dynamic_header(L, T, lut);
            if (ev != EV_NONE) {
                finish(L, ev, B);
                return LS_FETCH;
            }
            return LS_BUILD;
        }
        if (btype == 1) return LS_FIXED;
        // stored block (src/inflate.c:870-904)
        bi_drop(b, b.nbits & 7);
        const uint32_t bytepos = b.next - (b.nbits >> 3);
        if (b.len - bytepos < 4) {
            b.nbits = 8 * (b.len - bytepos);
            b.next = b.len;
            finish(L, EV_STARVE, B);
            return LS_FETCH;
        }
        const uint32_t v = le32(b.p + bytepos);
        if ((v & 0xffffu) != ((v >> 16) ^ 0xffffu)) {
            d.msg = M_STORED;
            d.extent = 32;
            finish(L, EV_ERR, B);
            return LS_FETCH;
        }
        const uint32_t spos = bytepos + 4;
        uint32_t slen = v & 0xffffu;
        const uint32_t have = b.len - spos, left = d.cap - d.pos;
        bool cut = false;
        if (slen > have) slen = have, cut = true;
        if (slen > left) slen = left, cut = true;
        push_raw(L, spos, slen);
        bi_seek(b, spos + slen);
        d.at_type = 1;
        if (!cut && L.q0 && L.seq + BIG_QMARGIN > L.qend) return LS_BLOCK;   // k_big: let the warp empty the queue
        if (cut) {                             // input or output ran out inside the block
            b.next = spos + slen;
            b.nbits = 0;
            b.hold = 0;
            d.at_type = 0;
            finish(L, EV_OUTFULL, B);
            return LS_FETCH;
        }
    }
}

// BIG: one member per warp (lane 0 decodes with the direct tables at `big_lut`, big_run); the other lanes
// only help to build tables.
template <int W, bool BIG = false, bool SEG = false>
__device__ __forceinline__ void warp_lanes(const Batch &B, uint8_t *slots, uint8_t *scr, uint32_t *rings, const uint32_t *lut, int lane,
                                           uint32_t *big_lut = nullptr, seq_t *big_q = nullptr) {
    Lane L;
    L.q0 = BIG ? big_q : nullptr;
    L.qend = BIG ? big_q + BIG_QCAP : nullptr;
    L.qbase = 0;
    L.ring = rings + lane * (RING_STRIDE == 1 ? RING_WORDS : 1);
    L.d.pos = 0; L.d.cap = 0; L.d.b.p = nullptr; L.d.b.len = 0; L.d.b.next = 0; L.d.b.nbits = 0; L.d.b.hold = 0;
    L.d.d64 = 0; L.d.msg = 0; L.nl = 0; L.lw = 0; L.seq = L.seq0 = nullptr; L.nlen = 0; L.ndist = 0;
    L.mm.idx = 0; L.mode = 0; L.pl = 0; L.hist = 0; L.seq_lim = nullptr;
#if defined(__CUDA_ARCH__)
    L.ring_sa = (uint32_t)__cvta_generic_to_shared(L.ring);
#else
    L.ring_sa = 0;
#endif
    uint8_t *T = slots + (size_t)lane * SLOT;
    if (BIG) L.seq = L.seq0 = L.q0;           // an empty queue
    int st = (BIG && lane != 0) ? LS_IDLE : LS_FETCH;
    const unsigned lt = (1u << lane) - 1u;

    for (;;) {
        if (SEG && st == LS_SYM && L.seq + 16 > L.seq_lim) {           // the segment's arena range is full: the member falls back
            finish(L, EV_SEGFULL, B);
            st = LS_FETCH;
        }
        // ---------------- service: until every lane is decoding symbols or out of work ----------------
        for (;;) {
            if (BIG) big_flush<W>(L, big_q, lane);       // before a new member is fetched, a block is read, the tail goes on
            if (!Wp<W>::any(st != LS_SYM && st != LS_IDLE)) break;
            const unsigned fm = Wp<W>::ballot(st == LS_FETCH);
            if (fm) {
                const int leader = __ffs(fm) - 1;
                uint32_t base = 0;
                if (lane == leader) base = ZSB_FETCH_ADD(B.counter, (uint32_t)__popc(fm));
                base = Wp<W>::shfl(base, leader);
                if (st == LS_FETCH) {
                    const uint32_t slot = base + __popc(fm & lt);
                    if (slot >= B.n) st = LS_IDLE;
                    else {
                        st = start_member(L, slot, B);
                        if (BIG) {                       // sequences go to the warp's queue, from output position 0
                            L.seq = L.seq0 = L.q0;
                            L.qbase = 0;
                        }
                    }
                }
            }
            if (st == LS_BLOCK) st = lane_block(L, T, lut, B);
            Wp<W>::sync();
            unsigned bm = Wp<W>::ballot(st == LS_BUILD || st == LS_FIXED);
            while (bm) {
                const int src = __ffs(bm) - 1;
                bm &= bm - 1;
                const uint32_t info = Wp<W>::shfl((uint32_t)(st == LS_FIXED) | (L.d.d64 << 1) | (L.nlen << 2) | (L.ndist << 12), src);
                uint8_t *Ts = slots + (size_t)src * SLOT;
                int msg = 0;
                if (info & 1u) {
                    const uint16_t *f = (const uint16_t *)B.fixed_tab;
                    uint16_t *t16 = (uint16_t *)Ts;
                    for (int i = lane; i < SLOT / 2; i += W) t16[i] = f[i];
                    Wp<W>::sync();
                } else {
                    msg = build_tables<W>(Ts, scr, (info >> 2) & 1023u, info >> 12, (info >> 1) & 1u, lane);
                }
                if (BIG && !msg) big_build<W>(big_lut, Ts, lut, (info >> 1) & 1u, lane);
                if (lane == src) {
                    if (msg) {
                        L.d.msg = msg;
                        finish(L, EV_ERR, B);
                        st = LS_FETCH;
                    } else {
                        st = LS_SYM;
                        if (!BIG) prime(L);
                    }
                }
            }
            if (st == LS_SYM && L.d.b.next + 8 > L.d.b.len) st = LS_TAIL;
            while (st == LS_TAIL) {             // the last bytes of a member: every step checked
                if (BIG && L.seq + 8 > L.qend) break;  // k_big: the queue is emptied at the top of the service loop
                if (SEG && L.seq + 16 > L.seq_lim) {
                    finish(L, EV_SEGFULL, B);
                    st = LS_FETCH;
                    break;
                }
                if (L.mode) L.d.mlen = L.pl;          // the loop stopped between a length and its distance
                int e2 = careful_symbol(L.d, T, lut, L.mode != 0);
                L.mode = 0;
                if (e2 == EV_LIT) {
                    if (L.d.pos >= L.d.cap) e2 = EV_OUTFULL;
                    else {
                        push_literal(L, L.d.mdist);
                        continue;
                    }
                } else if (e2 == EV_MATCH) {
                    e2 = accept_match(L);
                    if (e2 == EV_NONE) continue;
                }
                if (e2 == EV_EOB) {
                    L.d.at_type = 1;
                    st = LS_BLOCK;
                } else {
                    finish(L, e2, B);
                    st = LS_FETCH;
                }
            }
            Wp<W>::sync();
        }
        if (!Wp<W>::any(st == LS_SYM)) break;
        if constexpr (BIG) {         // lane 0 alone: plain symbols through the direct tables, the rest in the service section
            if (st == LS_SYM && !big_run(L, T, lut, big_lut)) st = LS_TAIL;
            Wp<W>::sync();
            big_flush<W>(L, big_q, lane);
        } else {
            // ---------------- the lock-step symbol loop ----------------
#pragma unroll 1
            for (;;) {           // two codes per vote: a lane that has to leave waits one step at most
                if (st == LS_SYM) {
                    if (L.d.b.next + 8 > L.d.b.len || !fast_step(L, T, lut)) st = LS_TAIL;
                }
                if (st == LS_SYM) {
                    if (L.d.b.next + 8 > L.d.b.len || !fast_step(L, T, lut)) st = LS_TAIL;
                }
                if (Wp<W>::any((st != LS_SYM && st != LS_IDLE) || (SEG && st == LS_SYM && L.seq + 16 > L.seq_lim))) break;
            }
        }
    }
}

}  // namespace lane
}  // namespace zsb
