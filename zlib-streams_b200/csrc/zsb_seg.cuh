// zsb_seg.cuh — parallel decode INSIDE one large member.
//
// A lane of k_lane decodes its member at a fixed pace, k_big gives a member one warp: either way a 16 MiB member
// is one serial Huffman decode (14-18 MB/s).  The bits of a deflate stream can only be decoded from a block start,
// so this path FINDS block starts inside the member, decodes the stretches between them as if they were members
// of their own, and then proves that the stretches fit together:
//
//   k_seg_sync   a lane per multiple of `segb` compressed bytes: inside runs of FIXED-code blocks a parse from any bit
//                position joins the true one (or dies on an invalid code), so the first end-of-block code behind 1 024
//                cleanly parsed codes is a block boundary (find_sync_fixed).
//   k_seg_find   a warp per start still missing: the first bit position at or behind the multiple where a block of
//                the DYNAMIC type starts.  The input is staged in shared memory; a word of 32 positions per lane goes
//                through a 13-bit test done with 64-bit logic (BFINAL = 0, BTYPE = 2, HLIT / HDIST in range), the
//                positions that pass are queued and checked 32 at a time for an exactly complete code-length code (a
//                table of partial Kraft sums), and what is left (about 1 position in 600) is validated eight at a time by
//                the decoder's own header reader (lane_dynamic_header_t<EARLY>) and then by the warp with its own table
//                builder (build_tables) — so a start is accepted exactly when the decoder would accept that header.
//                Stored blocks cannot be told from their three header bits; a stretch simply runs on through them.
//   k_seg_plan   a member with a stretch of more than 6 segments (192 KiB) without a start is not worth a lane: serial.
//   k_lane_seg   k_lane with segments as its slots (zsb_lane.cuh: start_member / lane_block / finish): segment 0
//                starts like the member itself (wrapper, capacity, distance checks exact); segment k > 0 starts at
//                its found bit position at a block header, with 64 KiB of assumed history and no capacity of its
//                own.  A segment ends at the first block boundary at or behind the next found start (EV_SEGEND), at
//                the final block, at an error, or when its range of the sequence arena is full (EV_SEGFULL).
//   k_seg_chain  one thread per member walks its segments in order: each must have ended where the next one began —
//                or, when it ran on past a FALSE start (a header-shaped bit pattern inside a block), where some later
//                one began; the false ones stay dead — and (k > 0) at least a window of output (32 / 64 KiB) must
//                precede it, which makes the assumed history true; output positions are the running sum.  The first
//                segment that ended otherwise — final block, error, input exhausted — is where the serial decoder would
//                have stopped in the same state, so its record, with the position added, gives the member's result
//                through the same write_result().  Anything that does not fit (a gap, a full arena range, output beyond
//                out_cap) marks the member for the fallback: k_big decodes it serially afterwards, so results never
//                depend on a guess.
//   k_seg_lz_ring / k_seg_lz   the verified segments' sequences become bytes, a warp per member, segment after segment:
//                with the member's last 64 KiB of output mirrored in a shared-memory ring (1.6x faster for one member,
//                three members per SM: the wave's largest members; four warps per member when there are few), or on
//                global memory (lz::resolve_range: the rest, and Deflate64 members), side by side.
#pragma once
#include "zsb_lz.cuh"

namespace zsb {
namespace seg {

using lane::Seg;
using lane::SEG_NONE;
using lane::Wp;

// Output that must precede a segment k > 0: the window, so that no distance it accepted could have been "too far
// back" for the serial decoder.  Starts are only looked for behind FIRST_START compressed bytes, where that much
// output exists unless the stream expands its data (then the chain fails and the member is decoded serially).
ZSB_DEV uint32_t min_history(bool d64) { return d64 ? 65536u : 32768u; }
ZSB_DEV uint32_t first_start(bool d64) { return d64 ? 81920u : 40960u; }

// ---- the finder's per-warp shared memory (bytes) ----
constexpr int F_STAGE_WORDS = 256;                    // input words staged per chunk (+ F_STAGE_PAD read-ahead)
constexpr int F_STAGE_PAD = 4;
constexpr int F_VSLOTS = 8;                           // headers validated side by side
constexpr int F_STAGE = 0;                            // u32[260]
constexpr int F_Q1_CAP = 1024 + 64;                   // u16: positions (within the chunk) that passed the 13-bit test
constexpr int F_Q1 = (F_STAGE_WORDS + F_STAGE_PAD) * 4;
constexpr int F_Q2 = F_Q1 + F_Q1_CAP * 2;             // u32[48]: positions whose code-length code is complete (relative to `from`)
constexpr int F_SLOTS = F_Q2 + 48 * 4;                // F_VSLOTS table slots of 480 bytes (16-byte aligned: 3408)
constexpr int F_SCR = F_SLOTS + F_VSLOTS * 480;       // the warp's build scratch
constexpr int F_BYTES = F_SCR + lane::WSCR;           // 8,240
constexpr int KTAB_ENTRIES = 4096;                    // u16 per 4 code-length-code lengths (12 bits): their sum of 2^(7-l)

// Could a dynamic block start here?  Stage 1, the first 13 bits: BFINAL = 0, BTYPE = 2, HLIT and HDIST in range.
ZSB_DEV bool maybe_header(uint32_t a, bool d64) {
    return (a & 7u) == 4u && ((a >> 3) & 31u) <= 29u && (d64 || ((a >> 8) & 31u) <= 29u);
}
// Stage 2, the 4 + 3 * ncode bits behind them: the code-length code exactly complete (an incomplete or
// over-subscribed one is "invalid code lengths set", reference src/inflate.c:936-941 with src/inftrees.c:126-142
// for type CODES).
ZSB_DEV bool complete_clen(uint32_t a, uint32_t b, uint32_t c) {
    const uint32_t ncode = ((a >> 13) & 15u) + 4u;
    const uint64_t x = (((uint64_t)b << 32) | a) >> 17;          // lengths 0..14
    const uint32_t y = (b >> 30) | (c << 2);                     // lengths 15..18
    uint32_t kraft = 0;
#pragma unroll
    for (uint32_t i = 0; i < 19; i++) {
        const uint32_t l = i < 15 ? (uint32_t)(x >> (3 * i)) & 7u : (y >> (3 * (i - 15))) & 7u;
        if (i < ncode) kraft += (128u >> l) & 127u;              // 2^(7-l), nothing for l = 0
    }
    return kraft == 128u;
}
ZSB_DEV uint32_t fsr32(uint32_t lo, uint32_t hi, uint32_t sh) { return sh ? (lo >> sh) | (hi << (32 - sh)) : lo; }
ZSB_DEV uint16_t ktab_entry(uint32_t idx) {
    uint32_t s = 0;
    for (int f = 0; f < 4; f++) s += (128u >> ((idx >> (3 * f)) & 7u)) & 127u;
    return (uint16_t)s;
}
// complete_clen() through the table: five lookups instead of nineteen shifts and adds
ZSB_DEV bool complete_clen_tab(uint32_t a, uint32_t b, uint32_t c, const uint16_t *K) {
    const uint32_t ncode = ((a >> 13) & 15u) + 4u;
    uint64_t v = (((((uint64_t)b << 32) | a) >> 17) & ((1ull << 45) - 1ull)) | ((uint64_t)(((b >> 30) | (c << 2)) & 0xfffu) << 45);
    v &= (1ull << (3u * ncode)) - 1ull;
    const uint32_t sum = (uint32_t)K[(uint32_t)v & 4095u] + K[(uint32_t)(v >> 12) & 4095u] + K[(uint32_t)(v >> 24) & 4095u] +
                         K[(uint32_t)(v >> 36) & 4095u] + K[(uint32_t)(v >> 48) & 4095u];
    return sum == 128u;
}
// the 13-bit test (maybe_header) on the 32 positions of one word at once: bit p of the result = position p passes
ZSB_DEV uint32_t header_mask(uint32_t w0, uint32_t w1, bool d64) {
    const uint64_t x = ((uint64_t)w1 << 32) | w0, nx = ~x;
    uint64_t M = nx & (nx >> 1) & (x >> 2);                              // BFINAL = 0, BTYPE = 2
    M &= ~((x >> 4) & (x >> 5) & (x >> 6) & (x >> 7));                   // HLIT <= 29
    if (!d64) M &= ~((x >> 9) & (x >> 10) & (x >> 11) & (x >> 12));      // HDIST <= 29
    return (uint32_t)M;
}

// The first accepted start in [from, to) bits of the member's input, or SEG_NONE.  All lanes call; sm = F_BYTES of the
// warp's shared memory.  The input is staged a chunk at a time (coalesced), positions go through two queues so that
// every stage runs with full lanes: 13-bit test on W positions per trip -> complete code-length code, W queued
// positions at a time -> the decoder's own header reader, F_VSLOTS at a time, and its table builder, one at a time.
template <int W>
struct Finder {
    const uint8_t *in;
    uint32_t len;
    bool d64;
    uint64_t from;
    uint32_t *stage, *q2;
    uint16_t *q1;
    const uint16_t *ktab;
    uint32_t h1;                 // head of q1
    uint8_t *slots, *scr;
    const uint32_t *lut;
    int lane;
    unsigned lt;
    uint32_t n1, n2;
    uint64_t found;

    // stage 3: the oldest (up to F_VSLOTS) positions of q2 through the decoder's header reader and table builder
    ZSB_DEV void validate() {
        constexpr uint32_t V = F_VSLOTS < W ? F_VSLOTS : W;         // (a warp of one lane in the host emulation)
        const uint32_t tk = n2 < V ? n2 : V;
        int ev = EV_ERR;
        lane::Lane L;
        uint64_t mine = 0;
        L.nlen = 0; L.ndist = 0;
        if ((uint32_t)lane < tk) {
            mine = from + q2[lane];
            L.d.b.p = in; L.d.b.len = len; L.d.d64 = d64 ? 1u : 0u; L.d.msg = 0; L.d.extent = 0;
            bi_seek(L.d.b, (uint32_t)(mine >> 3));
            bi_need(L.d.b, 8);
            bi_drop(L.d.b, (uint32_t)(mine & 7u));
            bi_need(L.d.b, 3);
            bi_drop(L.d.b, 3);
            ev = lane::lane_dynamic_header_t<true>(L, slots + (size_t)lane * 480, lut);
        }
        unsigned pass = Wp<W>::ballot(ev == EV_NONE);
        while (pass && found == SEG_NONE) {
            const int src = __ffs(pass) - 1;
            pass &= pass - 1;
            const uint32_t info = Wp<W>::shfl((L.nlen << 2) | (L.ndist << 12), src);
            const int msg = lane::build_tables<W>(slots + (size_t)src * 480, scr, (info >> 2) & 1023u, info >> 12, d64, lane);
            if (!msg) {
                const uint32_t lo = Wp<W>::shfl((uint32_t)mine, src), hi = Wp<W>::shfl((uint32_t)(mine >> 32), src);
                found = ((uint64_t)hi << 32) | lo;
            }
        }
        Wp<W>::sync();
        for (uint32_t i = (uint32_t)lane; i - (uint32_t)lane + tk < n2; i += W) {     // drop them (pass by pass: read, then write)
            const uint32_t kp = i + tk < n2 ? q2[i + tk] : 0u;
            Wp<W>::sync();
            if (i + tk < n2) q2[i] = kp;
            Wp<W>::sync();
        }
        n2 -= tk;
    }

    // stage 2: the oldest (up to W) positions of q1 (relative to the staged chunk at bit address Ac): complete code-length code?
    ZSB_DEV void kraft(uint64_t Ac, uint64_t A_from) {
        const uint32_t take = n1 < (uint32_t)W ? n1 : (uint32_t)W;
        bool ok = false;
        uint32_t r = 0;
        if ((uint32_t)lane < take) {
            r = q1[h1 + lane];
            const uint32_t wi = r >> 5, sh = r & 31u;
            const uint32_t s0 = stage[wi], s1 = stage[wi + 1], s2 = stage[wi + 2], s3 = stage[wi + 3];
            ok = complete_clen_tab(fsr32(s0, s1, sh), fsr32(s1, s2, sh), fsr32(s2, s3, sh), ktab);
        }
        h1 += take;
        n1 -= take;
        const unsigned m = Wp<W>::ballot(ok);
        if (m) {
            if (ok) q2[n2 + (uint32_t)__popc(m & lt)] = (uint32_t)(Ac + r - A_from);
            n2 += (uint32_t)__popc(m);
            Wp<W>::sync();
            while (n2 >= (uint32_t)F_VSLOTS && found == SEG_NONE) validate();
        }
    }

    ZSB_DEV uint64_t run(uint64_t to) {
        if (len < 64) return SEG_NONE;
        const uint64_t last = 8ull * (len - 24);                 // the 96-bit window and the header reader stay inside the input
        if (to > last) to = last;
        if (from >= to) return SEG_NONE;
        const uint32_t boff = 8u * (uint32_t)((uintptr_t)in & 3);    // bit address A = boff + p in the aligned words at abase
        const uint32_t *abase = (const uint32_t *)(in - (boff >> 3));
        const uint64_t full_words = ((uint64_t)boff + 8ull * len) >> 5;   // whole words inside the input
        const uint64_t A_from = boff + from, A_to = boff + to;
        n1 = n2 = 0;
        found = SEG_NONE;
        for (uint64_t Ac = A_from & ~31ull; Ac < A_to && found == SEG_NONE; Ac += 32ull * F_STAGE_WORDS) {
            const uint64_t w0 = Ac >> 5;
            for (uint32_t i = lane; i < (uint32_t)(F_STAGE_WORDS + F_STAGE_PAD); i += W)
                stage[i] = w0 + i < full_words ? __ldg(abase + w0 + i) : 0u;
            Wp<W>::sync();
            const uint64_t Aend = Ac + 32ull * F_STAGE_WORDS < A_to ? Ac + 32ull * F_STAGE_WORDS : A_to;
            // stage 1: a word (32 positions) per lane and trip; the positions that pass go to q1 in order
            const uint32_t r_lo = A_from > Ac ? (uint32_t)(A_from - Ac) : 0u, r_hi = (uint32_t)(Aend - Ac);
            h1 = 0;
            for (uint32_t j0 = 0; 32u * j0 < r_hi && found == SEG_NONE; j0 += W) {
                const uint32_t j = j0 + (uint32_t)lane;
                uint32_t m32 = 0;
                if (32u * j < r_hi) {
                    m32 = header_mask(stage[j], stage[j + 1], d64);
                    if (32u * j < r_lo) m32 &= r_lo - 32u * j >= 32u ? 0u : ~0u << (r_lo - 32u * j);
                    if (r_hi - 32u * j < 32u) m32 &= (1u << (r_hi - 32u * j)) - 1u;
                }
                uint32_t incl = (uint32_t)__popc(m32);
#pragma unroll
                for (int o = 1; o < W; o <<= 1) {
                    const uint32_t v = Wp<W>::shfl_up(incl, o);
                    if (lane >= o) incl += v;
                }
                const uint32_t total = Wp<W>::shfl(incl, W - 1);
                if (total) {
                    uint32_t at = h1 + n1 + incl - (uint32_t)__popc(m32);
                    while (m32) {
                        q1[at++] = (uint16_t)(32u * j + (uint32_t)__ffs(m32) - 1u);
                        m32 &= m32 - 1u;
                    }
                    n1 += total;
                    Wp<W>::sync();
                    while (n1 >= (uint32_t)W && found == SEG_NONE) kraft(Ac, A_from);
                    // the remainder (< W entries) back to the front of the queue
                    const uint16_t keep = (uint32_t)lane < n1 ? q1[h1 + lane] : (uint16_t)0;
                    Wp<W>::sync();
                    if ((uint32_t)lane < n1) q1[lane] = keep;
                    h1 = 0;
                    Wp<W>::sync();
                }
            }
            while (n1 && found == SEG_NONE) kraft(Ac, A_from);     // (q1 refers to this chunk)
            Wp<W>::sync();
        }
        while (n2 && found == SEG_NONE) validate();
        return found;
    }
};

template <int W>
ZSB_DEV uint64_t find_start(const uint8_t *in, uint32_t len, bool d64, uint64_t from, uint64_t to, uint8_t *sm, const uint16_t *ktab,
                            const uint32_t *lut, int lane) {
    Finder<W> f;
    f.in = in; f.len = len; f.d64 = d64; f.from = from;
    f.stage = (uint32_t *)(sm + F_STAGE); f.q1 = (uint16_t *)(sm + F_Q1); f.q2 = (uint32_t *)(sm + F_Q2);
    f.ktab = ktab; f.h1 = 0;
    f.slots = sm + F_SLOTS; f.scr = sm + F_SCR; f.lut = lut; f.lane = lane;
    f.lt = W == 32 ? ((1u << lane) - 1u) : 0u;
    return f.run(to);
}

// Inside a run of blocks coded with the FIXED code the bits can be parsed from any position: a wrong parse joins the
// true one within a few dozen codes (Huffman codes resynchronise) or meets an invalid code.  From bit `from`: parse
// with the fixed code, follow block headers as long as they say "fixed, not final"; the first end-of-block code behind
// SYNC_MIN codes parsed without an error is taken for a block boundary (the position behind it).  On other data —
// dynamic blocks, stored bytes — 2 % of the codes met are invalid (literal/length 286-287, distance 30-31), so a parse
// survives SYNC_MIN codes with probability < 1e-9.  A lane each, with the fixed code's direct tables (9-bit literal /
// length index, as k_big's); Deflate64 members are left to the dynamic search (their distance codes 30-31 are valid).
// A false boundary is caught by k_seg_chain like any other.
constexpr uint32_t SYNC_MIN = 1024, SYNC_TRIES = 12;
// LL / DD: the fixed code's direct tables (lane::big_build: 512 literal/length entries, 256 distance entries)
// All lanes of a warp call (active = this lane has a boundary to look for): the loop runs one code per trip for every
// lane that is still parsing, so the warp stays converged.
template <int W>
ZSB_DEV uint64_t find_sync_fixed(bool active, const uint8_t *in, uint32_t len, uint64_t from, uint64_t to, const uint32_t *LL) {
    const uint32_t *DD = LL + (1 << lane::BIG_LBITS);
    uint64_t result = SEG_NONE;
    uint64_t A_to = 0, A_from = 0, wi = 0, hold = 0;
    uint32_t boff = 0, nbits = 0, count = 0, attempt = 0;
    const uint32_t *w = nullptr;
    bool alive = active && len >= 64;
    bool fresh = true;                                           // (re)start a parse at A_from + 5 * attempt
    if (alive) {
        const uint64_t last = 8ull * (len - 24);
        if (to > last) to = last;
        boff = 8u * (uint32_t)((uintptr_t)in & 3);               // bit address A = boff + p in the aligned words at w
        w = (const uint32_t *)(in - (boff >> 3));
        A_to = boff + to;
        A_from = boff + from;
        alive = A_from < A_to;
    }
    while (Wp<W>::any(alive)) {
        if (alive) {
            if (fresh) {
                const uint64_t A = A_from + 5ull * attempt;
                if (attempt >= SYNC_TRIES || A >= A_to) alive = false;
                else {
                    wi = (A >> 5) + 1;
                    nbits = 32u - (uint32_t)(A & 31u);
                    hold = (uint64_t)(__ldg(w + wi - 1) >> (A & 31u));
                    count = 0;
                    fresh = false;
                }
            }
        }
        if (alive) {
            if (nbits <= 32u) {
                hold |= (uint64_t)__ldg(w + wi) << nbits;
                nbits += 32u;
                wi++;
            }
            if (32ull * wi - nbits >= A_to) alive = false;       // no boundary within reach
        }
        if (alive) {
            const uint32_t e = LL[(uint32_t)hold & ((1u << lane::BIG_LBITS) - 1u)];
            const uint32_t nb = e & 15u, kind = (e >> 9) & 3u;
            bool dead = false;
            if (kind == lane::BK_LIT) {
                hold >>= nb;
                nbits -= nb;
                count++;
            } else if (kind == lane::BK_BASE) {                  // a length (<= 8 + 5 bits), then its distance (<= 5 + 13 bits)
                const uint32_t xb = (e >> 4) & 31u;
                hold >>= nb + xb;
                nbits -= nb + xb;
                const uint32_t f = DD[(uint32_t)hold & ((1u << lane::BIG_DBITS) - 1u)];
                if (((f >> 9) & 3u) != lane::BK_BASE) dead = true;   // distance code 30 / 31
                const uint32_t n2 = (f & 15u) + ((f >> 4) & 31u);
                hold >>= n2;
                nbits -= n2;
                count++;
            } else if (kind == lane::BK_EOB) {
                hold >>= nb;
                nbits -= nb;
                if (count >= SYNC_MIN) {
                    result = 32ull * wi - nbits - boff;
                    alive = false;
                } else if (((uint32_t)hold & 7u) != 2u) dead = true; // the next block: BFINAL = 0, BTYPE = 1, or the parse is given up
                else {
                    hold >>= 3;
                    nbits -= 3;
                }
            } else dead = true;                                  // literal/length code 286 / 287
            if (dead) {
                attempt++;
                fresh = true;
            }
        }
    }
    return result;
}

// After the searches: a member with a stretch of more than MAX_SPAN wanted segments without a start (fixed or stored
// blocks the searches cannot see, or one huge block) would keep a lane busy far longer than the warp of the serial
// path needs for the whole member, so it is not decoded in segments at all (segment 0 is marked and the chain fails).
constexpr uint32_t MAX_SPAN = 6;
constexpr uint64_t MAX_SPAN_BYTES = 196608;   // ... or this many compressed bytes, whichever is more (small segments)
ZSB_DEV void plan_member(Seg *S, uint32_t segb, uint32_t len, bool d64) {
    const uint32_t nseg = S[0].nseg;
    uint64_t prev = 0;
    bool bad = false;
    for (uint32_t k = 1; k <= nseg && !bad; k++) {
        const uint64_t at = k < nseg ? S[k].start_bit : 8ull * len;
        if (at == SEG_NONE) continue;
        if (at < prev) bad = true;                                   // starts out of order: some start is false
        else if (at - prev > 8ull * (((uint64_t)MAX_SPAN * segb > MAX_SPAN_BYTES ? (uint64_t)MAX_SPAN * segb : MAX_SPAN_BYTES) + (prev ? 0u : first_start(d64)))) bad = true;   // (no start is looked for before first_start)
        prev = at;
    }
    if (bad) {
        S[0].start_bit = SEG_NONE;
        for (uint32_t k = 1; k < nseg; k++) S[k].start_bit = SEG_NONE;
    }
}

// One member's segments in order (see the head of the file).  Returns true when the member's result was written
// from the chain, false when it has to be decoded serially.
ZSB_DEV bool chain_member(const lane::Batch &B, uint32_t first) {
    Seg *S = B.segs + first;
    const uint32_t nseg = S[0].nseg, idx = S[0].member;
    const zsb_member m = B.members[idx];
    uint64_t base = 0;
    uint32_t k = 0;
    bool ok = false;
    for (;;) {
        Seg &g = S[k];
        if (k && base < min_history(m.window_bits == -16)) break;
        if (base > 0xfffffff0ull) break;
        g.base = (uint32_t)base;
        g.live = 1;
        if (g.ev == lane::EV_SEGEND) {
            const uint64_t endbit = 8ull * g.next - g.nbits;
            if (g.target == SEG_NONE || endbit < g.target) break;
            // It stopped at the first block boundary at or behind the next found start.  Behind it: that start was a false
            // one (a header-like bit pattern inside a block — about one per 8 MB searched); the chain goes on with the
            // segment that starts exactly where this one stopped, if there is one, and the false ones stay dead.
            uint32_t j = k + 1;
            while (j < nseg && (S[j].start_bit == SEG_NONE || S[j].start_bit < endbit)) j++;
            if (j >= nseg || S[j].start_bit != endbit) break;
            base += g.pos;
            k = j;
            continue;
        }
        if (g.ev == lane::EV_SEGFULL || g.ev == lane::EV_ABSENT) break;
        if (k && g.ev == EV_OUTFULL) break;
        const uint64_t total = base + g.pos;
        if (total > (uint64_t)m.out_cap) break;
        Dec d;
        Mem mm;
        d.b.p = (const uint8_t *)m.in; d.b.len = (uint32_t)m.in_len; d.b.next = g.next; d.b.nbits = g.nbits; d.b.hold = 0;
        d.out = (uint8_t *)m.out; d.pos = (uint32_t)total; d.cap = (uint32_t)m.out_cap; d.d64 = m.window_bits == -16;
        d.last = g.last; d.at_type = g.at_type; d.msg = g.msg; d.extent = g.extent; d.mlen = 0; d.mdist = 0; d.mtrunc = 0;
        mm.idx = idx; mm.kind = S[0].kind; mm.adler0 = S[0].adler0;
        mm.hdr_stop = k == 0 && g.hdr_stop != 0; mm.hdr_tin = g.hdr_tin; mm.hdr_bits = g.hdr_bits;
        write_result(d, mm, g.ev, B.results + idx, B.aux + idx);
        ok = true;
        break;
    }
    if (!ok)
        for (uint32_t j = 0; j < nseg; j++) S[j].live = 0;
    return ok;
}

// per segmented member, from the host
struct MemberPlan {
    uint32_t idx, first, nseg, pad;
    uint64_t seq_off;           // the member's range of the arena starts here (sequences); segment k at + k * seg_q
};

}  // namespace seg

#if defined(__CUDACC__)
// ------------------------------------------------------------------------------------------------
// Kernels
// ------------------------------------------------------------------------------------------------
// segment records and per-slot sequence offsets from the per-member plan (a thread per member)
__global__ void k_seg_init(const seg::MemberPlan *__restrict__ plan, uint32_t nm, lane::Seg *segs, uint64_t *seq_off, uint32_t seg_q) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    const seg::MemberPlan p = plan[i];
    for (uint32_t k = 0; k < p.nseg; k++) {
        lane::Seg g;
        g.member = p.idx; g.k = k; g.first = p.first; g.nseg = p.nseg;
        g.start_bit = k ? lane::SEG_NONE : 0ull;
        g.target = lane::SEG_NONE;
        g.ev = lane::EV_ABSENT; g.msg = 0; g.pos = 0; g.next = 0; g.nbits = 0; g.extent = 0; g.at_type = 0; g.last = 0;
        g.nseq = 0; g.kind = 0; g.adler0 = 0; g.hdr_tin = 0; g.hdr_bits = 0; g.hdr_stop = 0; g.base = 0; g.live = 0;
        segs[p.first + k] = g;
        seq_off[p.first + k] = p.seq_off + (uint64_t)k * seg_q;
    }
}

constexpr int FIND_WARPS = 12;
constexpr size_t FIND_SMEM = LUT_BYTES + seg::KTAB_ENTRIES * 2 + FIND_WARPS * (size_t)seg::F_BYTES;
__global__ void __launch_bounds__(FIND_WARPS * 32) k_seg_find(const zsb_member *__restrict__ members, lane::Seg *segs, uint32_t nsegs,
                                                              uint32_t segb, uint32_t *counter) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *lut = (uint32_t *)smem;
    init_lut(lut, threadIdx.x, blockDim.x);
    uint16_t *ktab = (uint16_t *)(smem + LUT_BYTES);
    for (uint32_t i = threadIdx.x; i < (uint32_t)seg::KTAB_ENTRIES; i += blockDim.x) ktab[i] = seg::ktab_entry(i);
    __syncthreads();
    const int ln = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t *sm = smem + LUT_BYTES + seg::KTAB_ENTRIES * 2 + (size_t)warp * seg::F_BYTES;
    for (;;) {
        uint32_t slot = 0;
        if (ln == 0) slot = atomicAdd(counter, 1u);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= nsegs) break;
        const uint32_t k = segs[slot].k;
        if (k == 0 || segs[slot].start_bit != lane::SEG_NONE) continue;          // (k_seg_sync found a boundary already)
        const zsb_member m = members[segs[slot].member];
        const uint64_t from = 8ull * k * segb;
        if ((uint64_t)k * segb < seg::first_start(m.window_bits == -16)) continue;        // (stays SEG_NONE)
        const uint64_t found = seg::find_start<32>((const uint8_t *)m.in, (uint32_t)m.in_len, m.window_bits == -16, from,
                                                   from + 8ull * segb, sm, ktab, lut, ln);
        if (ln == 0) segs[slot].start_bit = found;
        __syncwarp();
    }
}

// block boundaries inside runs of fixed-code blocks, a thread per wanted start (before k_seg_find)
constexpr int SYNC_THREADS = 128;
__global__ void __launch_bounds__(SYNC_THREADS) k_seg_sync(const zsb_member *__restrict__ members, lane::Seg *segs, uint32_t nsegs,
                                                            uint32_t segb, const uint8_t *__restrict__ fixed_tab) {
    __shared__ __align__(16) uint32_t lut[LUT_WORDS];
    __shared__ __align__(16) uint8_t T[lane::SLOT + 10];
    __shared__ __align__(16) uint32_t LL[lane::BIG_LUT_BYTES / 4];
    init_lut(lut, threadIdx.x, blockDim.x);
    for (int i = threadIdx.x; i < lane::SLOT; i += blockDim.x) T[i] = fixed_tab[i];
    __syncthreads();
    if (threadIdx.x < 32) lane::big_build<32>(LL, T, lut, 0u, threadIdx.x);
    __syncthreads();
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    bool active = slot < nsegs;
    uint32_t k = 0;
    zsb_member m;
    m.in = nullptr; m.in_len = 0; m.window_bits = 0;
    if (active) {
        k = segs[slot].k;
        m = members[segs[slot].member];
        active = k != 0 && m.window_bits != -16 && (uint64_t)k * segb >= seg::first_start(false);
    }
    const uint64_t from = 8ull * k * segb;
    const uint64_t found = seg::find_sync_fixed<32>(active, (const uint8_t *)m.in, (uint32_t)m.in_len, from, from + 16ull * segb, LL);
    if (active) segs[slot].start_bit = found;
}

__global__ void k_seg_plan(const zsb_member *__restrict__ members, const seg::MemberPlan *__restrict__ plan, uint32_t nm, lane::Seg *segs,
                           uint32_t segb) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    seg::plan_member(segs + plan[i].first, segb, (uint32_t)members[plan[i].idx].in_len, members[plan[i].idx].window_bits == -16);
}

// k_lane over segments: the same CTA layout, the sequence-range checks compiled in
__global__ void __launch_bounds__(lane::WARPS * 32, 1) k_lane_seg(lane::Batch B) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *lut = (uint32_t *)smem;
    init_lut(lut, threadIdx.x, blockDim.x);
    __syncthreads();
    const int ln = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t warps = blockDim.x >> 5;
    uint8_t *scr = smem + LUT_BYTES + (size_t)warp * lane::WSCR;
    uint8_t *slots = smem + LUT_BYTES + warps * lane::WSCR + (size_t)warp * 32 * lane::SLOT;
    uint32_t *rings = (uint32_t *)(smem + LUT_BYTES + warps * lane::WSCR + warps * 32 * lane::SLOT) + warp * 32 * lane::RING_WORDS;
    lane::warp_lanes<32, false, true>(B, slots, scr, rings, lut, ln);
}

// the chain of every segmented member; fail_order[i] = the member's index when it has to be decoded serially
__global__ void k_seg_chain(lane::Batch B, const seg::MemberPlan *__restrict__ plan, uint32_t nm, uint32_t *fail_order) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nm) return;
    const bool ok = seg::chain_member(B, plan[i].first);
    fail_order[i] = ok ? 0xffffffffu : plan[i].idx;
}

constexpr int SEGLZ_WARPS = 8, SEGLZ_THREADS = SEGLZ_WARPS * 32;
// The verified segments' sequences -> bytes, a warp per member, segment after segment, on global memory (Deflate64
// members; the others go through k_seg_lz_ring below).  (A CTA per member — 8 warps taking 32 sequences each per
// round, a frontier in shared memory for matches that reach into the round — resolved one member 1.3x faster but did
// a third of the work per SM, C3 64 ms against 23 ms: measured and dropped.)
__global__ void __launch_bounds__(SEGLZ_THREADS) k_seg_lz(lane::Batch B, const seg::MemberPlan *__restrict__ plan, uint32_t nm,
                                                          const uint64_t *__restrict__ seq_off, uint32_t *counter, uint32_t ring_k) {
    const int ln = threadIdx.x & 31;
    for (;;) {
        uint32_t i = 0;
        if (ln == 0) i = atomicAdd(counter, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= nm) break;
        const seg::MemberPlan p = plan[i];
        const zsb_member m = B.members[p.idx];
        if (i < ring_k && m.window_bits != -16) continue;      // (k_seg_lz_ring resolves those)
        for (uint32_t k = 0; k < p.nseg; k++) {
            const lane::Seg &g = B.segs[p.first + k];
            if (!g.live || !g.nseq) continue;
            lz::resolve_range<32>((uint8_t *)m.out, (const uint8_t *)m.in, B.seq_base + seq_off[p.first + k], g.nseq, ln, g.base);
            __syncwarp();
        }
    }
}

// the same with the member's recent output in a shared-memory ring (lz::resolve_range_ring): one warp per CTA, 64 KiB
// each, three members per SM at a time: 1.6x faster for ONE member (200 against 125 MB/s), an eighth of the
// throughput for many — so it takes the first `nm` (largest) members of a wave, k_seg_lz the rest beside it
// (the host balances the two, run_segmented).  Not for Deflate64 members (k_seg_lz takes those too).
__global__ void __launch_bounds__(32) k_seg_lz_ring(lane::Batch B, const seg::MemberPlan *__restrict__ plan, uint32_t nm,
                                                    const uint64_t *__restrict__ seq_off, uint32_t *counter) {
    extern __shared__ __align__(16) uint8_t ring[];
    const int ln = threadIdx.x;
    for (;;) {
        uint32_t i = 0;
        if (ln == 0) i = atomicAdd(counter, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= nm) break;
        const seg::MemberPlan p = plan[i];
        const zsb_member m = B.members[p.idx];
        if (m.window_bits == -16) continue;                    // (k_seg_lz resolves those)
        for (uint32_t k = 0; k < p.nseg; k++) {
            const lane::Seg &g = B.segs[p.first + k];
            if (!g.live || !g.nseq) continue;
            lz::resolve_range_ring<32>((uint8_t *)m.out, (const uint8_t *)m.in, B.seq_base + seq_off[p.first + k], g.nseq, ln, g.base, ring);
            __syncwarp();
        }
    }
}

// The ring resolver with SEGR_WARPS warps on ONE member and one ring: a ROUND is SEGR_WARPS trips of 32 sequences, a warp
// each.  The warps scan their byte counts and exchange the totals (positions), store their literals and copy every
// match whose source lies before the round at once; a match that reaches into the round waits until the warps before
// it have published that those bytes are final (done[w] = the position up to which warp w's trip is final, ~0 when it
// is complete) — lz::resolve_range's rule with the frontier shared through shared memory.  Sources come from the ring,
// so what a waiting warp waits for is a shared-memory round trip, not an L2 one.  Rounds are separated by a barrier.
constexpr int SEGR_WARPS = 4, SEGR_THREADS = SEGR_WARPS * 32;
__device__ __forceinline__ uint32_t resolve_range_ring_cta(uint8_t *out, const uint8_t *in, const lane::seq_t *seq, uint32_t nseq, uint32_t base,
                                                           uint8_t *ring, volatile uint32_t *done, uint32_t *tot) {
    const int ln = threadIdx.x & 31, w = threadIdx.x >> 5;
    lane::seq_t qn;
    qn.lo = 0;
    qn.hi = 0;
    {
        const uint32_t i0 = (uint32_t)w * 32u + (uint32_t)ln;
        if (i0 < nseq) qn = seq[i0];
    }
    for (uint32_t c = 0; c < nseq; c += 32 * SEGR_WARPS) {
        const lane::seq_t q = qn;
        {
            const uint32_t tn = c + 32 * SEGR_WARPS + (uint32_t)w * 32u + (uint32_t)ln;      // next round's sequence, requested now
            qn.lo = 0;
            qn.hi = 0;
            if (tn < nseq) qn = seq[tn];
        }
        const uint32_t nlit = (q.lo >> lane::Q_NLIT_SHIFT) & 3u;
        const bool is_match = (q.lo & lane::Q_MATCH) != 0, is_raw = (q.lo & lane::Q_RAW) != 0;
        uint32_t mlen = 0;
        if (is_match) mlen = (q.hi >> 16) + ((q.lo & lane::Q_ABSLEN) ? 0u : 3u);
        if (is_raw) mlen = q.lo & 0xffffu;
        const uint32_t dist = (q.hi & 0xffffu) + 1u;
        const uint32_t nbytes = nlit + mlen;
        uint32_t incl = nbytes;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (ln >= o) incl += v;
        }
        if (ln == 31) tot[w] = incl;
        if (ln == 0) done[w] = 0u;
        __syncthreads();
        uint32_t before = 0, all = 0;
#pragma unroll
        for (int k = 0; k < SEGR_WARPS; k++) {
            const uint32_t t = tot[k];
            before += k < w ? t : 0u;
            all += t;
        }
        const uint32_t mypos = base + before + incl - nbytes;
        if (nlit > 0) lz::ring_put(out, ring, mypos, q.lo);
        if (nlit > 1) lz::ring_put(out, ring, mypos + 1, q.lo >> 8);
        if (nlit > 2) lz::ring_put(out, ring, mypos + 2, q.lo >> 16);
        const uint32_t mp = mypos + nlit;
        const bool solo = is_match && mlen <= lz::SHORT && dist >= mlen;
        bool pending = mlen != 0;
        const uint64_t need = is_raw ? 0ull : (uint64_t)mp + (dist < mlen ? dist : mlen);
        __syncwarp();
        for (;;) {
            const unsigned pm = __ballot_sync(0xffffffffu, pending);
            const uint32_t own = pm ? __shfl_sync(0xffffffffu, mp, __ffs(pm) - 1) : 0xffffffffu;
            if (ln == 0) {
                __threadfence_block();                       // the bytes written so far are visible before the position is
                done[w] = own;
            }
            if (!pm) break;
            uint32_t F = own;
            const uint32_t dv = ln < w ? done[ln] : 0xffffffffu;
            __threadfence_block();
            const unsigned inc = __ballot_sync(0xffffffffu, dv != 0xffffffffu);
            if (inc) {
                const uint32_t f2 = __shfl_sync(0xffffffffu, dv, __ffs(inc) - 1);
                F = f2 < F ? f2 : F;
            }
            const bool ready = pending && need <= (uint64_t)F + dist;
            if (ready && solo) lz::lane_copy16_ring(out, ring, mp, mp - dist, mlen);
            unsigned cm = __ballot_sync(0xffffffffu, ready && !solo);
            while (cm) {
                const int src = __ffs(cm) - 1;
                cm &= cm - 1;
                const uint32_t p = __shfl_sync(0xffffffffu, mp, src), n = __shfl_sync(0xffffffffu, mlen, src);
                const uint32_t hi = __shfl_sync(0xffffffffu, q.hi, src);
                const bool raw = __shfl_sync(0xffffffffu, (uint32_t)is_raw, src) != 0;
                if (raw) lz::coop_raw_ring<32>(out, ring, p, in + hi, n, ln);
                else lz::coop_copy_ring<32>(out, ring, p, (hi & 0xffffu) + 1u, n, ln);
            }
            if (ready) pending = false;
            __syncwarp();
        }
        __syncthreads();                                  // the round's bytes are final (and in the ring)
        base += all;
    }
    return base;
}

__global__ void __launch_bounds__(SEGR_THREADS) k_seg_lz_ring_cta(lane::Batch B, const seg::MemberPlan *__restrict__ plan, uint32_t nm,
                                                                   const uint64_t *__restrict__ seq_off, uint32_t *counter) {
    extern __shared__ __align__(16) uint8_t ring[];
    __shared__ uint32_t s_i;
    __shared__ volatile uint32_t done[SEGR_WARPS];
    __shared__ uint32_t tot[SEGR_WARPS];
    for (;;) {
        if (threadIdx.x == 0) s_i = atomicAdd(counter, 1u);
        __syncthreads();
        const uint32_t i = s_i;
        __syncthreads();
        if (i >= nm) break;
        const seg::MemberPlan p = plan[i];
        const zsb_member m = B.members[p.idx];
        if (m.window_bits == -16) continue;                    // (k_seg_lz resolves those)
        for (uint32_t k = 0; k < p.nseg; k++) {
            const lane::Seg &g = B.segs[p.first + k];
            if (!g.live || !g.nseq) continue;
            resolve_range_ring_cta((uint8_t *)m.out, (const uint8_t *)m.in, B.seq_base + seq_off[p.first + k], g.nseq, g.base, ring, done, tot);
        }
    }
}
#endif  // __CUDACC__

}  // namespace zsb
