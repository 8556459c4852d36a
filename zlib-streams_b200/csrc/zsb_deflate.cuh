// zsb_deflate.cuh — batched deflate (compression) on the GPU: raw deflate, zlib and gzip streams, sm_100a.
// the compress half of the product (reference src/wasm/deflate_stream_wasm.c:22-69 drives
// zlib's deflate(); its block writer is src/trees.c).  Nothing is translated from there: the stream layout is
// RFC 1951's, the algorithms below are chosen for a warp.
//
// Unit of work: a BLOCK = up to 64 KiB of one member's input, compressed by one warp.  The window of a block is
// the member's INPUT before it (compression knows all of its history in advance), so the blocks of a large member
// are independent; every block but a member's last ends with an empty stored block, which byte-aligns it (what
// Z_SYNC_FLUSH emits), and k_dstitch concatenates the blocks behind the wrapper header.
//
//   k_deflate   per block: (1) LZ77 — the warp handles 32 consecutive positions at once: a 4-byte hash selects a
//               bucket of the 4 most recent positions with that hash (shared memory, 2^11 buckets x 4 x u16),
//               every lane measures its candidates (plus distance 1, which catches runs inside a tile), then the
//               parse of the tile — greedy with one step of lazy evaluation — is found by pointer doubling over
//               the lanes' "next position" links (five shuffle rounds, no serial walk) and the tokens appended; (2) code construction — symbol frequencies -> length-limited Huffman code
//               lengths (rank sort by the warp, two-queue merge, overflow repair), canonical codes, the run-length
//               coded header of RFC 1951 3.2.7, and the choice between a dynamic, a fixed and stored blocks by
//               exact size; (3) emission — 32 tokens per trip: bit lengths by a warp prefix sum, bits OR-ed into
//               a shared-memory stage, whole words stored coalesced.
//   k_dstitch   per member: wrapper header, its blocks' bytes, Adler-32 / CRC-32 trailer; reports total_out.
#pragma once
#if defined(__CUDACC__)
#include "zsb_kernels.cuh"
#endif
#include "zsb_lane.cuh"

namespace zsb {
namespace df {

using lane::Wp;

constexpr uint32_t BLK = 65536;             // input bytes per block
constexpr int HBITS = 11;                   // 2048 buckets of 4 ways (16 KiB: twelve warps per SM)
constexpr uint32_t WIN = 32768, MINM = 4, MAXM = 258, NICE = 128, LAZY = 32;
constexpr uint32_t TOK_MATCH = 0x80000000u; // token: literal byte, or TOK_MATCH | (len - 3) << 15 | (dist - 1)

// ---- per-warp shared memory (bytes) ----
constexpr int D_HASH = 0;                   // u64[2048]: the matcher's buckets; afterwards the code builder's scratch
constexpr int D_LFREQ = 16384;              // u32[288] literal/length symbol counts
constexpr int D_DFREQ = D_LFREQ + 288 * 4;  // u32[32]
constexpr int D_STAGE = D_DFREQ + 32 * 4;   // u32[64] emission stage
constexpr int D_TILE = D_STAGE + 64 * 4;    // u32[32] tokens of the current tile
constexpr int D_BYTES = D_TILE + 32 * 4;    // 18048
// builder scratch inside D_HASH (the buckets are dead by then)
constexpr int B_ORDER = 0;                  // u16[288] symbols by ascending count
constexpr int B_W = 1024;                   // u32[576] node weights (leaves in order, then internal nodes)
constexpr int B_PAR = 4096;                 // u16[576] parent of each node
constexpr int B_LLEN = 6144;                // u8[288] literal/length code lengths
constexpr int B_DLEN = 6464;                // u8[32]
constexpr int B_CLEN = 6528;                // u8[19] (+pad)
constexpr int B_LCODE = 6656;               // u16[288] codes, bit-reversed
constexpr int B_DCODE = 7296;               // u16[32]
constexpr int B_CCODE = 7424;               // u16[19]
constexpr int B_CFREQ = 7552;               // u32[19]
constexpr int B_ITEMS = 8192;               // u32[400] header items: value | nbits << 24
constexpr int B_RLE = 10240;                // u16[340] run-length coded code lengths: sym | extra << 8
constexpr int B_CNT = 11264;                // u32[16] codes per length

struct Member {                             // device copy of zsb_deflate_member
    const uint8_t *in;
    uint64_t in_len;
    uint8_t *out;
    uint64_t out_cap;
    int32_t window_bits, level;
};
struct Block {
    uint32_t member;
    uint32_t start;                         // offset of the block in the member's input
    uint64_t scratch_off;                   // where its bytes go in the scratch area (4-byte aligned)
    uint32_t last;                          // 1: the member's last block (BFINAL)
    uint32_t nbytes;                        // OUT: bytes produced
};

// unaligned little-endian 32-bit read from global memory through two aligned loads
ZSB_DEV uint32_t ld32u(const uint8_t *p) {
    const uint32_t s = (uint32_t)((uintptr_t)p & 3);
    const uint32_t *a = (const uint32_t *)(p - s);
    const uint32_t lo = __ldg(a);
    if (s == 0) return lo;
    const uint32_t hi = __ldg(a + 1);
    return (lo >> (8 * s)) | (hi << (32 - 8 * s));
}
ZSB_DEV uint32_t hash4(uint32_t v) { return (v * 2654435761u) >> (32 - HBITS); }

// length of the common prefix of a and b, at most `most` (>= 1); both readable for `most` bytes (+3 of slack
// inside the member's input, guarded by the caller through `most`)
ZSB_DEV uint32_t match_len(const uint8_t *a, const uint8_t *b, uint32_t most) {
    uint32_t n = 0;
    while (n + 4 <= most) {
        const uint32_t x = ld32u(a + n) ^ ld32u(b + n);
        if (x) return n + ((uint32_t)__ffs((int)x) - 1) / 8;
        n += 4;
    }
    while (n < most && a[n] == b[n]) n++;
    return n;
}

// ---- symbols of a match (RFC 1951 3.2.5) ----
ZSB_DEV void len_symbol(uint32_t len, uint32_t &sym, uint32_t &xb, uint32_t &xv) {
    const uint32_t l = len - 3;
    if (l < 8) { sym = 257 + l; xb = 0; xv = 0; return; }
    if (len == 258) { sym = 285; xb = 0; xv = 0; return; }
    const uint32_t hb = 31 - (uint32_t)__clz((int)l), g = hb - 2;
    sym = 261 + 4 * g + ((l - (1u << hb)) >> g);
    xb = g;
    xv = l & ((1u << g) - 1u);
}
ZSB_DEV void dist_symbol(uint32_t dist, uint32_t &sym, uint32_t &xb, uint32_t &xv) {
    const uint32_t d = dist - 1;
    if (d < 4) { sym = d; xb = 0; xv = 0; return; }
    const uint32_t hb = 31 - (uint32_t)__clz((int)d);
    sym = 2 * hb + ((d >> (hb - 1)) & 1u);
    xb = hb - 1;
    xv = d & ((1u << xb) - 1u);
}


#ifndef ZSB_SMEM_ADD
#define ZSB_SMEM_ADD(p, v) atomicAdd((p), (v))
#endif
#ifndef ZSB_SMEM_OR
#define ZSB_SMEM_OR(p, v) atomicOr((p), (v))
#endif

// ------------------------------------------------------------------------------------------------
// (1) LZ77 over one block: tokens to tok[], symbol counts to shared memory.  Returns the token count.
// ------------------------------------------------------------------------------------------------
template <int W>
ZSB_DEV void hash_insert(unsigned long long *buckets, bool canhash, uint32_t h, uint32_t cur, int lane) {
    const uint32_t key = canhash ? h : 0x10000u + (uint32_t)lane;        // lanes without a hash form groups of their own
    const unsigned m = Wp<W>::match(key);
    const uint32_t above = lane >= 31 ? 0u : (uint32_t)__popc(m >> (lane + 1));   // newer positions of the tile in this bucket
    const uint32_t cnt = (uint32_t)__popc(m);
    if (canhash && above == 0) {
        const uint32_t k = cnt < 4 ? cnt : 4;
        buckets[h] = k == 4 ? 0ull : buckets[h] << (16 * k);
    }
    Wp<W>::sync();
    if (canhash && above < 4) ((uint16_t *)&buckets[h])[above] = (uint16_t)cur;     // way 0 = the most recent position
    Wp<W>::sync();
}

template <int W>
ZSB_DEV uint32_t lz77_block(uint8_t *sm, const uint8_t *in, uint64_t in_len, uint32_t start, uint32_t n, uint32_t *tok, int lane) {
    unsigned long long *buckets = (unsigned long long *)(sm + D_HASH);
    uint32_t *lfreq = (uint32_t *)(sm + D_LFREQ), *dfreq = (uint32_t *)(sm + D_DFREQ);
    for (int i = lane; i < (1 << HBITS); i += W) buckets[i] = 0;
    for (int i = lane; i < 288 + 32; i += W) lfreq[i] = 0;             // (dfreq follows lfreq)
    Wp<W>::sync();
    // the window before the block: positions are entered, nothing is matched
    const uint32_t hs = start < WIN ? start : WIN;
    for (uint32_t q0 = start - hs; q0 < start; q0 += W) {
        const uint32_t cur = q0 + lane;
        const bool canhash = cur < start && (uint64_t)cur + 4 <= in_len;
        const uint32_t h = canhash ? hash4(ld32u(in + cur)) : 0u;
        hash_insert<W>(buckets, canhash, h, cur, lane);
    }
    uint32_t T = 0, skip = 0;
    for (uint32_t t0 = 0; t0 < n; t0 += W) {
        const uint32_t p = t0 + lane, cur = start + p;
        const bool valid = p < n;
        const bool canhash = valid && (uint64_t)cur + 4 <= in_len;
        uint32_t v4 = 0, h = 0, blen = 0, bdist = 0;
        if (canhash) {
            v4 = ld32u(in + cur);
            h = hash4(v4);
        } else if (valid) {
            v4 = in[cur];
        }
        if (canhash && (uint32_t)lane >= skip) {      // (positions a match of the previous tile covers need no search)
            const uint32_t most = n - p < MAXM ? n - p : MAXM;
            const unsigned long long b = buckets[h];
            if (most >= MINM) {
                // the four candidates' first words are requested together (independent loads), then the survivors
                // are measured
                long long cand[4];
                uint32_t cw[4];
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    const uint32_t e = (uint32_t)(b >> (16 * w)) & 0xffffu;
                    long long c = (long long)((cur & ~0xffffu) | e);
                    if (c >= (long long)cur) c -= 0x10000;
                    const bool okc = c >= 0 && (long long)cur - c <= (long long)WIN;
                    cand[w] = okc ? c : -1;
                    cw[w] = okc ? ld32u(in + c) : ~v4;
                }
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    if (cand[w] < 0 || cw[w] != v4 || blen >= NICE) continue;
                    const uint32_t dist = (uint32_t)((long long)cur - cand[w]);
                    const uint32_t l = 4 + (most > 4 ? match_len(in + cur + 4, in + cand[w] + 4, most - 4) : 0u);
                    if (l > blen || (l == blen && dist < bdist)) {
                        blen = l;
                        bdist = dist;
                    }
                }
                if (cur >= 1 && blen < most && in[cur - 1] == (uint8_t)v4) {     // a run: its source may lie inside this tile
                    const uint32_t l = match_len(in + cur, in + cur - 1, most);
                    if (l >= blen && l >= MINM) {
                        blen = l;
                        bdist = 1;
                    }
                }
            }
            if (blen < MINM) blen = 0;
        }
        hash_insert<W>(buckets, canhash, h, cur, lane);
        // The parse of the tile — greedy with one step of lazy evaluation — without a serial walk: every position knows
        // where the parse would go next from it (a match's end, or the next byte), and the positions the parse visits
        // from `skip` are found by pointer doubling (five rounds for 32 positions, one warp-wide OR each).
        const uint32_t lim = n - t0 < (uint32_t)W ? n - t0 : (uint32_t)W;
        const uint32_t L2 = Wp<W>::shfl_down(blen, 1);
        const bool take = blen >= MINM && !(blen < LAZY && (uint32_t)lane + 1 < lim && W > 1 && L2 > blen);
        uint32_t jump = (uint32_t)lane + (take ? blen : 1u);           // next position of the parse, tile-relative
        const uint32_t nxt = jump;
        if (jump > (uint32_t)W) jump = W;
        unsigned visited = skip < lim ? 1u << skip : 0u;
#pragma unroll
        for (int r = 1; r < W; r <<= 1) {
            const bool mine = (visited >> lane) & 1u;
            visited |= Wp<W>::reduce_or((mine && jump < (uint32_t)W) ? 1u << jump : 0u);
            const uint32_t j2 = Wp<W>::shfl(jump, (int)(jump < (uint32_t)W ? jump : 0u));
            jump = jump < (uint32_t)W ? j2 : (uint32_t)W;
        }
        visited &= lim >= 32 ? 0xffffffffu : (1u << lim) - 1u;
        const bool on = (visited >> lane) & 1u;
        const uint32_t count = (uint32_t)__popc(visited);
        // where the parse leaves the tile: the end of the last visited position's step
        const int lastv = visited ? 31 - __clz((int)visited) : -1;
        const uint32_t endpos = lastv >= 0 ? Wp<W>::shfl(nxt, lastv) : skip;
        skip = endpos > (uint32_t)W ? endpos - W : 0u;
        if (on) {
            const uint32_t t = take ? (TOK_MATCH | ((blen - 3) << 15) | (bdist - 1)) : (v4 & 0xffu);
            tok[T + (uint32_t)__popc(visited & ((1u << lane) - 1u))] = t;
            if (take) {
                uint32_t sym, xb, xv;
                len_symbol(blen, sym, xb, xv);
                ZSB_SMEM_ADD(&lfreq[sym], 1u);
                dist_symbol(bdist, sym, xb, xv);
                ZSB_SMEM_ADD(&dfreq[sym], 1u);
            } else {
                ZSB_SMEM_ADD(&lfreq[t], 1u);
            }
        }
        T += count;
        Wp<W>::sync();
    }
    if (lane == 0) lfreq[256] = 1;                  // end of block
    Wp<W>::sync();
    return T;
}

// ------------------------------------------------------------------------------------------------
// (2) Code construction.  Length-limited Huffman code lengths for n symbols with counts freq[] (shared memory):
// symbols ranked by the warp, the tree by the two-queue merge (lane 0; leaves and internal nodes are both met in
// non-decreasing weight), lengths handed out longest-first to the rarest symbols from the per-length census after
// the overflow repair every deflate encoder needs for the 15 / 7 bit limits.  At least two symbols get a code, so
// every code set written is complete (the decoder's incomplete-set rules never come into play).
// ------------------------------------------------------------------------------------------------
template <int W>
ZSB_DEV void build_lengths(uint8_t *scr, uint32_t *freq, uint32_t n, uint32_t maxbits, uint8_t *lens, int lane) {
    uint16_t *order = (uint16_t *)(scr + B_ORDER), *par = (uint16_t *)(scr + B_PAR);
    uint32_t *wt = (uint32_t *)(scr + B_W), *cnt = (uint32_t *)(scr + B_CNT);
    if (lane == 0) {
        uint32_t nz = 0;
        for (uint32_t s = 0; s < n; s++) nz += freq[s] != 0;
        for (uint32_t s = 0; nz < 2 && s < n; s++)
            if (!freq[s]) {
                freq[s] = 1;
                nz++;
            }
    }
    Wp<W>::sync();
    uint32_t m = 0;
    for (uint32_t s = 0; s < n; s++) m += freq[s] != 0;          // (every lane counts: a broadcast read)
    for (uint32_t s = lane; s < n; s += W) {
        lens[s] = 0;
        const uint32_t f = freq[s];
        if (!f) continue;
        uint32_t r = 0;
        for (uint32_t t = 0; t < n; t++) {
            const uint32_t g = freq[t];
            r += (g && (g < f || (g == f && t < s))) ? 1u : 0u;
        }
        order[r] = (uint16_t)s;
    }
    Wp<W>::sync();
    for (uint32_t i = lane; i < m; i += W) wt[i] = freq[order[i]];
    Wp<W>::sync();
    if (lane == 0) {
        uint32_t li = 0, ni = m, nn = m;
        while (nn < 2 * m - 1) {
            uint32_t pick[2];
            for (int k = 0; k < 2; k++) {
                const bool leaf = li < m && (ni >= nn || wt[li] <= wt[ni]);
                pick[k] = leaf ? li++ : ni++;
            }
            wt[nn] = wt[pick[0]] + wt[pick[1]];
            par[pick[0]] = par[pick[1]] = (uint16_t)nn;
            nn++;
        }
        // depths of the internal nodes (reusing wt), root first; then the census of the leaves' lengths
        // (a node below a clamped node is clamped too and counts as an overflow of its own, internal or leaf: the
        //  repair below moves two of them per round)
        const uint32_t root = 2 * m - 2;
        uint32_t overflow = 0;
        wt[root] = 0;
        for (uint32_t i = root; i-- > m;) {
            uint32_t bits = wt[par[i]] + 1;
            if (bits > maxbits) {
                bits = maxbits;
                overflow++;
            }
            wt[i] = bits;
        }
        for (uint32_t b = 0; b <= 15; b++) cnt[b] = 0;
        for (uint32_t i = 0; i < m; i++) {
            uint32_t bits = wt[par[i]] + 1;
            if (bits > maxbits) {
                bits = maxbits;
                overflow++;
            }
            cnt[bits]++;
        }
        while ((int)overflow > 0) {                  // make room: split a shorter code, drop a maxbits code
            uint32_t bits = maxbits - 1;
            while (cnt[bits] == 0) bits--;
            cnt[bits]--;
            cnt[bits + 1] += 2;
            cnt[maxbits]--;
            overflow -= 2;
        }
        uint32_t i = 0;                               // rarest symbols first, longest codes first
        for (uint32_t bits = maxbits; bits >= 1; bits--)
            for (uint32_t k = cnt[bits]; k; k--) lens[order[i++]] = (uint8_t)bits;
    }
    Wp<W>::sync();
}

// canonical codes (RFC 1951 3.2.2), stored bit-reversed: they are written starting from their most significant bit
template <int W>
ZSB_DEV void assign_codes(uint8_t *scr, const uint8_t *lens, uint32_t n, uint16_t *codes, int lane) {
    uint32_t *cnt = (uint32_t *)(scr + B_CNT);
    if (lane == 0) {
        uint32_t next[17];
        for (int b = 0; b < 16; b++) cnt[b] = 0, next[b] = 0;
        next[16] = 0;
        for (uint32_t s = 0; s < n; s++) cnt[lens[s]]++;
        uint32_t code = 0;
        cnt[0] = 0;
        for (int b = 1; b <= 15; b++) {
            code = (code + cnt[b - 1]) << 1;
            next[b] = code;
        }
        for (uint32_t s = 0; s < n; s++) {
            const uint32_t l = lens[s];
            codes[s] = l ? (uint16_t)(__brev(next[l]++) >> (32 - l)) : 0;
        }
    }
    Wp<W>::sync();
}

// ------------------------------------------------------------------------------------------------
// (3) Emission: `count` items of at most 48 bits each, 32 per trip.  The stage holds the word the bit position is
// in (stage[0], lower bits already written) and what this trip adds; whole words leave coalesced.
// ------------------------------------------------------------------------------------------------
template <int W, typename F>
ZSB_DEV void emit_items(uint8_t *sm, uint32_t *out32, uint32_t &bitpos, uint32_t count, F item, int lane) {
    uint32_t *stage = (uint32_t *)(sm + D_STAGE);
    for (uint32_t i0 = 0; i0 < count; i0 += W) {
        const uint32_t i = i0 + lane;
        unsigned long long v = 0;
        uint32_t nb = 0;
        if (i < count) item(i, v, nb);
        uint32_t incl = nb;
#pragma unroll
        for (int o = 1; o < W; o <<= 1) {
            const uint32_t u = Wp<W>::shfl_up(incl, o);
            if (lane >= o) incl += u;
        }
        const uint32_t total = Wp<W>::shfl(incl, W - 1);
        const uint32_t off = (bitpos & 31u) + incl - nb;
        if (nb) {
            const uint32_t w = off >> 5, sh = off & 31u;
            const uint32_t lo = (uint32_t)(v << sh);
            const uint32_t mid = sh ? (uint32_t)(v >> (32 - sh)) : (uint32_t)(v >> 32);
            const uint32_t top = sh ? (uint32_t)(v >> (64 - sh)) : 0u;
            ZSB_SMEM_OR(&stage[w], lo);
            if (mid) ZSB_SMEM_OR(&stage[w + 1], mid);
            if (top) ZSB_SMEM_OR(&stage[w + 2], top);
        }
        Wp<W>::sync();
        const uint32_t full = ((bitpos & 31u) + total) >> 5, w0 = bitpos >> 5;
        for (uint32_t k = lane; k < full; k += W) out32[w0 + k] = stage[k];
        const uint32_t carry = stage[full];
        Wp<W>::sync();
        for (uint32_t k = lane; k < 64; k += W) stage[k] = k == 0 ? carry : 0u;
        Wp<W>::sync();
        bitpos += total;
    }
}

ZSB_DEV uint32_t len_extra_bits(uint32_t sym) { return (sym < 265 || sym == 285) ? 0u : (sym - 261) >> 2; }
ZSB_DEV uint32_t dist_extra_bits(uint32_t sym) { return sym < 4 ? 0u : (sym >> 1) - 1; }

// One block: in[start, start + n) of member m -> its bytes in `dst` (4-byte aligned scratch).  Returns the byte count.
template <int W>
ZSB_DEV uint32_t deflate_block(uint8_t *sm, const uint8_t *in, uint64_t in_len, uint32_t start, uint32_t n, bool last, int level,
                               uint32_t *tok, uint8_t *dst, int lane) {
    uint8_t *scr = sm + D_HASH;
    uint32_t *lfreq = (uint32_t *)(sm + D_LFREQ), *dfreq = (uint32_t *)(sm + D_DFREQ), *stage = (uint32_t *)(sm + D_STAGE);
    uint8_t *llen = scr + B_LLEN, *dlen = scr + B_DLEN, *clen = scr + B_CLEN;
    uint16_t *lcode = (uint16_t *)(scr + B_LCODE), *dcode = (uint16_t *)(scr + B_DCODE), *ccode = (uint16_t *)(scr + B_CCODE);
    uint32_t *cfreq = (uint32_t *)(scr + B_CFREQ), *items = (uint32_t *)(scr + B_ITEMS);
    uint16_t *rle = (uint16_t *)(scr + B_RLE);
    uint32_t *out32 = (uint32_t *)dst;
    const uint32_t nsub = n ? (n + 65534u) / 65535u : 1u;
    const uint32_t stored_bytes = n + 5u * nsub;
    uint32_t T = 0;
    bool stored = level == 0;
    uint32_t kind = 0;                               // 1 fixed, 2 dynamic
    uint32_t hlit = 257, hdist = 1, hclen = 4, nrle = 0;
    if (!stored) {
        T = lz77_block<W>(sm, in, in_len, start, n, tok, lane);
        // ---- the block's own codes ----
        build_lengths<W>(scr, lfreq, 286, 15, llen, lane);
        build_lengths<W>(scr, dfreq, 30, 15, dlen, lane);
        if (lane == 0) {
            hlit = 286;
            while (hlit > 257 && llen[hlit - 1] == 0) hlit--;
            hdist = 30;
            while (hdist > 1 && dlen[hdist - 1] == 0) hdist--;
            for (int i = 0; i < 19; i++) cfreq[i] = 0;
            // run-length coding of the hlit + hdist code lengths (RFC 1951 3.2.7)
            const uint32_t tot = hlit + hdist;
            uint32_t i = 0, r = 0;
            while (i < tot) {
                const uint32_t v = i < hlit ? llen[i] : dlen[i - hlit];
                uint32_t run = 1;
                while (i + run < tot && (i + run < hlit ? llen[i + run] : dlen[i + run - hlit]) == v) run++;
                i += run;
                if (v == 0) {
                    while (run >= 11) {
                        const uint32_t k = run < 138 ? run : 138;
                        rle[r++] = (uint16_t)(18 | ((k - 11) << 8));
                        cfreq[18]++;
                        run -= k;
                    }
                    if (run >= 3) {
                        rle[r++] = (uint16_t)(17 | ((run - 3) << 8));
                        cfreq[17]++;
                        run = 0;
                    }
                } else {
                    rle[r++] = (uint16_t)v;
                    cfreq[v]++;
                    run--;
                    while (run >= 3) {
                        const uint32_t k = run < 6 ? run : 6;
                        rle[r++] = (uint16_t)(16 | ((k - 3) << 8));
                        cfreq[16]++;
                        run -= k;
                    }
                }
                for (; run; run--) {
                    rle[r++] = (uint16_t)v;
                    cfreq[v]++;
                }
            }
            stage[40] = hlit; stage[41] = hdist; stage[42] = r;      // (the stage is idle until emission)
        }
        Wp<W>::sync();
        hlit = stage[40]; hdist = stage[41]; nrle = stage[42];
        Wp<W>::sync();
        build_lengths<W>(scr, cfreq, 19, 7, clen, lane);
        // ---- exact sizes of the three encodings ----
        uint32_t dyn = 0, fix = 0;
        for (uint32_t s = lane; s < 286; s += W) {
            const uint32_t f = lfreq[s], x = s >= 257 ? len_extra_bits(s) : 0u;
            dyn += f * (llen[s] + x);
            fix += f * ((s < 144 ? 8u : s < 256 ? 9u : s < 280 ? 7u : 8u) + x);
        }
        for (uint32_t s = lane; s < 30; s += W) {
            const uint32_t f = dfreq[s], x = dist_extra_bits(s);
            dyn += f * (dlen[s] + x);
            fix += f * (5u + x);
        }
        for (uint32_t k = lane; k < nrle; k += W) {
            const uint32_t sym = rle[k] & 0xffu;
            dyn += clen[sym] + (sym == 16 ? 2u : sym == 17 ? 3u : sym == 18 ? 7u : 0u);
        }
#pragma unroll
        for (int o = W / 2; o; o >>= 1) {
            dyn += Wp<W>::shfl(dyn, lane ^ o);
            fix += Wp<W>::shfl(fix, lane ^ o);
        }
        dyn = Wp<W>::shfl(dyn, 0);
        fix = Wp<W>::shfl(fix, 0);
        if (lane == 0) {
            const uint8_t perm[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            uint32_t hc = 19;
            while (hc > 4 && clen[perm[hc - 1]] == 0) hc--;
            stage[43] = hc;
        }
        Wp<W>::sync();
        hclen = stage[43];
        Wp<W>::sync();
        dyn += 17 + 3 * hclen;
        fix += 3;
        const uint32_t tail_bits = last ? 0u : 3u + 7u + 32u;          // the empty stored block that aligns a non-final block
        const uint32_t best = dyn < fix ? dyn : fix;
        kind = dyn < fix ? 2u : 1u;
        if (((best + tail_bits + 7) >> 3) >= stored_bytes) stored = true;
    }
    if (stored) {
        // stored blocks of at most 65535 bytes (RFC 1951 3.2.4); the block starts byte aligned
        uint32_t o = 0, done = 0;
        for (uint32_t k = 0; k < nsub; k++) {
            const uint32_t len = n - done < 65535u ? n - done : 65535u;
            if (lane == 0) {
                dst[o] = (uint8_t)((last && k + 1 == nsub) ? 1 : 0);
                dst[o + 1] = (uint8_t)len; dst[o + 2] = (uint8_t)(len >> 8);
                dst[o + 3] = (uint8_t)~len; dst[o + 4] = (uint8_t)(~len >> 8);
            }
            for (uint32_t i = lane; i < len; i += W) dst[o + 5 + i] = in[start + done + i];
            o += 5 + len;
            done += len;
        }
        Wp<W>::sync();
        return o;
    }
    // ---- codes ----
    if (kind == 1) {
        for (uint32_t s = lane; s < 288; s += W) llen[s] = s < 144 ? 8 : s < 256 ? 9 : s < 280 ? 7 : 8;
        for (uint32_t s = lane; s < 32; s += W) dlen[s] = 5;
        Wp<W>::sync();
        assign_codes<W>(scr, llen, 288, lcode, lane);
        assign_codes<W>(scr, dlen, 32, dcode, lane);
    } else {
        assign_codes<W>(scr, llen, 286, lcode, lane);
        assign_codes<W>(scr, dlen, 30, dcode, lane);
        assign_codes<W>(scr, clen, 19, ccode, lane);
    }
    for (uint32_t k = lane; k < 64; k += W) stage[k] = 0;
    Wp<W>::sync();
    uint32_t bitpos = 0;
    // ---- block header ----
    uint32_t nitems = 0;
    if (lane == 0) {
        if (kind == 1) {
            items[0] = (last ? 1u : 0u) | (1u << 1) | (3u << 24);
            nitems = 1;
        } else {
            const uint8_t perm[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            items[0] = ((last ? 1u : 0u) | (2u << 1) | ((hlit - 257) << 3) | ((hdist - 1) << 8) | ((hclen - 4) << 13)) | (17u << 24);
            nitems = 1;
            for (uint32_t i = 0; i < hclen; i++) items[nitems++] = clen[perm[i]] | (3u << 24);
            for (uint32_t k = 0; k < nrle; k++) {
                const uint32_t sym = rle[k] & 0xffu, ex = rle[k] >> 8;
                const uint32_t xb = sym == 16 ? 2u : sym == 17 ? 3u : sym == 18 ? 7u : 0u;
                items[nitems++] = (ccode[sym] | (ex << clen[sym])) | ((clen[sym] + xb) << 24);
            }
        }
        stage[44] = nitems;
    }
    Wp<W>::sync();
    nitems = stage[44];
    Wp<W>::sync();
    if (lane == 0) stage[44] = 0;
    Wp<W>::sync();
    emit_items<W>(sm, out32, bitpos, nitems, [&](uint32_t i, unsigned long long &v, uint32_t &nb) {
        v = items[i] & 0xffffffu;
        nb = items[i] >> 24;
    }, lane);
    // ---- the tokens and the end-of-block code ----
    emit_items<W>(sm, out32, bitpos, T + 1, [&](uint32_t i, unsigned long long &v, uint32_t &nb) {
        if (i == T) {
            v = lcode[256];
            nb = llen[256];
            return;
        }
        const uint32_t t = tok[i];
        if (!(t & TOK_MATCH)) {
            v = lcode[t];
            nb = llen[t];
            return;
        }
        uint32_t ls, lxb, lxv, ds, dxb, dxv;
        len_symbol(((t >> 15) & 0xffu) + 3, ls, lxb, lxv);
        dist_symbol((t & 0x7fffu) + 1, ds, dxb, dxv);
        v = lcode[ls];
        nb = llen[ls];
        v |= (unsigned long long)lxv << nb;
        nb += lxb;
        v |= (unsigned long long)dcode[ds] << nb;
        nb += dlen[ds];
        v |= (unsigned long long)dxv << nb;
        nb += dxb;
    }, lane);
    if (!last) {            // empty stored block: 3 header bits, padding to the byte, LEN = 0, NLEN = 0xffff
        const uint32_t pad = (8u - ((bitpos + 3u) & 7u)) & 7u;
        emit_items<W>(sm, out32, bitpos, 2, [&](uint32_t i, unsigned long long &v, uint32_t &nb) {
            v = i ? 0xffff0000ull : 0ull;
            nb = i ? 32u : 3u + pad;
        }, lane);
    }
    if (lane == 0 && (bitpos & 31u)) out32[bitpos >> 5] = stage[0];
    Wp<W>::sync();
    return (bitpos + 7) >> 3;
}

}  // namespace df

#if defined(__CUDACC__)
// ------------------------------------------------------------------------------------------------
// Kernels
// ------------------------------------------------------------------------------------------------
constexpr int DF_TOK_WORDS = (int)df::BLK + 32;       // token scratch per CTA (a block has at most BLK tokens)

// one warp per CTA; persistent CTAs fetch blocks from a counter
__global__ void __launch_bounds__(32) k_deflate(const df::Member *__restrict__ members, df::Block *blocks, uint32_t nblocks,
                                                uint32_t *counter, uint32_t *tokbuf, uint8_t *scratch) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int ln = threadIdx.x;
    uint32_t *tok = tokbuf + (size_t)blockIdx.x * DF_TOK_WORDS;
    for (;;) {
        uint32_t b = 0;
        if (ln == 0) b = atomicAdd(counter, 1u);
        b = __shfl_sync(0xffffffffu, b, 0);
        if (b >= nblocks) break;
        const df::Block blk = blocks[b];
        const df::Member m = members[blk.member];
        const uint64_t left = m.in_len - blk.start;
        const uint32_t n = left < df::BLK ? (uint32_t)left : df::BLK;
        const uint32_t nb = df::deflate_block<32>(smem, m.in, m.in_len, blk.start, n, blk.last != 0, m.level, tok,
                                                  scratch + blk.scratch_off, ln);
        if (ln == 0) blocks[b].nbytes = nb;
        __syncwarp();
    }
}

struct DfResult {            // = zsb_deflate_result
    int32_t ret, blocks;
    uint64_t total_in, total_out;
    uint32_t check, reserved;
};

// per member: wrapper header + the blocks' bytes + trailer (RFC 1950 / RFC 1952); one CTA per member
__global__ void __launch_bounds__(CHK_THREADS) k_dstitch(const df::Member *__restrict__ members, const df::Block *__restrict__ blocks,
                                                         const uint32_t *__restrict__ first_block, uint32_t n, uint32_t *counter,
                                                         const uint8_t *__restrict__ scratch, DfResult *results) {
    __shared__ ChkShared sh;
    __shared__ uint32_t s_idx, s_check;
    const int tid = threadIdx.x;
    chk_init(sh);
    for (;;) {
        if (tid == 0) s_idx = atomicAdd(counter, 1u);
        __syncthreads();
        const uint32_t idx = s_idx;
        __syncthreads();
        if (idx >= n) break;
        const df::Member m = members[idx];
        const uint32_t b0 = first_block[idx], b1 = first_block[idx + 1];
        const int wb = m.window_bits;
        const uint32_t kind = wb < 0 ? 0u : wb > 15 ? 2u : 1u;
        const uint32_t hdr = kind == 0 ? 0u : kind == 1 ? 2u : 10u, trl = kind == 0 ? 0u : kind == 1 ? 4u : 8u;
        uint64_t body = 0;
        for (uint32_t b = b0; b < b1; b++) body += blocks[b].nbytes;
        const uint64_t total = hdr + body + trl;
        uint32_t check = 0;
        if (kind) check = cta_check(sh, kind, m.in, (uint32_t)m.in_len, kind == 1 ? 1u : 0u);
        if (tid == 0) s_check = check;
        __syncthreads();
        check = s_check;
        const bool fits = total <= m.out_cap;
        if (fits) {
            uint8_t *o = m.out;
            const int lv = m.level;
            if (tid == 0 && kind == 1) {
                const uint32_t fl = lv < 0 ? 2u : lv < 2 ? 0u : lv < 6 ? 1u : lv == 6 ? 2u : 3u;
                uint32_t h16 = (0x78u << 8) | (fl << 6);
                h16 += 31 - h16 % 31;
                o[0] = (uint8_t)(h16 >> 8);
                o[1] = (uint8_t)h16;
            }
            if (tid == 0 && kind == 2) {
                const uint8_t g[10] = {0x1f, 0x8b, 8, 0, 0, 0, 0, 0, (uint8_t)(lv == 9 ? 2 : lv == 1 ? 4 : 0), 3};
                for (int i = 0; i < 10; i++) o[i] = g[i];
            }
            uint64_t at = hdr;
            for (uint32_t b = b0; b < b1; b++) {
                const uint8_t *src = scratch + blocks[b].scratch_off;
                const uint32_t nb = blocks[b].nbytes;
                uint8_t *d = o + at;
                // destination words from (unaligned) source bytes; the edges bytewise
                uint32_t head = (uint32_t)((4 - ((uintptr_t)d & 3)) & 3);
                if (head > nb) head = nb;
                for (uint32_t i = tid; i < head; i += CHK_THREADS) d[i] = src[i];
                const uint32_t nw = (nb - head) >> 2;
                uint32_t *d32 = (uint32_t *)(d + head);
                for (uint32_t w = tid; w < nw; w += CHK_THREADS) d32[w] = df::ld32u(src + head + 4 * w);
                for (uint32_t i = head + 4 * nw + tid; i < nb; i += CHK_THREADS) d[i] = src[i];
                at += nb;
            }
            if (tid == 0 && kind == 1) {
                o[at] = (uint8_t)(check >> 24); o[at + 1] = (uint8_t)(check >> 16); o[at + 2] = (uint8_t)(check >> 8); o[at + 3] = (uint8_t)check;
            }
            if (tid == 0 && kind == 2) {
                const uint32_t isz = (uint32_t)m.in_len;
                for (int i = 0; i < 4; i++) {
                    o[at + i] = (uint8_t)(check >> (8 * i));
                    o[at + 4 + i] = (uint8_t)(isz >> (8 * i));
                }
            }
        }
        if (tid == 0) {
            DfResult r;
            r.ret = fits ? 1 : -5;                  // Z_STREAM_END / Z_BUF_ERROR (the output buffer is too small)
            r.blocks = (int32_t)(b1 - b0);
            r.total_in = fits ? m.in_len : 0;
            r.total_out = fits ? total : 0;
            r.check = check;
            r.reserved = 0;
            results[idx] = r;
        }
        __syncthreads();
    }
}

#endif  // __CUDACC__

}  // namespace zsb
