// zsb_lz.cuh — the LZ77 resolver ("phase B") of the lane path: one warp per member turns the 8-byte
// sequences written by zsb_lane.cuh into output bytes.  32 sequences per trip, one per lane: output
// positions by a prefix sum over the warp, literals stored at once, short matches copied by their own
// lane (word loads, byte stores), long / self-overlapping matches and stored runs by the whole warp
// with word-wide coalesced accesses.  A match is executed as soon as every byte of its source range
// is final, so independent matches of a trip overlap their memory latency.
//
// Semantics reproduced: out[pos+i] = out[pos-dist + (i mod dist)] — the byte-serial copy of the
// reference (src/inflate.c:1176-1187, src/inffast.c:196-260) including overlapping matches, the
// Deflate64 lengths up to 65538 and distances up to 65536, and stored runs (src/inflate.c:886-900).
#pragma once
#include "zsb_lane.cuh"

namespace zsb {
namespace lz {

using lane::seq_t;
using lane::Wp;

constexpr uint32_t SHORT = 16;            // longest match a single lane copies

ZSB_DEV uint32_t fsr(uint32_t lo, uint32_t hi, uint32_t sh) {   // (hi:lo) >> sh, sh in {0,8,16,24}
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(lo, hi, sh);
#else
    return sh ? (lo >> sh) | (hi << (32 - sh)) : lo;
#endif
}

// dst[i] = src[i], i < n, by the whole warp.  dst and src may overlap if dst - src >= 4*W + 4: the
// copy advances in chunks of W words and every chunk only reads bytes of earlier chunks.
template <int W>
ZSB_DEV void forward_copy(uint8_t *dst, const uint8_t *src, uint32_t n, int lane) {
    uint32_t h = (uint32_t)((4 - ((uintptr_t)dst & 3)) & 3);
    if (h > n) h = n;
    for (uint32_t i = lane; i < h; i += W) dst[i] = src[i];
    dst += h;
    src += h;
    n -= h;
    Wp<W>::sync();
    const uint32_t s = (uint32_t)((uintptr_t)src & 3), sh = 8 * s;
    const uint32_t *sa = (const uint32_t *)(src - s);
    uint32_t *da = (uint32_t *)dst;
    const uint32_t nw = n >> 2;
    for (uint32_t w0 = 0; w0 < nw; w0 += W) {
        const uint32_t w = w0 + lane;
        if (w < nw) {
            const uint32_t a = sa[w];
            const uint32_t b = s ? sa[w + 1] : 0u;
            da[w] = fsr(a, b, sh);
        }
        Wp<W>::sync();
    }
    const uint32_t r = n & 3u;
    for (uint32_t i = lane; i < r; i += W) dst[4 * nw + i] = src[4 * nw + i];
    Wp<W>::sync();
}

// dst[i] = src[i mod dist], i < n, src = dst - dist (a match) or dist >= n (disjoint source)
template <int W>
ZSB_DEV void coop_copy(uint8_t *dst, const uint8_t *src, uint32_t dist, uint32_t n, int lane) {
    if (dist >= n || dist >= 4u * W + 4u) {
        forward_copy<W>(dst, src, n, lane);
        return;
    }
    // short period: lay down a multiple D of the period that is at least 4*W+4 bytes, then the rest
    // is a forward copy at distance D (a sequence of period dist also has period D)
    const uint32_t D = dist * ((4u * W + 3u + dist) / dist);
    const uint32_t m = n < D ? n : D;
    uint32_t off = (uint32_t)lane % dist;
    const uint32_t step = (uint32_t)W % dist;
    for (uint32_t i = lane; i < m; i += W) {
        dst[i] = src[off];
        off += step;
        if (off >= dist) off -= dist;
    }
    Wp<W>::sync();
    if (n > m) forward_copy<W>(dst + m, dst + m - D, n - m, lane);
}

// one lane: n <= SHORT bytes, source and destination disjoint
ZSB_DEV void lane_copy16(uint8_t *dst, const uint8_t *src, uint32_t n) {
    const uint32_t s = (uint32_t)((uintptr_t)src & 3), sh = 8 * s;
    const uint32_t *sa = (const uint32_t *)(src - s);
    const uint32_t k = (s + n + 3) >> 2;           // words that hold source bytes (1..5)
    const uint32_t w0 = sa[0];
    const uint32_t w1 = k > 1 ? sa[1] : 0u;
    const uint32_t w2 = k > 2 ? sa[2] : 0u;
    const uint32_t w3 = k > 3 ? sa[3] : 0u;
    const uint32_t w4 = k > 4 ? sa[4] : 0u;
    const uint32_t d0 = fsr(w0, w1, sh), d1 = fsr(w1, w2, sh), d2 = fsr(w2, w3, sh), d3 = fsr(w3, w4, sh);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if ((uint32_t)i < n) dst[i] = (uint8_t)(d0 >> (8 * i));
        if ((uint32_t)(4 + i) < n) dst[4 + i] = (uint8_t)(d1 >> (8 * i));
        if ((uint32_t)(8 + i) < n) dst[8 + i] = (uint8_t)(d2 >> (8 * i));
        if ((uint32_t)(12 + i) < n) dst[12 + i] = (uint8_t)(d3 >> (8 * i));
    }
}

// nseq sequences of one member whose first byte goes to out[base]; returns the position after the last.
// (k_lz runs a member's whole array from 0; k_big runs its shared-memory queue chunk by chunk.)
template <int W>
ZSB_DEV uint32_t resolve_range(uint8_t *out, const uint8_t *in, const seq_t *seq, uint32_t nseq, int lane, uint32_t base) {
    seq_t qn;                                 // the next trip's sequence is requested one trip ahead
    qn.lo = 0;
    qn.hi = 0;
    if ((uint32_t)lane < nseq) qn = seq[lane];
    for (uint32_t c = 0; c < nseq; c += W) {
        const seq_t q = qn;
        const uint32_t tn = c + W + lane;
        qn.lo = 0;
        qn.hi = 0;
        if (tn < nseq) qn = seq[tn];
        const uint32_t nlit = (q.lo >> lane::Q_NLIT_SHIFT) & 3u;
        const bool is_match = (q.lo & lane::Q_MATCH) != 0, is_raw = (q.lo & lane::Q_RAW) != 0;
        uint32_t mlen = 0;
        if (is_match) mlen = (q.hi >> 16) + ((q.lo & lane::Q_ABSLEN) ? 0u : 3u);
        if (is_raw) mlen = q.lo & 0xffffu;
        const uint32_t dist = (q.hi & 0xffffu) + 1u;
        const uint32_t nbytes = nlit + mlen;
        uint32_t incl = nbytes;
#pragma unroll
        for (int o = 1; o < W; o <<= 1) {
            const uint32_t v = Wp<W>::shfl_up(incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t mypos = base + incl - nbytes;
        base += Wp<W>::shfl(incl, W - 1);
        if (nlit > 0) out[mypos] = (uint8_t)q.lo;
        if (nlit > 1) out[mypos + 1] = (uint8_t)(q.lo >> 8);
        if (nlit > 2) out[mypos + 2] = (uint8_t)(q.lo >> 16);
        const uint32_t mp = mypos + nlit;
        const bool solo = is_match && mlen <= SHORT && dist >= mlen;
        bool pending = mlen != 0;
        // end of the bytes this copy reads, shifted by dist so that it stays non-negative
        const uint64_t need = is_raw ? 0ull : (uint64_t)mp + (dist < mlen ? dist : mlen);
        Wp<W>::sync();
        for (;;) {
            const unsigned pm = Wp<W>::ballot(pending);
            if (!pm) break;
            const uint32_t frontier = Wp<W>::shfl(mp, __ffs(pm) - 1);   // every byte before it is final
            const bool ready = pending && need <= (uint64_t)frontier + dist;
            if (ready && solo) lane_copy16(out + mp, out + mp - dist, mlen);
            unsigned cm = Wp<W>::ballot(ready && !solo);
            while (cm) {
                const int src = __ffs(cm) - 1;
                cm &= cm - 1;
                const uint32_t p = Wp<W>::shfl(mp, src), n = Wp<W>::shfl(mlen, src);
                const uint32_t hi = Wp<W>::shfl(q.hi, src);
                const bool raw = Wp<W>::shfl((uint32_t)is_raw, src) != 0;
                if (raw) coop_copy<W>(out + p, in + hi, n, n, lane);
                else coop_copy<W>(out + p, out + p - ((hi & 0xffffu) + 1u), (hi & 0xffffu) + 1u, n, lane);
            }
            if (ready) pending = false;
            Wp<W>::sync();
        }
    }
    return base;
}

// ------------------------------------------------------------------------------------------------
// resolve_range with the member's recent output mirrored in a 64 KiB RING in shared memory (one warp, one member at
// a time; not for Deflate64, whose window alone is 64 KiB): every byte is written to the output AND to the ring,
// every match source is read from the ring.  A trip's dependent copies then cost a shared-memory round trip instead
// of an L2 round trip, which is what bounds the resolve of a single large member (k_seg_lz_ring, zsb_seg.cuh).
// The ring must have seen every byte of the member from position 0 (segments are resolved in order).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t RING_BYTES = 65536, RING_MASK = RING_BYTES - 1, RING_WMASK = RING_BYTES / 4 - 1;

ZSB_DEV void ring_put(uint8_t *out, uint8_t *ring, uint32_t pos, uint32_t b) {
    out[pos] = (uint8_t)b;
    ring[pos & RING_MASK] = (uint8_t)b;
}
// one lane: n <= SHORT bytes from position src to position dst (dst - src >= n)
ZSB_DEV void lane_copy16_ring(uint8_t *out, uint8_t *ring, uint32_t dst, uint32_t src, uint32_t n) {
    const uint32_t *r32 = (const uint32_t *)ring;
    const uint32_t s = src & 3u, sh = 8 * s, wi = src >> 2;
    const uint32_t k = (s + n + 3) >> 2;           // words that hold source bytes (1..5)
    const uint32_t w0 = r32[wi & RING_WMASK];
    const uint32_t w1 = k > 1 ? r32[(wi + 1) & RING_WMASK] : 0u;
    const uint32_t w2 = k > 2 ? r32[(wi + 2) & RING_WMASK] : 0u;
    const uint32_t w3 = k > 3 ? r32[(wi + 3) & RING_WMASK] : 0u;
    const uint32_t w4 = k > 4 ? r32[(wi + 4) & RING_WMASK] : 0u;
    const uint32_t d0 = fsr(w0, w1, sh), d1 = fsr(w1, w2, sh), d2 = fsr(w2, w3, sh), d3 = fsr(w3, w4, sh);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if ((uint32_t)i < n) ring_put(out, ring, dst + i, d0 >> (8 * i));
        if ((uint32_t)(4 + i) < n) ring_put(out, ring, dst + 4 + i, d1 >> (8 * i));
        if ((uint32_t)(8 + i) < n) ring_put(out, ring, dst + 8 + i, d2 >> (8 * i));
        if ((uint32_t)(12 + i) < n) ring_put(out, ring, dst + 12 + i, d3 >> (8 * i));
    }
}
// the whole warp: byte i of the match comes from position p - dist + (i mod dist), all of which lie before p
template <int W>
ZSB_DEV void coop_copy_ring(uint8_t *out, uint8_t *ring, uint32_t p, uint32_t dist, uint32_t n, int lane) {
    if (dist >= n) {
        for (uint32_t i = lane; i < n; i += W) ring_put(out, ring, p + i, ring[(p - dist + i) & RING_MASK]);
    } else {
        uint32_t off = (uint32_t)lane % dist;
        const uint32_t step = (uint32_t)W % dist;
        for (uint32_t i = lane; i < n; i += W) {
            ring_put(out, ring, p + i, ring[(p - dist + off) & RING_MASK]);
            off += step;
            if (off >= dist) off -= dist;
        }
    }
    Wp<W>::sync();
}
template <int W>
ZSB_DEV void coop_raw_ring(uint8_t *out, uint8_t *ring, uint32_t p, const uint8_t *src, uint32_t n, int lane) {
    for (uint32_t i = lane; i < n; i += W) ring_put(out, ring, p + i, src[i]);
    Wp<W>::sync();
}

template <int W>
ZSB_DEV uint32_t resolve_range_ring(uint8_t *out, const uint8_t *in, const seq_t *seq, uint32_t nseq, int lane, uint32_t base, uint8_t *ring) {
    seq_t qn;
    qn.lo = 0;
    qn.hi = 0;
    if ((uint32_t)lane < nseq) qn = seq[lane];
    for (uint32_t c = 0; c < nseq; c += W) {
        const seq_t q = qn;
        const uint32_t tn = c + W + lane;
        qn.lo = 0;
        qn.hi = 0;
        if (tn < nseq) qn = seq[tn];
        const uint32_t nlit = (q.lo >> lane::Q_NLIT_SHIFT) & 3u;
        const bool is_match = (q.lo & lane::Q_MATCH) != 0, is_raw = (q.lo & lane::Q_RAW) != 0;
        uint32_t mlen = 0;
        if (is_match) mlen = (q.hi >> 16) + ((q.lo & lane::Q_ABSLEN) ? 0u : 3u);
        if (is_raw) mlen = q.lo & 0xffffu;
        const uint32_t dist = (q.hi & 0xffffu) + 1u;
        const uint32_t nbytes = nlit + mlen;
        uint32_t incl = nbytes;
#pragma unroll
        for (int o = 1; o < W; o <<= 1) {
            const uint32_t v = Wp<W>::shfl_up(incl, o);
            if (lane >= o) incl += v;
        }
        const uint32_t mypos = base + incl - nbytes;
        base += Wp<W>::shfl(incl, W - 1);
        if (nlit > 0) ring_put(out, ring, mypos, q.lo);
        if (nlit > 1) ring_put(out, ring, mypos + 1, q.lo >> 8);
        if (nlit > 2) ring_put(out, ring, mypos + 2, q.lo >> 16);
        const uint32_t mp = mypos + nlit;
        const bool solo = is_match && mlen <= SHORT && dist >= mlen;
        bool pending = mlen != 0;
        const uint64_t need = is_raw ? 0ull : (uint64_t)mp + (dist < mlen ? dist : mlen);
        Wp<W>::sync();
        for (;;) {
            const unsigned pm = Wp<W>::ballot(pending);
            if (!pm) break;
            const uint32_t frontier = Wp<W>::shfl(mp, __ffs(pm) - 1);   // every byte before it is final
            const bool ready = pending && need <= (uint64_t)frontier + dist;
            if (ready && solo) lane_copy16_ring(out, ring, mp, mp - dist, mlen);
            unsigned cm = Wp<W>::ballot(ready && !solo);
            while (cm) {
                const int src = __ffs(cm) - 1;
                cm &= cm - 1;
                const uint32_t p = Wp<W>::shfl(mp, src), n = Wp<W>::shfl(mlen, src);
                const uint32_t hi = Wp<W>::shfl(q.hi, src);
                const bool raw = Wp<W>::shfl((uint32_t)is_raw, src) != 0;
                if (raw) coop_raw_ring<W>(out, ring, p, in + hi, n, lane);
                else coop_copy_ring<W>(out, ring, p, (hi & 0xffffu) + 1u, n, lane);
            }
            if (ready) pending = false;
            Wp<W>::sync();
        }
    }
    return base;
}

// all sequences of one member
// (A variant with S = 2 / 4 consecutive sequences per lane and trip — one prefix sum and frontier vote for S times the
// sequences — was measured and dropped: k_lz on C2 13.6 / 24.3 ms against 9.5 ms, 64 / 104 registers against 40.)
template <int W>
ZSB_DEV void resolve_member(uint8_t *out, const uint8_t *in, const seq_t *seq, uint32_t nseq, int lane) {
    resolve_range<W>(out, in, seq, nseq, lane, 0u);
}

}  // namespace lz
}  // namespace zsb
