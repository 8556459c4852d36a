// zsb_kernels.cuh — the three kernels of the batched inflate engine (sm_100a):
//   k_inflate<G>   decode: wrapper header, block headers + table build in shared memory, symbol loop,
//                  LZ77 copies, stored blocks.  One group of G lanes per member, persistent CTAs that
//                  fetch members from a global counter (largest first when an order is supplied).
//   k_check        Adler-32 / CRC-32 of each wrapped member's output and the trailer checks
//                  (reference src/inflate.c:1194-1232 and UPDATE_CHECK :424-429), one CTA per member.
//   k_build_fixed  the fixed-code tables (reference src/inflate.c:260-298), once per context.
#pragma once
#include "zsb_device.cuh"

namespace zsb {

// CTA size per group width: as many member slots as shared memory holds (96 x 2384 B), one CTA per SM
// (two for G = 32, which is limited by the 1024-thread CTA instead).
template <int G> struct Geo {
    static constexpr int S = 32 / G;
    static constexpr int threads = G == 1 ? 96 : G == 2 ? 192 : G == 4 ? 384 : G == 8 ? 768 : G == 16 ? 1024 : 512;
    static constexpr int min_ctas = G == 32 ? 2 : 1;
    static constexpr int members = threads / 32 * S;
    static constexpr size_t smem = LUT_BYTES + (size_t)members * STREAM_BYTES;
};

template <int G>
__global__ void __launch_bounds__(Geo<G>::threads, Geo<G>::min_ctas) k_inflate(const zsb_member *__restrict__ members, zsb_result *__restrict__ results,
                          Aux *__restrict__ aux, const uint32_t *__restrict__ order, uint32_t n,
                          uint32_t *counter, const uint16_t *__restrict__ fixed_tab) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *lut = (uint32_t *)smem;
    init_lut(lut, threadIdx.x, blockDim.x);
    __syncthreads();
    constexpr int S = 32 / G;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % G, grp = lane / G;
    uint16_t *T = (uint16_t *)(smem + LUT_BYTES) + (size_t)(warp * S + grp) * STREAM_U16;
    const unsigned gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (grp * G));
    warp_inflate<G>(members, results, aux, order, n, counter, T, lut, fixed_tab, sub, gmask);
}

__global__ void k_build_fixed(uint16_t *fixed_tab) {
    __shared__ __align__(16) uint16_t T[STREAM_U16];
    if (threadIdx.x == 0) build_fixed(T);
    __syncthreads();
    for (int i = threadIdx.x; i < STREAM_U16; i += blockDim.x) fixed_tab[i] = T[i];
}

// ------------------------------------------------------------------------------------------------
// Checksums.  CRC-32 (ISO-HDLC, reflected 0xEDB88320) and Adler-32 (RFC 1950) are the published
// definitions the reference takes from madler/zlib crc32.c / adler32.c (submodule absent there).
// CRC: 128 lanes own 256-byte spans of a 32 KiB tile, tiles are aligned to the END of the data so the
// ragged part is a run of leading zeros (which do not change a zero register); a lane's register is
// carried from tile to tile by multiplying with x^(8*(TILE-SPAN)), and lanes are merged by a tree of
// multiplications with x^(8*SPAN*2^k).  Adler: s1 = 1 + sum(b), s2 = n + n*sum(b) - sum(i*b_i).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t CRC_POLY = 0xEDB88320u;
constexpr int CHK_THREADS = 128;
constexpr uint32_t CHK_SPAN = 256;
constexpr uint32_t CHK_TILE = CHK_THREADS * CHK_SPAN;

// a(x)*b(x) mod P in the reflected representation (bit 31 is x^0)
__host__ __device__ inline uint32_t gf2_mul(uint32_t a, uint32_t b) {
    uint32_t p = 0;
    for (int i = 0; i < 32; i++) {
        if (a & (0x80000000u >> i)) p ^= b;
        b = (b >> 1) ^ (CRC_POLY & (0u - (b & 1u)));
    }
    return p;
}

struct CrcConsts {
    uint32_t x2n[40];      // x^(2^k) mod P
    uint32_t tile_skip;    // x^(8*(TILE-SPAN))
    uint32_t span_pow[8];  // x^(8*SPAN*2^k)
};
__constant__ CrcConsts c_crc;

__device__ __forceinline__ uint32_t crc_xpow(uint64_t nbits) {
    uint32_t r = 0x80000000u;
    for (int k = 0; nbits; k++, nbits >>= 1)
        if (nbits & 1) r = gf2_mul(r, c_crc.x2n[k]);
    return r;
}

__device__ __forceinline__ uint32_t crc_span(uint32_t s, const uint8_t *p, uint32_t n, const uint32_t (*tab)[256]) {
    while (n && ((uintptr_t)p & 3)) {
        s = tab[0][(s ^ *p++) & 0xff] ^ (s >> 8);
        n--;
    }
    const uint32_t *w = (const uint32_t *)p;
    uint32_t nw = n >> 2;
    uint32_t i = 0;
    // Every lane walks a span of its own, so a 32-bit load touches 32 cache lines for 128 bytes and the L1's
    // line-request rate bounds the kernel (1.2 TB/s over the GPU); 128-bit loads bring four times the bytes per request.
    while (i < nw && ((uintptr_t)(w + i) & 15)) {
        s ^= w[i++];
        s = tab[3][s & 0xff] ^ tab[2][(s >> 8) & 0xff] ^ tab[1][(s >> 16) & 0xff] ^ tab[0][s >> 24];
    }
    for (; i + 8 <= nw; i += 8) {
        const uint4 a = *(const uint4 *)(w + i), b = *(const uint4 *)(w + i + 4);
        const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
        for (int k = 0; k < 8; k++) {
            s ^= v[k];
            s = tab[3][s & 0xff] ^ tab[2][(s >> 8) & 0xff] ^ tab[1][(s >> 16) & 0xff] ^ tab[0][s >> 24];
        }
    }
    for (; i < nw; i++) {
        s ^= w[i];
        s = tab[3][s & 0xff] ^ tab[2][(s >> 8) & 0xff] ^ tab[1][(s >> 16) & 0xff] ^ tab[0][s >> 24];
    }
    p += nw * 4;
    n &= 3;
    while (n--) s = tab[0][(s ^ *p++) & 0xff] ^ (s >> 8);
    return s;
}

struct ChkShared {
    uint32_t tab[4][256];
    uint32_t red[CHK_THREADS / 32][2];
    unsigned long long red64[CHK_THREADS / 32][2];
};

__device__ __forceinline__ void chk_init(ChkShared &sh) {
    const int tid = threadIdx.x;
    for (int i = tid; i < 256; i += CHK_THREADS) {
        uint32_t c = i;
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (CRC_POLY & (0u - (c & 1u)));
        sh.tab[0][i] = c;
    }
    __syncthreads();
    for (int k = 1; k < 4; k++) {
        for (int i = tid; i < 256; i += CHK_THREADS) {
            uint32_t c = sh.tab[k - 1][i];
            sh.tab[k][i] = (c >> 8) ^ sh.tab[0][c & 0xff];
        }
        __syncthreads();
    }
}

// The part of a check value that depends on the data alone, over out[0..E) (E > 0) by the whole CTA; valid on
// thread 0.  kind 2 (CRC-32): pa = the data's polynomial remainder (register started at zero).  Else (Adler-32):
// pa = sum of the bytes mod 65521, pb = sum of i * b_i mod 65521 (i from 0).  Parts of a buffer combine:
// remainder(A || B) = remainder(A) * x^(8|B|) + remainder(B); the sums add, the second one with offset * sum.
__device__ __forceinline__ void cta_partial(ChkShared &sh, uint32_t kind, const uint8_t *out, uint32_t E, uint32_t &pa, uint32_t &pb) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pa = 0;
    pb = 0;
    if (kind == 2) {
        uint32_t s = 0;
        const uint32_t ntiles = (E + CHK_TILE - 1) / CHK_TILE;
        for (uint32_t t = ntiles; t-- > 0;) {            // t = tile index counted from the end
            long long lo = (long long)E - (long long)(t + 1) * CHK_TILE + (long long)tid * CHK_SPAN;
            long long hi = lo + CHK_SPAN;
            if (hi > 0) {
                if (lo < 0) lo = 0;
                s = crc_span(s, out + lo, (uint32_t)(hi - lo), sh.tab);
            }
            if (t) s = gf2_mul(c_crc.tile_skip, s);
        }
        // merge lanes: value(tid) * x^(8*SPAN*(127-tid))
        for (int k = 0; k < 5; k++) {
            uint32_t o = __shfl_down_sync(0xffffffffu, s, 1 << k);
            if ((lane & ((2 << k) - 1)) == 0) s = gf2_mul(c_crc.span_pow[k], s) ^ o;
        }
        if (lane == 0) sh.red[warp][0] = s;
        __syncthreads();
        if (tid == 0) {
            uint32_t z = sh.red[0][0];
            for (int wv = 1; wv < CHK_THREADS / 32; wv++) z = gf2_mul(c_crc.span_pow[5], z) ^ sh.red[wv][0];
            pa = z;
        }
        __syncthreads();
    } else {
        const uint32_t P = 65521u;
        unsigned long long S = 0, W = 0;
        uint32_t head = (uint32_t)((16 - ((uintptr_t)out & 15)) & 15);
        if (head > E) head = E;
        const uint32_t nv = (E - head) >> 4;                 // 16-byte vectors: four words per load
        const uint4 *v4 = (const uint4 *)(out + head);
        uint32_t im = (head + 16u * tid) % P;                // byte index of this thread's vector, mod P
        const uint32_t step = (16u * CHK_THREADS) % P;
        for (uint32_t w = tid; w < nv; w += CHK_THREADS) {
            const uint4 v = v4[w];
            const uint32_t s0 = __dp4a(v.x, 0x01010101u, 0u), s1 = __dp4a(v.y, 0x01010101u, 0u);
            const uint32_t s2 = __dp4a(v.z, 0x01010101u, 0u), s3 = __dp4a(v.w, 0x01010101u, 0u);
            const uint32_t q = __dp4a(v.x, 0x03020100u, 0u) + __dp4a(v.y, 0x03020100u, 0u) + __dp4a(v.z, 0x03020100u, 0u) +
                               __dp4a(v.w, 0x03020100u, 0u);
            const uint32_t sb = s0 + s1 + s2 + s3;
            S += sb;
            W += (unsigned long long)im * sb + 4u * (s1 + 2u * s2 + 3u * s3) + q;
            im += step;
            if (im >= P) im -= P;
        }
        if (tid == 0) {
            for (uint32_t i = 0; i < head; i++) { S += out[i]; W += (unsigned long long)i * out[i]; }
            for (uint32_t i = head + 16 * nv; i < E; i++) { S += out[i]; W += (unsigned long long)(i % P) * out[i]; }
        }
        S %= P; W %= P;
        for (int o = 16; o; o >>= 1) {
            S += __shfl_down_sync(0xffffffffu, S, o);
            W += __shfl_down_sync(0xffffffffu, W, o);
        }
        if (lane == 0) { sh.red64[warp][0] = S; sh.red64[warp][1] = W; }
        __syncthreads();
        if (tid == 0) {
            unsigned long long s = 0, wsum = 0;
            for (int wv = 0; wv < CHK_THREADS / 32; wv++) { s += sh.red64[wv][0]; wsum += sh.red64[wv][1]; }
            pa = (uint32_t)(s % P);
            pb = (uint32_t)(wsum % P);
        }
        __syncthreads();
    }
}
// the check value `init` continued over E bytes whose partial values are (pa, pb)
__device__ __forceinline__ uint32_t check_finish(uint32_t kind, uint32_t pa, uint32_t pb, uint32_t E, uint32_t init) {
    if (kind == 2) return ~(gf2_mul(crc_xpow(8ull * E), ~init) ^ pa);
    const uint32_t P = 65521u;
    // s1' = s1 + S;  s2' = s2 + E*s1 + E*S - W   (W = sum of i*b_i, i from 0)
    const unsigned long long nm = E % P, s1 = init & 0xffffu, s2 = init >> 16, s = pa, wsum = pb;
    const uint32_t r1 = (uint32_t)((s1 + s) % P);
    const uint32_t r2 = (uint32_t)((s2 + nm * s1 + nm * s + (unsigned long long)P - wsum) % P);
    return (r2 << 16) | r1;
}
// (pa, pb) of A || B from those of A and of B (B is nb bytes long and starts at offset |A| = off)
__device__ __forceinline__ void check_append(uint32_t kind, uint32_t &pa, uint32_t &pb, uint32_t qa, uint32_t qb, uint32_t off, uint32_t nb) {
    if (kind == 2) {
        pa = gf2_mul(crc_xpow(8ull * nb), pa) ^ qa;
        return;
    }
    const uint32_t P = 65521u;
    pa = (pa + qa) % P;
    pb = (uint32_t)((pb + qb + (unsigned long long)(off % P) * qa) % P);
}

// Check value of `init` continued over out[0..E) by the whole CTA (kind 2 = CRC-32, else Adler-32);
// valid on thread 0.  init is the running value: crc32(0)=0 / adler32(0)=1 for a fresh stream.
__device__ __forceinline__ uint32_t cta_check(ChkShared &sh, uint32_t kind, const uint8_t *out, uint32_t E, uint32_t init) {
    if (E == 0) return init;
    uint32_t pa, pb;
    cta_partial(sh, kind, out, E, pa, pb);
    return threadIdx.x == 0 ? check_finish(kind, pa, pb, E, init) : init;
}

// Large wrapped members: their output is checked in PARTS of CHK_PART bytes by a CTA each (k_check_parts), and k_check
// only puts the parts together — one CTA per member made the 16 MiB members of C4 decide k_check's length.
constexpr uint32_t CHK_PART = 1u << 20;
struct ChkPart {
    uint32_t member, part;        // member index; the part covers [part * CHK_PART, ...) of its output
    uint32_t pa, pb;              // OUT: cta_partial of the part (zero when the member's output ends before it)
};
__global__ void __launch_bounds__(CHK_THREADS)
k_check_parts(const zsb_member *__restrict__ members, const zsb_result *__restrict__ results, const Aux *__restrict__ aux,
              ChkPart *parts, uint32_t nparts) {
    __shared__ ChkShared sh;
    const int tid = threadIdx.x;
    chk_init(sh);
    for (uint32_t i = blockIdx.x; i < nparts; i += gridDim.x) {
        const ChkPart p = parts[i];
        const uint32_t E = (uint32_t)results[p.member].total_out, kind = aux[p.member].kind;
        const uint64_t from = (uint64_t)p.part * CHK_PART;
        uint32_t pa = 0, pb = 0;
        if (kind != 0 && from < E) {                                     // (uniform over the CTA)
            const uint32_t len = E - (uint32_t)from < CHK_PART ? E - (uint32_t)from : CHK_PART;
            cta_partial(sh, kind, (const uint8_t *)members[p.member].out + from, len, pa, pb);
        }
        if (tid == 0) {
            parts[i].pa = pa;
            parts[i].pb = pb;
        }
    }
}

__global__ void __launch_bounds__(CHK_THREADS)
k_check(const zsb_member *__restrict__ members, zsb_result *__restrict__ results, const Aux *__restrict__ aux,
        uint32_t n, uint32_t *counter, const uint32_t *__restrict__ part_first = nullptr, const ChkPart *__restrict__ parts = nullptr) {
    __shared__ ChkShared sh;
    __shared__ uint32_t s_idx;
    const int tid = threadIdx.x;
    chk_init(sh);
    for (;;) {
        if (tid == 0) s_idx = atomicAdd(counter, 1u);
        __syncthreads();
        const uint32_t idx = s_idx;
        __syncthreads();
        if (idx >= n) break;
        const Aux a = aux[idx];
        const int32_t mflags = members[idx].flags;
        if (mflags & ZSB_M_STORED) {                 // ZIP method 0: the CTA copies the bytes (the decoders skipped it)
            const zsb_member m = members[idx];
            const uint8_t *src = (const uint8_t *)m.in;
            uint8_t *dst = (uint8_t *)m.out;
            const bool bad = (src == nullptr && m.in_len != 0) || (dst == nullptr && m.out_cap != 0) ||
                             m.in_len > 0xfffffff0ull || m.out_cap > 0xfffffff0ull;
            const uint32_t nb = bad ? 0u : (uint32_t)(m.in_len < m.out_cap ? m.in_len : m.out_cap);
            if (((uintptr_t)src & 15) == ((uintptr_t)dst & 15)) {
                uint32_t head = (uint32_t)((16 - ((uintptr_t)dst & 15)) & 15);
                if (head > nb) head = nb;
                for (uint32_t i = tid; i < head; i += CHK_THREADS) dst[i] = src[i];
                const uint32_t nv = (nb - head) >> 4;
                const uint4 *s4 = (const uint4 *)(src + head);
                uint4 *d4 = (uint4 *)(dst + head);
                for (uint32_t i = tid; i < nv; i += CHK_THREADS) d4[i] = s4[i];
                for (uint32_t i = head + 16 * nv + tid; i < nb; i += CHK_THREADS) dst[i] = src[i];
            } else {
                for (uint32_t i = tid; i < nb; i += CHK_THREADS) dst[i] = src[i];
            }
            __syncthreads();
            uint32_t crc = 0;
            if (mflags & ZSB_M_CRC32) crc = cta_check(sh, 2u, dst, nb, 0u);
            if (tid == 0) {
                zsb_result r;
                r.ret = bad ? -2 : (m.in_len <= m.out_cap ? 1 : -5);
                r.msg_id = 0;
                r.total_in = nb;
                r.total_out = nb;
                r.check = crc;
                r.data_type = 0;
                results[idx] = r;
            }
            continue;
        }
        if (a.kind == 0) {
            if (mflags & ZSB_M_CRC32) {              // raw member of a container that carries a CRC-32 (ZIP)
                const uint32_t E0 = (uint32_t)results[idx].total_out;
                const uint32_t crc = cta_check(sh, 2u, (const uint8_t *)members[idx].out, E0, 0u);
                if (tid == 0) results[idx].check = crc;
            }
            continue;
        }
        const uint8_t *out = (const uint8_t *)members[idx].out;
        const uint32_t E = (uint32_t)results[idx].total_out;
        uint32_t check = a.init_check;
        if (part_first && part_first[idx] != 0xffffffffu) {       // checked in parts (k_check_parts): put them together
            if (tid == 0 && E) {
                const ChkPart *pp = parts + part_first[idx];
                uint32_t pa = 0, pb = 0;
                for (uint32_t k = 0; (uint64_t)k * CHK_PART < E; k++) {
                    const uint32_t off = k * CHK_PART, nb = E - off < CHK_PART ? E - off : CHK_PART;
                    check_append(a.kind, pa, pb, pp[k].pa, pp[k].pb, off, nb);
                }
                check = check_finish(a.kind, pa, pb, E, a.init_check);
            }
        } else {
            check = cta_check(sh, a.kind, out, E, a.init_check);
        }
        if (tid == 0) {
            zsb_result r = results[idx];
            finish_member(r, a, (const uint8_t *)members[idx].in, (uint32_t)members[idx].in_len, check, E);
            results[idx] = r;
        }
    }
}

}  // namespace zsb

// ------------------------------------------------------------------------------------------------
// Lane path: k_lane (one member per lane, symbols -> LZ sequences) then k_lz (one warp per member,
// sequences -> bytes).  See zsb_lane.cuh / zsb_lz.cuh.
// ------------------------------------------------------------------------------------------------
#include "zsb_lz.cuh"

namespace zsb {

// The CTA's warps come from the launch (1..lane::WARPS); shared memory is laid out for that many, so that
// small CTAs of different launches (the chunks of a pipelined host batch) share an SM.
__global__ void __launch_bounds__(lane::WARPS * 32, 1) k_lane(lane::Batch B) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *lut = (uint32_t *)smem;
    init_lut(lut, threadIdx.x, blockDim.x);
    __syncthreads();
    const int ln = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t warps = blockDim.x >> 5;
    uint8_t *scr = smem + LUT_BYTES + (size_t)warp * lane::WSCR;
    uint8_t *slots = smem + LUT_BYTES + warps * lane::WSCR + (size_t)warp * 32 * lane::SLOT;
    uint32_t *rings = (uint32_t *)(smem + LUT_BYTES + warps * lane::WSCR + warps * 32 * lane::SLOT) + warp * 32 * lane::RING_WORDS;
    lane::warp_lanes<32>(B, slots, scr, rings, lut, ln);
}

// Large members: one member per WARP; lane 0 decodes with direct-indexed tables (lane::big_run) into a
// sequence queue in shared memory, the whole warp resolves the queue (lz::resolve_range): same sequences and
// results as k_lane + k_lz, without the sequence arena.  Per warp: build scratch, one table slot, 3 KB of
// direct tables, a 4 KB queue.
constexpr int BIG_WARPS = 16;
constexpr int BIG_SLOT = 480;                                  // lane::SLOT rounded up to 16
constexpr size_t BIG_PER_WARP = lane::WSCR + BIG_SLOT + lane::BIG_LUT_BYTES + lane::BIG_QCAP * sizeof(lane::seq_t);
constexpr size_t BIG_SMEM = LUT_BYTES + (size_t)BIG_WARPS * BIG_PER_WARP;
__global__ void __launch_bounds__(BIG_WARPS * 32, 1) k_big(lane::Batch B) {
    extern __shared__ __align__(16) uint8_t smem[];
    uint32_t *lut = (uint32_t *)smem;
    init_lut(lut, threadIdx.x, blockDim.x);
    __syncthreads();
    const int ln = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t *base = smem + LUT_BYTES + (size_t)warp * BIG_PER_WARP;
    uint8_t *scr = base, *slot = base + lane::WSCR;
    uint32_t *big_lut = (uint32_t *)(base + lane::WSCR + BIG_SLOT);
    // (the input ring of k_lane is not used here; lane 0's ring pointer only has to be a shared address)
    lane::seq_t *big_q = (lane::seq_t *)(base + lane::WSCR + BIG_SLOT + lane::BIG_LUT_BYTES);
    lane::warp_lanes<32, true>(B, slot, scr, (uint32_t *)scr, lut, ln, big_lut, big_q);
}

constexpr int LZ_THREADS = 256;
__global__ void __launch_bounds__(LZ_THREADS) k_lz(lane::Batch B, uint32_t *counter) {
    const int ln = threadIdx.x & 31;
    for (;;) {
        uint32_t slot = 0;
        if (ln == 0) slot = atomicAdd(counter, 1u);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (slot >= B.n) break;
        const uint32_t idx = B.order[slot];
        const uint32_t nseq = B.aux[idx].nseq;
        if (!nseq) continue;
        const zsb_member m = B.members[idx];
        lz::resolve_member<32>((uint8_t *)m.out, (const uint8_t *)m.in, B.seq_base + B.seq_off[slot], nseq, ln);
    }
}

// tables of the fixed code (reference src/inflate.c:260-298) in the lane path's slot format
__global__ void k_build_fixed_lane(uint8_t *fixed_tab) {
    __shared__ __align__(16) uint8_t T[lane::SLOT];
    __shared__ __align__(16) uint8_t scr[lane::WSCR];
    const int ln = threadIdx.x;
    for (int i = ln; i < 288; i += 32) T[lane::H_LENS + i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
    for (int i = ln; i < 32; i += 32) T[lane::H_LENS + 288 + i] = 5;
    __syncwarp();
    lane::build_tables<32>(T, scr, 288, 32, false, ln);
    __syncwarp();
    for (int i = ln; i < lane::SLOT; i += 32) fixed_tab[i] = T[i];
}

}  // namespace zsb
