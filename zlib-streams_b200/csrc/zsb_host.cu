// zsb_host.cu — host side of libzsb.so: contexts, the length-aware scheduler, staging, kernel launches.
// C over the CUDA runtime; the only entry points are the extern "C" functions of include/zsb.h.
// There is no CPU decode path in this library: every member is decoded by k_inflate on the GPU.
#include <cuda_runtime.h>
#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>
#include "zsb_kernels.cuh"
#include "zsb_seg.cuh"
#include "zsb_internal.h"

using namespace zsb;

static thread_local char g_create_err[256] = "";
constexpr int MAX_CHUNKS_DEV = 64;
constexpr int BASE_PIPE_STREAMS = 6;    // decode streams that, with the two copy streams, fit the default 8 hardware queues      // sets of work counters per device (option "chunks" <= 64)

#define CK(ctx, call)                                                                              \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d %s: %s", __FILE__, __LINE__, #call,    \
                     cudaGetErrorString(e_));                                                      \
            return ZSB_E_CUDA;                                                                     \
        }                                                                                          \
    } while (0)

static const char *const k_messages[ZSB_NUM_MESSAGES] = {
    nullptr,
    "incorrect header check",
    "unknown compression method",
    "invalid window size",
    "unknown header flags set",
    "header crc mismatch",
    "invalid block type",
    "invalid stored block lengths",
    "too many length or distance symbols",
    "invalid code lengths set",
    "invalid bit length repeat",
    "invalid code -- missing end-of-block",
    "invalid literal/lengths set",
    "invalid distances set",
    "invalid literal/length code",
    "invalid distance code",
    "invalid distance too far back",
    "incorrect data check",
    "incorrect length check",
};

extern "C" const char *zsb_message(int id) {
    return (id > 0 && id < ZSB_NUM_MESSAGES) ? k_messages[id] : nullptr;
}
extern "C" const char *zsb_create_error(void) { return g_create_err; }
extern "C" const char *zsb_last_error(const zsb_ctx *ctx) { return ctx ? ctx->err : g_create_err; }
extern "C" int zsb_num_devices(const zsb_ctx *ctx) { return ctx ? (int)ctx->devs.size() : 0; }
extern "C" long zsb_last_launches(const zsb_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" void *zsb_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void zsb_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

// Stable sort of the indices idx[0..n) by key, largest first (LSD radix, 11 bits a pass over the bits the keys
// use): the schedulers order 10^6 members per call, where a comparison sort costs ~100 ms of host time per batch.
template <typename KeyOf> static void sort_largest_first(uint32_t *idx, size_t n, KeyOf key_of) {
    if (n < 2) return;
    std::vector<uint64_t> key(n);
    uint64_t all = 0;
    for (size_t i = 0; i < n; i++) {
        key[i] = key_of(idx[i]);
        all |= key[i];
    }
    int bits = 0;
    while (bits < 64 && (all >> bits)) bits++;
    std::vector<uint32_t> tmp_i(n);
    std::vector<uint64_t> tmp_k(n);
    uint32_t *src_i = idx, *dst_i = tmp_i.data();
    uint64_t *src_k = key.data(), *dst_k = tmp_k.data();
    for (int sh = 0; sh < bits; sh += 11) {
        size_t cnt[2049] = {0};
        for (size_t i = 0; i < n; i++) cnt[2047 - ((src_k[i] >> sh) & 2047) + 1]++;      // descending digit order
        for (int b = 0; b < 2048; b++) cnt[b + 1] += cnt[b];
        for (size_t i = 0; i < n; i++) {
            const size_t at = cnt[2047 - ((src_k[i] >> sh) & 2047)]++;
            dst_i[at] = src_i[i];
            dst_k[at] = src_k[i];
        }
        std::swap(src_i, dst_i);
        std::swap(src_k, dst_k);
    }
    if (src_i != idx) memcpy(idx, src_i, n * sizeof(uint32_t));
}

// ---- length-aware sharding: longest processing time first --------------------------------------
extern "C" int zsb_plan_shards(size_t n, const uint64_t *cost, int nshards, int32_t *shard_of, uint64_t *load) {
    if (nshards < 1 || (n && (!cost || !shard_of))) return ZSB_E_ARG;
    std::vector<uint64_t> ld((size_t)nshards, 0);
    std::vector<uint32_t> idx(n);
    for (size_t i = 0; i < n; i++) idx[i] = (uint32_t)i;
    sort_largest_first(idx.data(), n, [&](uint32_t a) { return cost[a]; });
    for (size_t k = 0; k < n; k++) {
        int best = 0;
        for (int s = 1; s < nshards; s++)
            if (ld[s] < ld[best]) best = s;
        shard_of[idx[k]] = best;
        ld[best] += cost[idx[k]] ? cost[idx[k]] : 1;
    }
    if (load)
        for (int s = 0; s < nshards; s++) load[s] = ld[s];
    return 0;
}

// Scratch of one decode (run_device).  NULL = the device's own buffers, grown as needed: only one such
// decode may be in flight per device.  The chunks of a pipelined host batch run concurrently on several
// streams, so each gets a preallocated set.
struct Scratch {
    Aux *aux;              // n entries
    uint8_t *plan;         // n * 12 + 64 bytes: sequence offsets (u64) + slot order (u32)
    uint8_t *seq;          // sequence arena
    size_t seq_bytes;
    uint32_t *counters;    // 8 work counters (4 for the large members' part)
    // Planned by the caller (plan_parts), who has also sent the slot order and the sequence offsets to `plan`
    // and zeroed the counters: run_device then issues no copy of its own (it would queue behind every input
    // copy already submitted).  NULL cuts = not planned.
    const std::vector<size_t> *cuts, *cuts_big;
    size_t nbig;           // the first nbig slots are the large members (split_big)
    uint64_t big_seqs;     // sequences at the start of `seq` that belong to the large members' part
    bool small_ctas;       // k_lane in CTAs of lane::PIPE_WARPS warps, one per 32*PIPE_WARPS members, so that
                           // launches on different streams share the SMs
};

template <int G>
static int launch_inflate(zsb_ctx *ctx, zsb_dev &dv, size_t n, const zsb_member *dm, zsb_result *dr,
                          const uint32_t *dorder, cudaStream_t st, const Scratch *sc) {
    using GG = Geo<G>;
    auto kern = k_inflate<G>;
    if (!dv.configured[G]) {
        CK(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GG::smem));
        int occ = 0;
        CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, GG::threads, GG::smem));
        if (occ < 1) {
            snprintf(ctx->err, sizeof ctx->err, "k_inflate<%d> does not fit an SM", G);
            return ZSB_E_CUDA;
        }
        dv.occ[G] = occ;
        dv.configured[G] = true;
    }
    size_t want = (n + GG::members - 1) / GG::members;
    size_t cap = (size_t)dv.sm_count * dv.occ[G];
    unsigned grid = (unsigned)std::max<size_t>(1, std::min(want, cap));
    kern<<<grid, GG::threads, GG::smem, st>>>(dm, dr, sc ? sc->aux : dv.aux, dorder, (uint32_t)n, (sc ? sc->counters : dv.counters) + 3, dv.fixed);
    ctx->launches++;
    CK(ctx, cudaGetLastError());
    return 0;
}

static int ensure_aux(zsb_ctx *ctx, zsb_dev &dv, size_t n) {
    if (n <= dv.aux_cap) return 0;
    if (dv.aux) CK(ctx, cudaFree(dv.aux));
    dv.aux = nullptr;
    dv.aux_cap = 0;
    size_t cap = std::max<size_t>(n, 1024);
    CK(ctx, cudaMalloc(&dv.aux, cap * sizeof(Aux)));
    dv.aux_cap = cap;
    return 0;
}

// sequences a member can produce at most (zsb_lane.cuh): every sequence but a few carries at least
// three output bytes; stored blocks add at most two short ones each; each costs at least one input bit
static inline uint64_t seq_bound(const zsb_member &m) {
    const uint64_t a = (uint64_t)m.out_cap / 3 + (uint64_t)m.in_len / 2 + 6;
    const uint64_t b = 8 * (uint64_t)m.in_len + 4;
    const uint64_t c = (uint64_t)m.out_cap + 2;
    return std::min(a, std::min(b, c));
}

template <typename T> static int grow(zsb_ctx *ctx, T *&p, size_t &cap, size_t need);

static size_t zsb_seq_bound_sizes(size_t in_len, size_t out_cap) {
    zsb_member m;
    m.in_len = in_len;
    m.out_cap = out_cap;
    return (size_t)seq_bound(m);
}

static int run_group(zsb_ctx *ctx, zsb_dev &dv, size_t n, const zsb_member *dm, zsb_result *dr,
                     const uint32_t *dorder, cudaStream_t st, const Scratch *sc, int lanes) {
    // (the counters were zeroed by run_device, or for a planned chunk by the caller ahead of its input copies)
    switch (lanes) {
    case 1: return launch_inflate<1>(ctx, dv, n, dm, dr, dorder, st, sc);
    case 2: return launch_inflate<2>(ctx, dv, n, dm, dr, dorder, st, sc);
    case 4: return launch_inflate<4>(ctx, dv, n, dm, dr, dorder, st, sc);
    case 8: return launch_inflate<8>(ctx, dv, n, dm, dr, dorder, st, sc);
    case 16: return launch_inflate<16>(ctx, dv, n, dm, dr, dorder, st, sc);
    default: return launch_inflate<32>(ctx, dv, n, dm, dr, dorder, st, sc);
    }
}

// The length-aware split of a batch whose slot order is largest first.  In the lane path a member is
// decoded by ONE lane at a fixed pace, so a member much larger than the rest keeps the batch waiting; k_big
// gives a member a warp of its own and decodes it 2-4 times faster, but the whole GPU then decodes only
// ~26 bytes/ns of in + out against the lane path's ~190.  Both parts run side by side, so the first nbig slots
// go to k_big where that minimises max(time of the big part, time of the rest) under this model (measured,
// scripts/big_pace.py and bench.py; bytes are in_len + out_cap):
//   lane path:  whole part 190 B/ns; its largest member alone 150 ns/B
//   k_big:      whole part 26 B/ns; its largest member alone 40 ns/B with <= 4 members per SM, rising to
//               90 ns/B at 16 per SM (the ALU pipe of an SM sub-partition is shared by its warps)
// Returns nbig (0 for a uniform batch).
static size_t split_big(const zsb_member *hm, const uint32_t *order, size_t n, int sm_count) {
    auto cost = [&](size_t s) {
        const zsb_member &x = hm[order ? order[s] : s];
        return (double)x.in_len + (double)x.out_cap;
    };
    double total = 0;
    for (size_t s = 0; s < n; s++) total += cost(s);
    if (n == 0) return 0;
    const double c0 = cost(0);
    const double lo = 4.0 * sm_count, hi = 16.0 * sm_count;
    double best = std::max(total / 190.0, c0 * 150.0), big_bytes = 0;
    size_t best_n = 0;
    const size_t most = std::min<size_t>(n, (size_t)(2 * hi));          // at most two rounds of warps
    for (size_t k = 1; k <= most; k++) {
        big_bytes += cost(k - 1);
        const double dens = k <= lo ? 0.0 : std::min(1.0, (k - lo) / (hi - lo));
        const double rounds = std::max(1.0, k / hi);
        const double tb = std::max(big_bytes / 26.0, c0 * (40.0 + 50.0 * dens) * rounds);
        const double tl = k < n ? std::max((total - big_bytes) / 190.0, cost(k) * 150.0) : 0.0;
        const double t = std::max(tb, tl);
        if (t < 0.9 * best) {            // worth a second pair of launches only for a clear gain
            best = t;
            best_n = k;
        } else if (best_n && t < best) {
            best = t;
            best_n = k;
        }
        if (tb > best) break;
    }
    return best_n;
}

// the split as a pure host function (include/zsb.h): members given largest first
extern "C" size_t zsb_plan_big(size_t n, const uint64_t *in_len, const uint64_t *out_cap, int sm_count) {
    if (!n || !in_len || !out_cap || sm_count < 1) return 0;
    std::vector<zsb_member> hm(n);
    for (size_t i = 0; i < n; i++) {
        hm[i].in_len = (size_t)in_len[i];
        hm[i].out_cap = (size_t)out_cap[i];
    }
    return split_big(hm.data(), nullptr, n, sm_count);
}

// Per-slot sequence offsets (off[n]) and the slices [cuts[k], cuts[k+1]) of the slot order, cut where the
// arena of arena_seqs sequences would overflow.  Returns the most sequences a slice needs.
static uint64_t lane_plan(const zsb_member *hm, const uint32_t *horder, size_t n, uint64_t arena_seqs, uint64_t *off,
                          std::vector<size_t> &cuts) {
    if (arena_seqs < 1) arena_seqs = 1;
    uint64_t run = 0, most = 0;
    cuts.clear();
    cuts.push_back(0);
    for (size_t s = 0; s < n; s++) {
        const uint64_t b = seq_bound(hm[horder ? horder[s] : s]);
        if (run && run + b > arena_seqs) {
            cuts.push_back(s);
            run = 0;
        }
        off[s] = run;
        run += b;
        most = std::max(most, run);
    }
    cuts.push_back(n);
    return most;
}

// Most sequence scratch a batch may hold at once: option "seq_arena_bytes", or (0, the default) half of what
// the device has free: the worst-case bound of seq_bound is ~3x what members write, and every
// extra slice of a batch of mixed sizes ends with its own tail (C4: 1.76 s in four slices, 1.27 s in one).
// (cudaMemGetInfo costs milliseconds, so the answer is kept until a batch would not fit it.)
static size_t arena_bytes(const zsb_ctx *ctx, zsb_dev &dv, bool refresh = false) {
    if (ctx->seq_arena_bytes) return ctx->seq_arena_bytes;
    if (!dv.arena_auto || refresh) {
        size_t fr = 0, tot = 0;
        if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) {
            cudaGetLastError();
            fr = 0;
        }
        // half of what is free (our own arena counts as free); when the query fails, a modest 4 GiB
        dv.arena_auto = fr ? std::max<size_t>((size_t)64 << 20, (fr + dv.d_seq_cap) / 2) : ((size_t)4 << 30);
    }
    return dv.arena_auto;
}

// The plan of a batch whose first nbig slots are its large members: both parts' sequence offsets (off[n]; the
// small part's start after the big part's big_seqs sequences) and slices.  Returns the sequences needed.
static uint64_t plan_parts(const zsb_member *hm, const uint32_t *order, size_t n, size_t nbig, uint64_t arena_seqs,
                           uint64_t *off, std::vector<size_t> &cuts_big, std::vector<size_t> &cuts, uint64_t &big_seqs) {
    // the large members' part needs no arena (k_big keeps its sequences in shared memory): one slice
    big_seqs = 0;
    cuts_big.clear();
    cuts_big.push_back(0);
    cuts_big.push_back(nbig);
    for (size_t s = 0; s < nbig; s++) off[s] = 0;
    const uint64_t left = arena_seqs;
    const uint64_t most = n > nbig ? lane_plan(hm, order ? order + nbig : nullptr, n - nbig, left, off + nbig, cuts) : 0;
    for (size_t s = nbig; s < n; s++) off[s] += big_seqs;
    return big_seqs + most;
}

// One part of a batch on the lane path: k_lane (or k_big: a warp per member, for the large members) + k_lz
// over the slices `cuts` of the part's slot order.  dorder / dseqoff point at the part's first slot.
struct LanePart {
    const uint32_t *dorder;
    const uint64_t *dseqoff;
    const std::vector<size_t> *cuts;
    Aux *aux;
    lane::seq_t *seq_base;
    uint32_t *counters;        // 4 of them, zero for the first slice
    cudaStream_t st;
    bool small_ctas, big, timed;
};

static int lane_configure(zsb_ctx *ctx, zsb_dev &dv) {
    if (dv.lane_configured) return 0;
    CK(ctx, cudaFuncSetAttribute(k_lane, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lane::SMEM));
    // all of the SM's memory as shared memory also for the small CTAs (the default carve-out fits fewer of them)
    CK(ctx, cudaFuncSetAttribute(k_lane, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    int occ = 0;
    CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_lane, lane::WARPS * 32, lane::SMEM));
    if (occ < 1) {
        snprintf(ctx->err, sizeof ctx->err, "k_lane does not fit an SM");
        return ZSB_E_CUDA;
    }
    CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dv.lane_small_occ, k_lane, lane::PIPE_WARPS * 32,
                                                          lane::smem_for(lane::PIPE_WARPS)));
    CK(ctx, cudaFuncSetAttribute(k_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BIG_SMEM));
    CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dv.big_occ, k_big, BIG_WARPS * 32, BIG_SMEM));
    if (dv.big_occ < 1) dv.big_occ = 1;
    CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dv.lz_occ, k_lz, LZ_THREADS, 0));
    if (dv.lz_occ < 1) dv.lz_occ = 1;
    dv.lane_configured = true;
    return 0;
}

static int launch_lane(zsb_ctx *ctx, zsb_dev &dv, const zsb_member *dm, zsb_result *dr, const LanePart &P) {
    int rc0 = lane_configure(ctx, dv);
    if (rc0) return rc0;
    cudaStream_t st = P.st;
    uint32_t *counters = P.counters;
    const std::vector<size_t> &cuts = *P.cuts;
    for (size_t k = 0; k + 1 < cuts.size(); k++) {
        const size_t first = cuts[k], cnt = cuts[k + 1] - cuts[k];
        // (the caller zeroed the counters for the first slice; [1] is k_check's and stays)
        if (k > 0) {
            CK(ctx, cudaMemsetAsync(counters, 0, sizeof(uint32_t), st));
            CK(ctx, cudaMemsetAsync(counters + 2, 0, sizeof(uint32_t), st));
        }
        lane::Batch B;
        B.members = dm;
        B.results = dr;
        B.aux = P.aux;
        B.order = P.dorder + first;
        B.seq_off = P.dseqoff + first;
        B.seq_base = P.seq_base;
        B.hist = nullptr;
        B.n = (uint32_t)cnt;
        B.counter = counters;
        B.fixed_tab = dv.fixed_lane;
        size_t warps, per_cta;
        unsigned grid;
        if (P.big) {
            // a warp per member; with few members, CTAs of just the warps needed so that they spread over the SMs
            const size_t per_sm = (cnt + dv.sm_count - 1) / dv.sm_count;
            warps = std::min<size_t>(BIG_WARPS, std::max<size_t>(1, per_sm));
            per_cta = warps;
            grid = (unsigned)std::max<size_t>(1, std::min((cnt + warps - 1) / warps, (size_t)dv.sm_count * dv.big_occ));
            cudaEvent_t *pb = nullptr;
            if (ctx->profile) {
                if (dv.prof_used + 3 > dv.prof_ev.size()) {
                    const size_t old = dv.prof_ev.size();
                    dv.prof_ev.resize(old + 96);
                    for (size_t i = old; i < dv.prof_ev.size(); i++) CK(ctx, cudaEventCreate(&dv.prof_ev[i]));
                }
                pb = &dv.prof_ev[dv.prof_used];
                dv.prof_used += 3;
                CK(ctx, cudaEventRecord(pb[0], st));
            }
            k_big<<<grid, (unsigned)(warps * 32), LUT_BYTES + warps * BIG_PER_WARP, st>>>(B);
            ctx->launches++;
            CK(ctx, cudaGetLastError());
            if (pb) {                                  // (k_big resolves its own sequences: no k_lz for this part)
                CK(ctx, cudaEventRecord(pb[1], st));
                CK(ctx, cudaEventRecord(pb[2], st));
            }
            continue;
        }
        if (P.small_ctas || ctx->force_small_ctas) {
            warps = lane::PIPE_WARPS;
            per_cta = warps * 32;
            grid = (unsigned)std::max<size_t>(1, std::min((cnt + per_cta - 1) / per_cta, (size_t)dv.sm_count * dv.lane_small_occ));
        } else {
            // one CTA per SM, with just the warps the slice needs so that a small slice spreads over all SMs
            const size_t per_sm = (cnt + dv.sm_count - 1) / dv.sm_count;
            warps = std::min<size_t>(lane::WARPS, std::max<size_t>(1, (per_sm + 31) / 32));
            per_cta = warps * 32;
            grid = (unsigned)std::max<size_t>(1, std::min((cnt + per_cta - 1) / per_cta, (size_t)dv.sm_count));
        }
        cudaEvent_t *pe = nullptr;
        if (ctx->profile && P.timed) {
            if (dv.prof_used + 3 > dv.prof_ev.size()) {
                const size_t old = dv.prof_ev.size();
                dv.prof_ev.resize(old + 96);
                for (size_t i = old; i < dv.prof_ev.size(); i++) CK(ctx, cudaEventCreate(&dv.prof_ev[i]));
            }
            pe = &dv.prof_ev[dv.prof_used];
            dv.prof_used += 3;
            CK(ctx, cudaEventRecord(pe[0], st));
        }
        k_lane<<<grid, (unsigned)per_cta, lane::smem_for(warps), st>>>(B);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        if (pe) CK(ctx, cudaEventRecord(pe[1], st));
        const size_t wpc = LZ_THREADS / 32;
        // (fewer resident CTAs to keep the members' windows in L2 were measured slower: 1/2/3/4/5 CTAs per SM
        // = 26.4/15.2/11.8/10.3/9.8 ms on C2 against 9.5 ms at the full 6; 8 CTAs of 32 registers: 11.2 ms)
        grid = (unsigned)std::max<size_t>(1, std::min((cnt + wpc - 1) / wpc, (size_t)dv.sm_count * dv.lz_occ));
        k_lz<<<grid, LZ_THREADS, 0, st>>>(B, counters + 2);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        if (pe) CK(ctx, cudaEventRecord(pe[2], st));
    }
    return 0;
}


// ---- large members in segments (zsb_seg.cuh) ----------------------------------------------------------------------
// Compressed bytes per segment.  Option "seg_bytes", or (0, the default) from the batch: a lane decodes its segment at
// a fixed pace, so the segments should fill the GPU's lanes in whole rounds — one round of larger segments rather than
// a round and a third of 32 KiB ones (C3) — and a batch of few members gets small segments (8 KiB: their decode is
// what a single member waits for).
static uint32_t seg_bytes_for(const zsb_ctx *ctx, const zsb_dev &dv, const zsb_member *hm, const std::vector<uint32_t> &segm) {
    if (ctx->seg_bytes) return (uint32_t)ctx->seg_bytes;
    double total = 0;
    for (uint32_t i : segm) total += (double)hm[i].in_len;
    const double lanes = 0.94 * dv.sm_count * lane::WARPS * 32;      // (every member rounds its last segment up)
    const double rounds = std::max(1.0, std::floor(total / (lanes * 32768.0) + 0.35));
    const double sb = std::min(65536.0, std::max(8192.0, total / (lanes * rounds)));
    return ((uint32_t)sb + 255u) & ~255u;
}
static inline uint32_t seg_q_of(uint32_t segb) { return segb / 4 * 3; }   // sequences of arena per segment

static int seg_configure(zsb_ctx *ctx, zsb_dev &dv) {
    if (dv.seg_configured) return 0;
    CK(ctx, cudaFuncSetAttribute(k_lane_seg, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lane::SMEM));
    CK(ctx, cudaFuncSetAttribute(k_lane_seg, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    CK(ctx, cudaFuncSetAttribute(k_seg_find, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FIND_SMEM));
    CK(ctx, cudaFuncSetAttribute(k_seg_lz_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lz::RING_BYTES));
    CK(ctx, cudaFuncSetAttribute(k_seg_lz_ring_cta, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lz::RING_BYTES));
    CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dv.find_occ, k_seg_find, FIND_WARPS * 32, FIND_SMEM));
    if (dv.find_occ < 1) dv.find_occ = 1;
    dv.seg_configured = true;
    return 0;
}

// The members `segm` (indices into dm / hm) through find -> k_lane_seg -> chain -> k_seg_lz -> fallback, in waves that
// fit the sequence arena (dv.d_seq, already allocated by the caller).  Everything is queued on st.
static int run_segmented(zsb_ctx *ctx, zsb_dev &dv, const zsb_member *dm, zsb_result *dr, const zsb_member *hm,
                         const std::vector<uint32_t> &segm, cudaStream_t st, lane::seq_t *arena, uint64_t arena_seqs, uint32_t segb) {
    int rc;
    if ((rc = lane_configure(ctx, dv)) || (rc = seg_configure(ctx, dv))) return rc;
    const uint32_t seg_q = seg_q_of(segb);
    const size_t nm = segm.size();
    // waves: whole members, first-fit in order; slots and arena offsets restart in every wave
    std::vector<seg::MemberPlan> plan(nm);
    std::vector<size_t> wave_m, wave_slot;          // first member / first slot of each wave (+ end)
    size_t nsegs = 0;
    uint64_t run = 0;
    uint32_t wslot = 0;
    wave_m.push_back(0);
    wave_slot.push_back(0);
    for (size_t i = 0; i < nm; i++) {
        const zsb_member &x = hm[segm[i]];
        const uint32_t nseg = (uint32_t)((x.in_len + segb - 1) / segb);
        const uint64_t want = (uint64_t)nseg * seg_q;
        if (want > arena_seqs) {
            snprintf(ctx->err, sizeof ctx->err, "sequence arena too small for a member of %zu bytes in segments", (size_t)x.in_len);
            return ZSB_E_NOMEM;
        }
        if (run + want > arena_seqs) {
            wave_m.push_back(i);
            wave_slot.push_back(nsegs);
            run = 0;
            wslot = 0;
        }
        plan[i].idx = segm[i]; plan[i].first = wslot; plan[i].nseg = nseg; plan[i].pad = 0; plan[i].seq_off = run;
        run += want;
        wslot += nseg;
        nsegs += nseg;
    }
    wave_m.push_back(nm);
    wave_slot.push_back(nsegs);
    if (nsegs > 0x7fffffffull) return ZSB_E_ARG;
    // device buffers: plan | fail order | counters | per-slot sequence offsets | segment records
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o_plan = 0, o_fail = o_plan + up(nm * sizeof(seg::MemberPlan)), o_cnt = o_fail + up(nm * 4),
                 o_off = o_cnt + up((wave_m.size() * 8 + 8) * 4), o_seg = o_off + up(nsegs * 8), total = o_seg + up(nsegs * sizeof(lane::Seg));
    if ((rc = grow(ctx, dv.d_segbuf, dv.d_segbuf_cap, total))) return rc;
    uint8_t *base = dv.d_segbuf;
    seg::MemberPlan *dplan = (seg::MemberPlan *)(base + o_plan);
    uint32_t *dfail = (uint32_t *)(base + o_fail), *dcnt = (uint32_t *)(base + o_cnt);
    uint64_t *doff = (uint64_t *)(base + o_off);
    lane::Seg *dsegs = (lane::Seg *)(base + o_seg);
    CK(ctx, cudaMemcpyAsync(dplan, plan.data(), nm * sizeof(seg::MemberPlan), cudaMemcpyHostToDevice, st));
    CK(ctx, cudaMemsetAsync(dcnt, 0, (wave_m.size() * 8 + 8) * 4, st));
    static const bool trace = getenv("ZSB_TRACE") != nullptr;
    if (trace) fprintf(stderr, "zsb trace segments: %zu members, %zu segments of %u bytes, %zu wave(s)\n", nm, nsegs, segb, wave_m.size() - 1);
    for (size_t w = 0; w + 1 < wave_m.size(); w++) {
        const size_t m0 = wave_m[w], mc = wave_m[w + 1] - m0, s0 = wave_slot[w], sn = wave_slot[w + 1] - s0;
        uint32_t *cnt = dcnt + 8 * w;               // [0] find, [1] lanes, [2] resolve (ring), [3] fallback, [4] resolve (Deflate64)
        k_seg_init<<<(unsigned)((mc + 127) / 128), 128, 0, st>>>(dplan + m0, (uint32_t)mc, dsegs + s0, doff + s0, seg_q);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        k_seg_sync<<<(unsigned)((sn + SYNC_THREADS - 1) / SYNC_THREADS), SYNC_THREADS, 0, st>>>(dm, dsegs + s0, (uint32_t)sn, segb, dv.fixed_lane);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        k_seg_find<<<(unsigned)std::min<size_t>((sn + FIND_WARPS - 1) / FIND_WARPS, (size_t)dv.sm_count * dv.find_occ), FIND_WARPS * 32, FIND_SMEM, st>>>(
            dm, dsegs + s0, (uint32_t)sn, segb, cnt);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        k_seg_plan<<<(unsigned)((mc + 127) / 128), 128, 0, st>>>(dm, dplan + m0, (uint32_t)mc, dsegs + s0, segb);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        lane::Batch B;
        B.segs = dsegs + s0;
        B.seg_q = seg_q;
        B.members = dm;
        B.results = dr;
        B.aux = dv.aux;
        B.order = nullptr;
        B.seq_off = doff + s0;
        B.seq_base = arena;
        B.hist = nullptr;
        B.n = (uint32_t)sn;
        B.counter = cnt + 1;
        B.fixed_tab = dv.fixed_lane;
        const size_t per_sm = (sn + dv.sm_count - 1) / dv.sm_count;
        const size_t warps = std::min<size_t>(lane::WARPS, std::max<size_t>(1, (per_sm + 31) / 32)), per_cta = warps * 32;
        k_lane_seg<<<(unsigned)std::max<size_t>(1, std::min((sn + per_cta - 1) / per_cta, (size_t)dv.sm_count)), (unsigned)per_cta,
                     lane::smem_for(warps), st>>>(B);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        k_seg_chain<<<(unsigned)((mc + 127) / 128), 128, 0, st>>>(B, dplan + m0, (uint32_t)mc, dfail + m0);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        // Resolve: the ring kernel (fast for one member, 3 per SM) takes the first ring_k members — the wave's largest
        // — and the global-memory kernel the rest beside it on a second stream; ring_k balances the two under the
        // measured paces (ring: 200 MB/s per member; global: 125 MB/s per member, ~700 GB/s for the whole GPU).
        size_t ring_k = 0;
        {
            const double slots = 3.0 * dv.sm_count, r = 200e6, g = 125e6, gall = 700e9;
            double all = 0, head = 0, best = 1e300;
            for (size_t i = m0; i < m0 + mc; i++) all += (double)hm[segm[i]].out_cap;
            for (size_t k = 0; k <= mc; k++) {
                // k members to the ring kernel (the order is largest first)
                const double first = k ? (double)hm[segm[m0]].out_cap / r : 0.0;
                const double t_ring = std::max(first, head / (slots * r));
                const double next = k < mc ? (double)hm[segm[m0 + k]].out_cap / g : 0.0;
                const double t_glob = std::max(next, (all - head) / gall);
                const double t = std::max(t_ring, t_glob);
                if (t < best) { best = t; ring_k = k; }
                if (k < mc) head += (double)hm[segm[m0 + k]].out_cap;
                if (t_ring > best) break;
            }
        }
        if (!dv.seg_stream2) {
            CK(ctx, cudaStreamCreateWithFlags(&dv.seg_stream2, cudaStreamNonBlocking));
            CK(ctx, cudaEventCreateWithFlags(&dv.seg_fork2, cudaEventDisableTiming));
            CK(ctx, cudaEventCreateWithFlags(&dv.seg_join2, cudaEventDisableTiming));
        }
        CK(ctx, cudaEventRecord(dv.seg_fork2, st));
        CK(ctx, cudaStreamWaitEvent(dv.seg_stream2, dv.seg_fork2, 0));
        k_seg_lz<<<(unsigned)std::max<size_t>(1, std::min((mc + SEGLZ_WARPS - 1) / SEGLZ_WARPS, (size_t)dv.sm_count * 8)), SEGLZ_THREADS, 0, dv.seg_stream2>>>(
            B, dplan + m0, (uint32_t)mc, doff + s0, cnt + 4, (uint32_t)ring_k);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        CK(ctx, cudaEventRecord(dv.seg_join2, dv.seg_stream2));
        if (ring_k) {
            // few members: four warps on each member's ring (270 MB/s per member; one 16 MiB member 83 -> 62 ms); many:
            // one warp each (C4: 293 ms against 338 ms — with every slot busy the four warps get in each other's way)
            if (ring_k <= (size_t)dv.sm_count)
                k_seg_lz_ring_cta<<<(unsigned)std::min(ring_k, (size_t)dv.sm_count * 3), SEGR_THREADS, lz::RING_BYTES, st>>>(B, dplan + m0, (uint32_t)ring_k,
                                                                                                                         doff + s0, cnt + 2);
            else
            k_seg_lz_ring<<<(unsigned)std::min(ring_k, (size_t)dv.sm_count * 3), 32, lz::RING_BYTES, st>>>(B, dplan + m0, (uint32_t)ring_k, doff + s0, cnt + 2);
            ctx->launches++;
            CK(ctx, cudaGetLastError());
        }
        CK(ctx, cudaStreamWaitEvent(st, dv.seg_join2, 0));
        // the members whose chain did not hold: serially, a warp each (slots of the others are empty)
        lane::Batch F;
        F.members = dm;
        F.results = dr;
        F.aux = dv.aux;
        F.order = dfail + m0;
        F.seq_off = doff + s0;                      // (read by no one: k_big keeps its sequences in shared memory)
        F.seq_base = arena;
        F.hist = nullptr;
        F.n = (uint32_t)mc;
        F.counter = cnt + 3;
        F.fixed_tab = dv.fixed_lane;
        const size_t bper_sm = (mc + dv.sm_count - 1) / dv.sm_count;
        const size_t bwarps = std::min<size_t>(BIG_WARPS, std::max<size_t>(1, bper_sm));
        k_big<<<(unsigned)std::max<size_t>(1, std::min((mc + bwarps - 1) / bwarps, (size_t)dv.sm_count * dv.big_occ)), (unsigned)(bwarps * 32),
                LUT_BYTES + bwarps * BIG_PER_WARP, st>>>(F);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        if (trace) {                                // (diagnostic only: waits for the wave) members that took the serial fallback
            std::vector<uint32_t> hf(mc);
            CK(ctx, cudaMemcpyAsync(hf.data(), dfail + m0, mc * 4, cudaMemcpyDeviceToHost, st));
            CK(ctx, cudaStreamSynchronize(st));
            size_t nf = 0, bytes = 0;
            for (size_t i = 0; i < mc; i++)
                if (hf[i] != 0xffffffffu) {
                    nf++;
                    bytes = std::max<size_t>(bytes, hm[hf[i]].out_cap);
                }
            fprintf(stderr, "zsb trace segments: wave %zu: %zu of %zu members fell back to the serial path (largest %zu bytes of output)\n", w, nf, mc, bytes);
        }
    }
    return 0;
}

static int pipe_stream_limit();

// dm/dr: device arrays; hm: host copy of the descriptors; horder: host slot order (NULL = identity)
static int run_device(zsb_ctx *ctx, zsb_dev &dv, size_t n, const zsb_member *dm, zsb_result *dr,
                      const zsb_member *hm, const uint32_t *horder, unsigned flags, cudaStream_t st,
                      const Scratch *sc = nullptr) {
    if (n == 0) return 0;
    if (n > 0xffffff00ull) return ZSB_E_ARG;
    int rc = 0;
    if (!sc && (rc = ensure_aux(ctx, dv, n))) return rc;
    // slot order and sequence offsets travel in one per-call plan buffer
    const size_t plan_bytes = n * (sizeof(uint64_t) + sizeof(uint32_t)) + 64;
    if (!sc && (rc = grow(ctx, dv.d_plan, dv.d_plan_cap, plan_bytes))) return rc;
    uint64_t *dseqoff = (uint64_t *)(sc ? sc->plan : dv.d_plan);
    uint32_t *dorder = (uint32_t *)(dseqoff + n);
    const bool planned = sc && sc->cuts;
    std::vector<uint32_t> own;
    std::vector<uint64_t> off_own;
    std::vector<size_t> cuts_own, cuts_big_own;
    const std::vector<size_t> *cuts = planned ? sc->cuts : &cuts_own, *cuts_big = planned ? sc->cuts_big : &cuts_big_own;
    size_t nbig = planned ? sc->nbig : 0;
    uint64_t big_seqs = planned ? sc->big_seqs : 0;
    uint32_t *counters = sc ? sc->counters : dv.counters;
    size_t ns = n;                         // slots of the lane / group paths (n less the segmented members)
    std::vector<uint32_t> segm;            // members decoded in segments
    uint64_t seg_want = 0, seg_least = 0;  // sequences of arena the segmented part would like / needs for its largest member
    uint32_t segb = 32768;                 // compressed bytes per segment (seg_bytes_for)
    uint64_t normal_need = 0;              // sequences the lane path's part uses (the segmented part's range starts behind them)
    if (!planned) {
        if (!horder) {
            own.resize(n);
            for (size_t i = 0; i < n; i++) own[i] = (uint32_t)i;
            if (ctx->path == 0)          // largest first: lanes finish together, and the split needs it
                sort_largest_first(own.data(), n, [&](uint32_t a) { return (uint64_t)hm[a].in_len + hm[a].out_cap; });
            // members large enough to be decoded in segments (zsb_seg.cuh) leave the slot order
            if (!sc && ctx->path == 0 && ctx->segment && ctx->split_big != 2) {
                size_t keep = 0;
                for (size_t i = 0; i < n; i++) {
                    const zsb_member &x = hm[own[i]];
                    // (a member that does not shrink is stored blocks or as good as: the serial path copies those at memory speed)
                    if (x.in && x.out && x.in_len >= ctx->seg_min && x.in_len <= 0xfffffff0ull && x.out_cap <= 0xfffffff0ull &&
                        !(x.flags & ZSB_M_STORED) && wbits_valid(x.window_bits) && (uint64_t)x.in_len * 16 < (uint64_t)x.out_cap * 15)
                        segm.push_back(own[i]);
                    else own[keep++] = own[i];
                }
                ns = keep;
            }
            horder = own.data();
        }
        if (ctx->path == 0 && ctx->split_big) nbig = ctx->split_big == 2 ? ns : split_big(hm, horder, ns, dv.sm_count);
        if (ns) CK(ctx, cudaMemcpyAsync(dorder, horder, ns * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
        CK(ctx, cudaMemsetAsync(counters, 0, 8 * sizeof(uint32_t), st));
        if (ctx->path != 2) {
            off_own.resize(n);
            if (!segm.empty()) {               // the segmented part shares the arena: as much as it can use, within the budget
                segb = seg_bytes_for(ctx, dv, hm, segm);
                for (uint32_t i : segm) {
                    const uint64_t one = ((hm[i].in_len + segb - 1) / segb) * (uint64_t)seg_q_of(segb);
                    seg_want += one;
                    seg_least = std::max(seg_least, one);
                }
                seg_want = std::max(seg_least, std::min<uint64_t>(seg_want, arena_bytes(ctx, dv) / sizeof(lane::seq_t)));
            }
            uint64_t need = plan_parts(hm, horder, ns, nbig, (sc ? sc->seq_bytes : arena_bytes(ctx, dv)) / sizeof(lane::seq_t),
                                       off_own.data(), cuts_big_own, cuts_own, big_seqs);
            if (!sc && cuts_own.size() > 2 && !ctx->seq_arena_bytes)      // sliced: is more memory free by now?
                need = plan_parts(hm, horder, ns, nbig, arena_bytes(ctx, dv, true) / sizeof(lane::seq_t), off_own.data(),
                                  cuts_big_own, cuts_own, big_seqs);
            need += seg_want;
            if (!sc) {
                // the arena: when the device cannot give it (another process took the memory since the last query),
                // decode in more slices of half the size rather than fail
                uint64_t arena = (uint64_t)arena_bytes(ctx, dv) / sizeof(lane::seq_t);
                for (;;) {
                    if ((size_t)need * sizeof(lane::seq_t) <= dv.d_seq_cap) break;
                    if (dv.d_seq) CK(ctx, cudaFree(dv.d_seq));
                    dv.d_seq = nullptr;
                    dv.d_seq_cap = 0;
                    const size_t want = (size_t)need * sizeof(lane::seq_t) + (size_t)need + 4096;
                    if (cudaMalloc(&dv.d_seq, want) == cudaSuccess) {
                        dv.d_seq_cap = want;
                        break;
                    }
                    cudaGetLastError();
                    uint64_t largest = 0;
                    for (size_t s = nbig; s < ns; s++) largest = std::max(largest, seq_bound(hm[horder[s]]));
                    if (seg_want > seg_least) {                        // the segmented part can do with less (more waves)
                        seg_want = std::max(seg_least, seg_want / 2);
                        need = plan_parts(hm, horder, ns, nbig, arena, off_own.data(), cuts_big_own, cuts_own, big_seqs) + seg_want;
                        continue;
                    }
                    if (arena <= largest || ctx->seq_arena_bytes) {
                        snprintf(ctx->err, sizeof ctx->err, "out of device memory for %zu bytes of sequence scratch", want);
                        return ZSB_E_CUDA;
                    }
                    arena = std::max(largest, std::min(arena, need) / 2);
                    dv.arena_auto = (size_t)arena * sizeof(lane::seq_t);
                    need = plan_parts(hm, horder, ns, nbig, arena, off_own.data(), cuts_big_own, cuts_own, big_seqs) + seg_want;
                }
            }
            normal_need = need - seg_want;
            // (a pageable source is staged by the runtime before cudaMemcpyAsync returns, so off_own may go)
            if (ns) CK(ctx, cudaMemcpyAsync(dseqoff, off_own.data(), ns * sizeof(uint64_t), cudaMemcpyHostToDevice, st));
        }
    }
    {
        static const bool trace = getenv("ZSB_TRACE") != nullptr;
        if (trace && !planned)
            fprintf(stderr, "zsb trace decode: %zu members, %zu to k_big (%zu slices, %llu sequences), rest %zu slices\n", n,
                    nbig, cuts_big->size() ? cuts_big->size() - 1 : 0, (unsigned long long)big_seqs,
                    cuts->size() ? cuts->size() - 1 : 0);
    }
    if (ctx->path == 2) {
        if ((rc = run_group(ctx, dv, n, dm, dr, dorder, st, sc, ctx->lanes))) return rc;
    } else {
        // the segmented members on a stream of their own, beside the lane path's part (their range of the arena is behind it)
        if (!segm.empty()) {
            if (!dv.seg_stream) {
                CK(ctx, cudaStreamCreateWithFlags(&dv.seg_stream, cudaStreamNonBlocking));
                CK(ctx, cudaEventCreateWithFlags(&dv.seg_fork, cudaEventDisableTiming));
                CK(ctx, cudaEventCreateWithFlags(&dv.seg_join, cudaEventDisableTiming));
            }
            CK(ctx, cudaEventRecord(dv.seg_fork, st));
            CK(ctx, cudaStreamWaitEvent(dv.seg_stream, dv.seg_fork, 0));
            cudaEvent_t *se = nullptr;
            if (ctx->profile) {                       // the segmented part's own duration (zsb_profile_read_seg)
                if (dv.seg_used + 2 > dv.seg_ev.size()) {
                    const size_t old = dv.seg_ev.size();
                    dv.seg_ev.resize(old + 32);
                    for (size_t i = old; i < dv.seg_ev.size(); i++) CK(ctx, cudaEventCreate(&dv.seg_ev[i]));
                }
                se = &dv.seg_ev[dv.seg_used];
                dv.seg_used += 2;
                CK(ctx, cudaEventRecord(se[0], dv.seg_stream));
            }
            if ((rc = run_segmented(ctx, dv, dm, dr, hm, segm, dv.seg_stream, (lane::seq_t *)dv.d_seq + normal_need, seg_want, segb))) return rc;
            if (se) CK(ctx, cudaEventRecord(se[1], dv.seg_stream));
            CK(ctx, cudaEventRecord(dv.seg_join, dv.seg_stream));
        }
        LanePart P;
        P.aux = sc ? sc->aux : dv.aux;
        P.small_ctas = sc && sc->small_ctas;
        lane::seq_t *seq0 = (lane::seq_t *)(sc ? sc->seq : dv.d_seq);
        // the large members, a warp each, on the group stream beside the rest on the caller's (fork / join by events)
        const bool fork = nbig && nbig < ns;
        if (nbig) {
            P.dorder = dorder; P.dseqoff = dseqoff; P.cuts = cuts_big; P.seq_base = seq0;
            P.counters = counters + 4; P.big = true; P.timed = false;
            P.st = fork ? dv.group_stream : st;
            if (fork) {
                CK(ctx, cudaEventRecord(dv.ev_fork, st));
                CK(ctx, cudaStreamWaitEvent(dv.group_stream, dv.ev_fork, 0));
            }
            if ((rc = launch_lane(ctx, dv, dm, dr, P))) return rc;
            if (fork) CK(ctx, cudaEventRecord(dv.ev_join, dv.group_stream));
        }
        if (nbig < ns) {
            P.dorder = dorder + nbig; P.dseqoff = dseqoff + nbig; P.cuts = cuts; P.seq_base = seq0;
            P.counters = counters; P.big = false; P.timed = true; P.st = st;
            if ((rc = launch_lane(ctx, dv, dm, dr, P))) return rc;
        }
        if (fork) CK(ctx, cudaStreamWaitEvent(st, dv.ev_join, 0));
        if (!segm.empty()) CK(ctx, cudaStreamWaitEvent(st, dv.seg_join, 0));
    }
    if (!(flags & ZSB_RAW_ONLY)) {
        unsigned grid = (unsigned)std::min<size_t>(n, (size_t)dv.sm_count * 8);
        cudaEvent_t *ce = nullptr;
        if (ctx->profile) {                           // k_check's own duration (zsb_profile_read_check)
            if (dv.chk_used + 2 > dv.chk_ev.size()) {
                const size_t old = dv.chk_ev.size();
                dv.chk_ev.resize(old + 32);
                for (size_t i = old; i < dv.chk_ev.size(); i++) CK(ctx, cudaEventCreate(&dv.chk_ev[i]));
            }
            ce = &dv.chk_ev[dv.chk_used];
            dv.chk_used += 2;
            CK(ctx, cudaEventRecord(ce[0], st));
        }
        // wrapped members of 2 MiB of output or more are checked in parts of 1 MiB, a CTA each, and k_check combines them
        const uint32_t *d_first = nullptr;
        const ChkPart *d_parts = nullptr;
        if (!sc) {
            std::vector<ChkPart> hp;
            std::vector<uint32_t> first;
            for (size_t i = 0; i < n; i++) {
                const zsb_member &x = hm[i];
                if (x.window_bits > 0 && x.out_cap >= 2ull * CHK_PART && x.out_cap <= 0xfffffff0ull && !(x.flags & ZSB_M_STORED)) {
                    if (first.empty()) first.assign(n, 0xffffffffu);
                    first[i] = (uint32_t)hp.size();
                    const uint32_t np = (uint32_t)((x.out_cap + CHK_PART - 1) / CHK_PART);
                    for (uint32_t k = 0; k < np; k++) hp.push_back(ChkPart{(uint32_t)i, k, 0u, 0u});
                }
            }
            if (!hp.empty()) {
                const size_t bytes = n * sizeof(uint32_t) + 256 + hp.size() * sizeof(ChkPart);
                if ((rc = grow(ctx, dv.d_chkparts, dv.d_chkparts_cap, bytes))) return rc;
                uint32_t *df = (uint32_t *)dv.d_chkparts;
                ChkPart *dp = (ChkPart *)(dv.d_chkparts + ((n * sizeof(uint32_t) + 255) & ~(size_t)255));
                CK(ctx, cudaMemcpyAsync(df, first.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
                CK(ctx, cudaMemcpyAsync(dp, hp.data(), hp.size() * sizeof(ChkPart), cudaMemcpyHostToDevice, st));
                k_check_parts<<<(unsigned)std::min<size_t>(hp.size(), (size_t)dv.sm_count * 8), CHK_THREADS, 0, st>>>(dm, dr, dv.aux, dp, (uint32_t)hp.size());
                ctx->launches++;
                CK(ctx, cudaGetLastError());
                d_first = df;
                d_parts = dp;
            }
        }
        k_check<<<grid, CHK_THREADS, 0, st>>>(dm, dr, sc ? sc->aux : dv.aux, (uint32_t)n, (sc ? sc->counters : dv.counters) + 1, d_first, d_parts);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
        if (ce) CK(ctx, cudaEventRecord(ce[1], st));
    }
    return 0;
}

// ---- contexts ----------------------------------------------------------------------------------
static int init_device(zsb_ctx *ctx, zsb_dev &dv, int device) {
    dv.device = device;
    CK(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(ctx, cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        snprintf(ctx->err, sizeof ctx->err, "device %d is sm_%d%d; libzsb is built for sm_100a only", device,
                 prop.major, prop.minor);
        return ZSB_E_CUDA;
    }
    dv.sm_count = prop.multiProcessorCount;
    CK(ctx, cudaStreamCreateWithFlags(&dv.stream, cudaStreamNonBlocking));
    CK(ctx, cudaStreamCreateWithFlags(&dv.copy_stream, cudaStreamNonBlocking));
    CK(ctx, cudaStreamCreateWithFlags(&dv.back_stream, cudaStreamNonBlocking));
    // The decode streams of pipelined host batches.  Streams share the device's hardware work queues
    // (CUDA_DEVICE_MAX_CONNECTIONS, 8 unless the process raised it before CUDA started), and two streams
    // on one queue wait for each other.  Created back to back, these eight streams get eight different
    // queues; with more queues available a batch may use more decode streams (pipe_stream_limit).
    dv.pipe_streams.push_back(dv.stream);
    for (int i = 1; i < BASE_PIPE_STREAMS; i++) {
        cudaStream_t st;
        CK(ctx, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        dv.pipe_streams.push_back(st);
    }
    CK(ctx, cudaStreamCreateWithFlags(&dv.group_stream, cudaStreamNonBlocking));
    CK(ctx, cudaEventCreateWithFlags(&dv.ev_fork, cudaEventDisableTiming));
    CK(ctx, cudaEventCreateWithFlags(&dv.ev_join, cudaEventDisableTiming));
    CK(ctx, cudaMalloc(&dv.fixed, STREAM_BYTES));
    CK(ctx, cudaMalloc(&dv.fixed_lane, lane::SLOT));
    CK(ctx, cudaMalloc(&dv.counters, 8 * MAX_CHUNKS_DEV * sizeof(uint32_t)));
    CrcConsts cc;
    cc.x2n[0] = 0x40000000u;
    for (int k = 1; k < 40; k++) cc.x2n[k] = gf2_mul(cc.x2n[k - 1], cc.x2n[k - 1]);
    auto xpow = [&](uint64_t nbits) {
        uint32_t r = 0x80000000u;
        for (int k = 0; nbits; k++, nbits >>= 1)
            if (nbits & 1) r = gf2_mul(r, cc.x2n[k]);
        return r;
    };
    cc.tile_skip = xpow(8ull * (CHK_TILE - CHK_SPAN));
    for (int k = 0; k < 8; k++) cc.span_pow[k] = xpow((8ull * CHK_SPAN) << k);
    CK(ctx, cudaMemcpyToSymbol(c_crc, &cc, sizeof cc));
    k_build_fixed<<<1, 128, 0, dv.stream>>>(dv.fixed);
    CK(ctx, cudaGetLastError());
    k_build_fixed_lane<<<1, 32, 0, dv.stream>>>(dv.fixed_lane);
    CK(ctx, cudaGetLastError());
    CK(ctx, cudaStreamSynchronize(dv.stream));
    return 0;
}

static void free_device(zsb_dev &dv) {
    cudaSetDevice(dv.device);
    if (dv.stream) cudaStreamDestroy(dv.stream);
    if (dv.copy_stream) cudaStreamDestroy(dv.copy_stream);
    if (dv.back_stream) cudaStreamDestroy(dv.back_stream);
    if (dv.group_stream) cudaStreamDestroy(dv.group_stream);
    if (dv.seg_stream) cudaStreamDestroy(dv.seg_stream);
    if (dv.seg_stream2) cudaStreamDestroy(dv.seg_stream2);
    if (dv.seg_fork2) cudaEventDestroy(dv.seg_fork2);
    if (dv.seg_join2) cudaEventDestroy(dv.seg_join2);
    if (dv.seg_fork) cudaEventDestroy(dv.seg_fork);
    if (dv.seg_join) cudaEventDestroy(dv.seg_join);
    if (dv.ev_fork) cudaEventDestroy(dv.ev_fork);
    if (dv.ev_join) cudaEventDestroy(dv.ev_join);
    for (size_t i = 1; i < dv.pipe_streams.size(); i++) cudaStreamDestroy(dv.pipe_streams[i]);   // [0] is dv.stream
    dv.pipe_streams.clear();
    for (cudaEvent_t e : dv.pipe_ev) cudaEventDestroy(e);
    dv.pipe_ev.clear();
    cudaFree(dv.fixed);
    cudaFree(dv.fixed_lane);
    cudaFree(dv.d_seq);
    cudaFree(dv.d_plan);
    cudaFree(dv.counters);
    cudaFree(dv.aux);
    cudaFree(dv.d_in);
    cudaFree(dv.d_out);
    cudaFree(dv.d_desc);
    cudaFree(dv.d_df);
    cudaFree(dv.d_chkparts);
    cudaFree(dv.d_segbuf);
    if (dv.h_desc) cudaFreeHost(dv.h_desc);
    for (cudaEvent_t e : dv.prof_ev) cudaEventDestroy(e);
    dv.prof_ev.clear();
    for (cudaEvent_t e : dv.chk_ev) cudaEventDestroy(e);
    dv.chk_ev.clear();
    for (cudaEvent_t e : dv.seg_ev) cudaEventDestroy(e);
    dv.seg_ev.clear();
}

extern "C" zsb_ctx *zsb_create(const int *devices, int ndev) {
    g_create_err[0] = 0;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        snprintf(g_create_err, sizeof g_create_err, "no CUDA device: %s (libzsb has no CPU path)",
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return nullptr;
    }
    zsb_ctx *ctx = new zsb_ctx();
    int cur = 0;
    cudaGetDevice(&cur);
    std::vector<int> devs;
    if (!devices || ndev <= 0) devs.push_back(cur);
    else devs.assign(devices, devices + ndev);
    ctx->devs.resize(devs.size());
    for (size_t i = 0; i < devs.size(); i++) {
        if (devs[i] < 0 || devs[i] >= count || init_device(ctx, ctx->devs[i], devs[i]) != 0) {
            if (!ctx->err[0]) snprintf(ctx->err, sizeof ctx->err, "bad device ordinal %d", devs[i]);
            snprintf(g_create_err, sizeof g_create_err, "%s", ctx->err);
            for (size_t k = 0; k <= i; k++) free_device(ctx->devs[k]);
            delete ctx;
            cudaSetDevice(cur);
            return nullptr;
        }
    }
    cudaSetDevice(cur);
    return ctx;
}

extern "C" void zsb_destroy(zsb_ctx *ctx) {
    if (!ctx) return;
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto &dv : ctx->devs) free_device(dv);
    cudaSetDevice(cur);
    delete ctx;
}

extern "C" int zsb_set_option(zsb_ctx *ctx, const char *name, long value) {
    if (!ctx || !name) return ZSB_E_ARG;
    if (!strcmp(name, "lanes_per_member")) {
        if (value != 1 && value != 2 && value != 4 && value != 8 && value != 16 && value != 32) return ZSB_E_ARG;
        ctx->lanes = (int)value;
        return 0;
    }
    if (!strcmp(name, "slab_bytes")) {
        if (value < (1 << 20)) return ZSB_E_ARG;
        ctx->slab_bytes = (size_t)value;
        return 0;
    }
    if (!strcmp(name, "profile")) {
        ctx->profile = value != 0;
        return 0;
    }
    if (!strcmp(name, "chunks")) {
        if (value < 0 || value > 64) return ZSB_E_ARG;
        ctx->chunks = (int)value;
        return 0;
    }
    if (!strcmp(name, "split_big")) {         // 1 (default): members far larger than the rest of their batch go to k_big
        if (value < 0 || value > 2) return ZSB_E_ARG;
        ctx->split_big = (int)value;          // 2: every member (tests of k_big)
        return 0;
    }
    if (!strcmp(name, "small_ctas")) {        // diagnostic: k_lane in the small CTAs of the pipelined host path
        ctx->force_small_ctas = value != 0;
        return 0;
    }
    if (!strcmp(name, "path")) {
        if (value < 0 || value > 2) return ZSB_E_ARG;
        ctx->path = (int)value;
        return 0;
    }
    if (!strcmp(name, "segment")) {               // 1 (default): members of seg_min compressed bytes or more are decoded in segments
        ctx->segment = value != 0;
        return 0;
    }
    if (!strcmp(name, "seg_bytes")) {             // compressed bytes between the block starts looked for
        if (value != 0 && (value < 256 || value > (1 << 24))) return ZSB_E_ARG;      // 0 = choose from the batch
        ctx->seg_bytes = (size_t)value;
        return 0;
    }
    if (!strcmp(name, "seg_min")) {               // smallest member (compressed bytes) that is segmented
        if (value < 64) return ZSB_E_ARG;
        ctx->seg_min = (size_t)value;
        return 0;
    }
    if (!strcmp(name, "seq_arena_bytes")) {       // 0 = choose from the device's free memory
        if (value != 0 && value < 4096) return ZSB_E_ARG;
        ctx->seq_arena_bytes = (size_t)value;
        return 0;
    }
    return ZSB_E_ARG;
}
extern "C" long zsb_get_option(const zsb_ctx *ctx, const char *name) {
    if (!ctx || !name) return ZSB_E_ARG;
    if (!strcmp(name, "lanes_per_member")) return ctx->lanes;
    if (!strcmp(name, "slab_bytes")) return (long)ctx->slab_bytes;
    if (!strcmp(name, "path")) return ctx->path;
    if (!strcmp(name, "chunks")) return ctx->chunks;
    if (!strcmp(name, "segment")) return ctx->segment;
    if (!strcmp(name, "seg_bytes")) return (long)ctx->seg_bytes;
    if (!strcmp(name, "seg_min")) return (long)ctx->seg_min;
    if (!strcmp(name, "seq_arena_bytes")) return (long)ctx->seq_arena_bytes;
    if (!strcmp(name, "sm_count")) return ctx->devs.empty() ? 0 : ctx->devs[0].sm_count;
    return ZSB_E_ARG;
}

// Sums the event-timed durations (ms) of the k_lane and k_lz launches recorded on device dev_index since
// the last read (option "profile" = 1); waits for the last recorded event.  Returns the number of
// launch pairs, or a negative error.
extern "C" long zsb_profile_read(zsb_ctx *ctx, int dev_index, double *lane_ms, double *lz_ms) {
    if (!ctx || dev_index < 0 || dev_index >= (int)ctx->devs.size()) return ZSB_E_ARG;
    zsb_dev &dv = ctx->devs[dev_index];
    double a = 0, b = 0;
    long pairs = 0;
    for (size_t i = 0; i + 3 <= dv.prof_used; i += 3) {
        float t0 = 0, t1 = 0;
        CK(ctx, cudaEventSynchronize(dv.prof_ev[i + 2]));
        CK(ctx, cudaEventElapsedTime(&t0, dv.prof_ev[i], dv.prof_ev[i + 1]));
        CK(ctx, cudaEventElapsedTime(&t1, dv.prof_ev[i + 1], dv.prof_ev[i + 2]));
        a += t0;
        b += t1;
        pairs++;
    }
    dv.prof_used = 0;
    if (lane_ms) *lane_ms = a;
    if (lz_ms) *lz_ms = b;
    return pairs;
}

// Sum of the event-timed durations (ms) of the k_check launches recorded since the last read (option "profile").
extern "C" long zsb_profile_read_check(zsb_ctx *ctx, int dev_index, double *check_ms) {
    if (!ctx || dev_index < 0 || dev_index >= (int)ctx->devs.size()) return ZSB_E_ARG;
    zsb_dev &dv = ctx->devs[dev_index];
    double a = 0;
    long k = 0;
    for (size_t i = 0; i + 2 <= dv.chk_used; i += 2) {
        float t = 0;
        CK(ctx, cudaEventSynchronize(dv.chk_ev[i + 1]));
        CK(ctx, cudaEventElapsedTime(&t, dv.chk_ev[i], dv.chk_ev[i + 1]));
        a += t;
        k++;
    }
    dv.chk_used = 0;
    if (check_ms) *check_ms = a;
    return k;
}

// The same for the segmented part of the batches (find + k_lane_seg + chain + resolve + fallback, on its own stream).
extern "C" long zsb_profile_read_seg(zsb_ctx *ctx, int dev_index, double *seg_ms) {
    if (!ctx || dev_index < 0 || dev_index >= (int)ctx->devs.size()) return ZSB_E_ARG;
    zsb_dev &dv = ctx->devs[dev_index];
    double a = 0;
    long k = 0;
    for (size_t i = 0; i + 2 <= dv.seg_used; i += 2) {
        float t = 0;
        CK(ctx, cudaEventSynchronize(dv.seg_ev[i + 1]));
        CK(ctx, cudaEventElapsedTime(&t, dv.seg_ev[i], dv.seg_ev[i + 1]));
        a += t;
        k++;
    }
    dv.seg_used = 0;
    if (seg_ms) *seg_ms = a;
    return k;
}

extern "C" int zsb_inflate_batch_device(zsb_ctx *ctx, int dev_index, size_t n, const zsb_member *d_members,
                                        const zsb_member *h_members, zsb_result *d_results,
                                        const uint32_t *h_order, unsigned flags, void *cuda_stream) {
    if (!ctx || dev_index < 0 || dev_index >= (int)ctx->devs.size() || (n && (!d_members || !d_results || !h_members)))
        return ZSB_E_ARG;
    zsb_dev &dv = ctx->devs[dev_index];
    int cur = 0;
    cudaGetDevice(&cur);
    if (cur != dv.device) CK(ctx, cudaSetDevice(dv.device));
    // what the other batch paths check: sizes the 32-bit decoders cannot hold are an argument error here (the
    // device descriptors are the caller's, so they cannot be rewritten to answer Z_STREAM_ERROR per member), and
    // any member that asks for a copy or a CRC (ZSB_M_*) needs k_check whatever the caller said about wrappers
    for (size_t i = 0; i < n; i++) {
        if (h_members[i].in_len > 0xfffffff0ull || h_members[i].out_cap > 0xfffffff0ull) {
            snprintf(ctx->err, sizeof ctx->err, "member %zu: in_len / out_cap above 0xfffffff0", i);
            return ZSB_E_ARG;
        }
        if (h_members[i].flags) flags = (flags & ~ZSB_RAW_ONLY) | ZSB_HAS_WRAPPED;
    }
    ctx->launches = 0;
    int rc = run_device(ctx, dv, n, d_members, d_results, h_members, h_order, flags, (cudaStream_t)cuda_stream);
    if (cur != dv.device) cudaSetDevice(cur);
    return rc;
}

// ---- host-pointer batches: shard, stage, decode, return ------------------------------------------
template <typename T> static int grow(zsb_ctx *ctx, T *&p, size_t &cap, size_t need) {
    if (need <= cap) return 0;
    if (p) CK(ctx, cudaFree(p));
    p = nullptr;
    cap = 0;
    size_t want = need + need / 8 + 4096;
    CK(ctx, cudaMalloc(&p, want));
    cap = want;
    return 0;
}

// A run = consecutive members whose host buffers follow each other closely enough to be moved with one
// copy; the device arena mirrors the host layout inside a run.  Inputs may be merged across small gaps
// (a gap shorter than a page only touches pages that hold member bytes, so reading it is safe);
// outputs only when exactly adjacent, because the gap is not ours to write.
struct Run {
    uintptr_t host;
    size_t dev_off, len;
};
struct Shard {
    std::vector<uint32_t> idx;
    std::vector<size_t> in_off, out_off;
    std::vector<Run> in_runs, out_runs;
    size_t in_span = 0, out_span = 0;
};

static size_t layout_runs(const zsb_member *m, const std::vector<uint32_t> &idx, bool out, size_t gap_limit,
                          std::vector<size_t> &off, std::vector<Run> &runs) {
    const size_t k = idx.size();
    off.assign(k, 0);
    runs.clear();
    size_t total = 0;
    bool open = false;
    uintptr_t host_base = 0, end = 0;
    size_t dev_base = 0;
    for (size_t j = 0; j < k; j++) {
        const zsb_member &x = m[idx[j]];
        uintptr_t p = out ? (uintptr_t)x.out : (uintptr_t)x.in;
        size_t len = out ? x.out_cap : x.in_len;
        if (len == 0 || p == 0) continue;
        if (x.in_len > 0xfffffff0ull || x.out_cap > 0xfffffff0ull) continue;   // answered Z_STREAM_ERROR, never staged
        if (open && p >= end && p - end <= gap_limit) {
            off[j] = dev_base + (p - host_base);
        } else {
            if (open) runs.push_back({host_base, dev_base, (size_t)(end - host_base)});
            dev_base = (total + 255) & ~(size_t)255;
            dev_base += p & 15;               // keep the host's 16-byte phase
            host_base = p;
            off[j] = dev_base;
            open = true;
        }
        end = p + len;
        total = dev_base + (end - host_base);
    }
    if (open) runs.push_back({host_base, dev_base, (size_t)(end - host_base)});
    return total;
}

// Host-pointer batches are pipelined: a shard is cut into chunks of consecutive members and chunk c's
// host->device copies (copy_stream), decode (stream) and device->host copies (back_stream) overlap
// with the other chunks', so that for large batches the call costs about the slower PCIe direction.
constexpr int MAX_CHUNKS = 32;
// decode streams a pipelined host batch may use: the hardware queues left beside the two copy streams
static int pipe_stream_limit() {
    const char *e = getenv("CUDA_DEVICE_MAX_CONNECTIONS");
    const int conn = e ? atoi(e) : 8;
    return std::max(BASE_PIPE_STREAMS, std::min(conn, 32) - 2);
}
struct Chunk {
    size_t first = 0, count = 0;       // range of the shard's member list
    std::vector<uint32_t> idx;         // member indices of the range
    std::vector<size_t> in_off, out_off;
    std::vector<Run> in_runs, out_runs;
    size_t in_base = 0, out_base = 0;  // where the chunk's spans start in the device arenas
};

static int choose_chunks(const zsb_ctx *ctx, const zsb_dev &dv, const zsb_member *m, const Shard &sh) {
    if (ctx->chunks > 0) return (int)std::min<size_t>(ctx->chunks, std::max<size_t>(sh.idx.size(), 1));
    size_t bytes = 0;
    for (uint32_t i : sh.idx) bytes += m[i].in_len + m[i].out_cap;
    // A member decodes at a fixed pace however few are in flight (one lane each), so the chunks' decodes
    // must overlap each other: every chunk has its own stream and k_lane runs in small CTAs that share the
    // SMs (Scratch::small_ctas).  A chunk is then only bounded below by the cost of its ~12 API calls.
    // Every stream decodes its chunks one after the other, so at least two rounds of chunks keep the copy
    // engines busy while a round decodes.
    (void)dv;
    const size_t by_members = sh.idx.size() / 2048;
    const size_t by_bytes = bytes / (128u << 20);
    const size_t by_streams = (size_t)std::min(MAX_CHUNKS, 2 * pipe_stream_limit());
    return (int)std::max<size_t>(1, std::min(by_streams, std::min(by_members, by_bytes)));
}

static int batch_host(zsb_ctx *ctx, size_t n, const zsb_member *m, zsb_result *r) {
    const auto t_call = std::chrono::steady_clock::now();
    const int nd = (int)ctx->devs.size();
    std::vector<int32_t> shard_of(n, 0);
    if (nd > 1) {
        std::vector<uint64_t> cost(n);
        for (size_t i = 0; i < n; i++) cost[i] = (uint64_t)m[i].in_len + m[i].out_cap;
        zsb_plan_shards(n, cost.data(), nd, shard_of.data(), nullptr);
    }
    std::vector<Shard> shards(nd);
    for (size_t i = 0; i < n; i++) shards[shard_of[i]].idx.push_back((uint32_t)i);
    bool raw_only = true;
    for (size_t i = 0; i < n; i++) raw_only &= m[i].window_bits < 0 && m[i].flags == 0;
    const unsigned flags = raw_only ? ZSB_RAW_ONLY : ZSB_HAS_WRAPPED;
    ctx->launches = 0;

    // enqueue everything on every device, then drain
    for (int s = 0; s < nd; s++) {
        Shard &sh = shards[s];
        zsb_dev &dv = ctx->devs[s];
        const size_t k = sh.idx.size();
        if (!k) continue;
        CK(ctx, cudaSetDevice(dv.device));
        const int K = choose_chunks(ctx, dv, m, sh);
        std::vector<Chunk> chunks(K);
        std::vector<uint64_t> seq_need(K);
        size_t in_total = 0, out_total = 0;
        const uint64_t arena_seqs = arena_bytes(ctx, dv) / sizeof(lane::seq_t);
        for (int c = 0; c < K; c++) {
            Chunk &ch = chunks[c];
            ch.first = k * c / K;
            ch.count = k * (c + 1) / K - ch.first;
            ch.idx.assign(sh.idx.begin() + ch.first, sh.idx.begin() + ch.first + ch.count);
            ch.in_base = in_total;
            ch.out_base = out_total;
            in_total += (layout_runs(m, ch.idx, false, 4095, ch.in_off, ch.in_runs) + 511) & ~(size_t)255;
            out_total += (layout_runs(m, ch.idx, true, 0, ch.out_off, ch.out_runs) + 511) & ~(size_t)255;
            // a region holds the chunk's whole bound, or (sliced by plan_parts) the arena option — but never less than
            // the largest single member's bound: lane_plan places such a member alone in a slice at offset 0
            uint64_t sq = 0, largest = 0;
            for (uint32_t i : ch.idx) {
                const uint64_t b = seq_bound(m[i]);
                sq += b;
                largest = std::max(largest, b);
            }
            seq_need[c] = std::max(std::min<uint64_t>(sq, arena_seqs), largest);
        }
        // Chunk c decodes on stream c % NS with sequence region c % NS; the regions together stay inside
        // the arena (fewer streams if they would not).
        int NS = std::min(K, pipe_stream_limit());
        std::vector<uint64_t> region;
        for (;; NS = (NS + 1) / 2) {
            region.assign(NS, 0);
            for (int c = 0; c < K; c++) region[c % NS] = std::max(region[c % NS], seq_need[c]);
            uint64_t sum = 0;
            for (uint64_t v : region) sum += v;
            if (sum <= std::max<uint64_t>(arena_seqs, region[0]) || NS == 1) break;
        }
        std::vector<uint64_t> region_off(NS, 0);
        uint64_t seq_total = 0;
        for (int q = 0; q < NS; q++) {
            region_off[q] = seq_total;
            seq_total += (region[q] + 31) & ~(uint64_t)31;
        }
        int rc;
        if ((rc = grow(ctx, dv.d_in, dv.d_in_cap, in_total + 256))) return rc;
        if ((rc = grow(ctx, dv.d_out, dv.d_out_cap, out_total + 256))) return rc;
        const size_t desc_bytes = k * (sizeof(zsb_member) + sizeof(zsb_result)) + 64;
        if ((rc = grow(ctx, dv.d_desc, dv.d_desc_cap, desc_bytes))) return rc;
        // scratch of the decode kernels, allocated before anything is enqueued: growing it between chunks
        // would free memory under the running pipeline
        if ((rc = ensure_aux(ctx, dv, k))) return rc;
        if ((rc = grow(ctx, dv.d_plan, dv.d_plan_cap, k * (sizeof(uint64_t) + sizeof(uint32_t)) + (size_t)K * 64))) return rc;
        if (ctx->path != 2 && (rc = grow(ctx, dv.d_seq, dv.d_seq_cap, (size_t)seq_total * sizeof(lane::seq_t)))) return rc;
        const size_t hdesc_bytes = k * (sizeof(zsb_member) + sizeof(zsb_result) + sizeof(uint64_t) + sizeof(uint32_t)) + 64;
        if (hdesc_bytes > dv.h_desc_cap) {
            if (dv.h_desc) CK(ctx, cudaFreeHost(dv.h_desc));
            dv.h_desc = nullptr;
            CK(ctx, cudaHostAlloc(&dv.h_desc, hdesc_bytes + hdesc_bytes / 8, cudaHostAllocDefault));
            dv.h_desc_cap = hdesc_bytes + hdesc_bytes / 8;
        }
        while (dv.pipe_ev.size() < (size_t)2 * K) {
            cudaEvent_t e;
            CK(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            dv.pipe_ev.push_back(e);
        }
        while (dv.pipe_streams.size() < (size_t)NS) {
            cudaStream_t st;
            CK(ctx, cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            dv.pipe_streams.push_back(st);
        }
        // ZSB_TRACE=1: per-chunk timeline (inputs arrived, decoded, outputs back; ms since the first copy) on stderr
        static const bool trace = getenv("ZSB_TRACE") != nullptr;
        std::vector<cudaEvent_t> tev;
        if (trace) {
            fprintf(stderr, "zsb trace host: planning done %.2f ms after the call started\n",
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_call).count());
            tev.resize(3 * K + 1);
            for (auto &e : tev) CK(ctx, cudaEventCreate(&e));
            CK(ctx, cudaEventRecord(tev[3 * K], dv.copy_stream));
        }
        zsb_member *hm = (zsb_member *)dv.h_desc;
        zsb_result *hr = (zsb_result *)(hm + k);
        uint64_t *hs = (uint64_t *)(hr + k);
        uint32_t *ho = (uint32_t *)(hs + k);
        zsb_member *dm = (zsb_member *)dv.d_desc;
        zsb_result *dr = (zsb_result *)(dm + k);
        // the three stages of every chunk: inputs on copy_stream, decode on the chunk's stream, results and
        // outputs on back_stream
        // Which members are large is decided for the shard as a whole: the chunks of a pipeline overlap, so what
        // counts is the batch's time, not a chunk's own (k_big would finish a lone chunk sooner, but the whole
        // GPU decodes only ~26 bytes/ns that way).  The sort is skipped when the largest member's own lane time
        // is within the batch's.
        double big_threshold = 1e300;
        if (ctx->path == 0 && ctx->split_big == 1 && K > 1) {
            double total = 0, most = 0;
            for (uint32_t i : sh.idx) {
                const double c = (double)m[i].in_len + (double)m[i].out_cap;
                total += c;
                most = std::max(most, c);
            }
            if (most * 150.0 > total / 190.0) {
                std::vector<uint32_t> ord(sh.idx);
                sort_largest_first(ord.data(), ord.size(), [&](uint32_t a) { return (uint64_t)m[a].in_len + m[a].out_cap; });
                const size_t nb = split_big(m, ord.data(), ord.size(), dv.sm_count);
                if (nb) big_threshold = (double)m[ord[nb - 1]].in_len + (double)m[ord[nb - 1]].out_cap;
            }
        }
        size_t plan_off = 0;
        std::vector<std::vector<size_t>> cuts_of(K), cuts_big_of(K);
        CK(ctx, cudaMemsetAsync(dv.counters, 0, 8 * (size_t)K * sizeof(uint32_t), dv.copy_stream));
        for (int c = 0; c < K; c++) {
            Chunk &ch = chunks[c];
            zsb_member *hmc = hm + ch.first;
            uint32_t *hoc = ho + ch.first;
            for (size_t j = 0; j < ch.count; j++) {
                const zsb_member &x = m[ch.idx[j]];
                hmc[j] = x;
                hmc[j].in = x.in ? dv.d_in + ch.in_base + ch.in_off[j] : nullptr;
                hmc[j].out = x.out ? dv.d_out + ch.out_base + ch.out_off[j] : nullptr;
                if (x.in_len > 0xfffffff0ull || x.out_cap > 0xfffffff0ull) hmc[j].window_bits = 1000;  // -> Z_STREAM_ERROR
                hoc[j] = (uint32_t)j;
            }
            // largest first: the decoder's lanes then finish together
            sort_largest_first(hoc, ch.count, [&](uint32_t a) { return (uint64_t)hmc[a].in_len + hmc[a].out_cap; });
            cudaEvent_t ev_in = dv.pipe_ev[2 * c], ev_dec = dv.pipe_ev[2 * c + 1];
            cudaStream_t dst = dv.pipe_streams[c % NS];
            CK(ctx, cudaMemcpyAsync(dm + ch.first, hm + ch.first, ch.count * sizeof(zsb_member), cudaMemcpyHostToDevice,
                                    dv.copy_stream));
            for (const Run &rn : ch.in_runs)
                CK(ctx, cudaMemcpyAsync(dv.d_in + ch.in_base + rn.dev_off, (const void *)rn.host, rn.len,
                                        cudaMemcpyHostToDevice, dv.copy_stream));
            Scratch sc;
            sc.aux = dv.aux + ch.first;
            sc.plan = dv.d_plan + plan_off;
            plan_off += (ch.count * (sizeof(uint64_t) + sizeof(uint32_t)) + 63) & ~(size_t)63;
            sc.seq = dv.d_seq + region_off[c % NS] * sizeof(lane::seq_t);
            sc.seq_bytes = (size_t)region[c % NS] * sizeof(lane::seq_t);
            sc.counters = dv.counters + 8 * c;
            sc.small_ctas = K > 1;
            // the chunk's plan travels with its inputs: slot order, and for the lane path the sequence offsets
            sc.nbig = 0;
            if (ctx->path == 0 && ctx->split_big == 2) sc.nbig = ch.count;
            else if (ctx->path == 0 && ctx->split_big && K == 1) sc.nbig = split_big(hmc, hoc, ch.count, dv.sm_count);
            else if (ctx->path == 0 && ctx->split_big)          // hoc is largest first: the prefix at or above the shard's threshold
                while (sc.nbig < ch.count && (double)hmc[hoc[sc.nbig]].in_len + (double)hmc[hoc[sc.nbig]].out_cap >= big_threshold) sc.nbig++;
            plan_parts(hmc, hoc, ch.count, sc.nbig, region[c % NS], hs + ch.first, cuts_big_of[c], cuts_of[c], sc.big_seqs);
            sc.cuts = &cuts_of[c];
            sc.cuts_big = &cuts_big_of[c];
            CK(ctx, cudaMemcpyAsync(sc.plan, hs + ch.first, ch.count * sizeof(uint64_t), cudaMemcpyHostToDevice, dv.copy_stream));
            CK(ctx, cudaMemcpyAsync(sc.plan + ch.count * sizeof(uint64_t), hoc, ch.count * sizeof(uint32_t),
                                    cudaMemcpyHostToDevice, dv.copy_stream));
            CK(ctx, cudaEventRecord(ev_in, dv.copy_stream));
            if (trace) CK(ctx, cudaEventRecord(tev[3 * c], dv.copy_stream));
            CK(ctx, cudaStreamWaitEvent(dst, ev_in, 0));
            if ((rc = run_device(ctx, dv, ch.count, dm + ch.first, dr + ch.first, hm + ch.first, ho + ch.first, flags,
                                 dst, &sc)))
                return rc;
            CK(ctx, cudaEventRecord(ev_dec, dst));
            if (trace) CK(ctx, cudaEventRecord(tev[3 * c + 1], dst));
            CK(ctx, cudaStreamWaitEvent(dv.back_stream, ev_dec, 0));
            CK(ctx, cudaMemcpyAsync(hr + ch.first, dr + ch.first, ch.count * sizeof(zsb_result), cudaMemcpyDeviceToHost,
                                    dv.back_stream));
            // whole buffers come back (bytes past total_out are unspecified by contract): no host sync needed
            for (const Run &rn : ch.out_runs)
                CK(ctx, cudaMemcpyAsync((void *)rn.host, dv.d_out + ch.out_base + rn.dev_off, rn.len,
                                        cudaMemcpyDeviceToHost, dv.back_stream));
            if (trace) CK(ctx, cudaEventRecord(tev[3 * c + 2], dv.back_stream));
        }
        if (trace) {
            fprintf(stderr, "zsb trace host: all %d chunks enqueued %.2f ms after the call started\n", K,
                    std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_call).count());
            CK(ctx, cudaStreamSynchronize(dv.back_stream));
            for (int c = 0; c < K; c++) {
                float a = 0, b = 0, d = 0;
                cudaEventElapsedTime(&a, tev[3 * K], tev[3 * c]);
                cudaEventElapsedTime(&b, tev[3 * K], tev[3 * c + 1]);
                cudaEventElapsedTime(&d, tev[3 * K], tev[3 * c + 2]);
                fprintf(stderr, "zsb trace dev %d chunk %2d/%d members %zu: in %.2f dec %.2f out %.2f ms\n", s, c, K,
                        chunks[c].count, a, b, d);
            }
            for (auto &e : tev) cudaEventDestroy(e);
        }
    }
    for (int s = 0; s < nd; s++) {
        Shard &sh = shards[s];
        zsb_dev &dv = ctx->devs[s];
        const size_t k = sh.idx.size();
        if (!k) continue;
        CK(ctx, cudaSetDevice(dv.device));
        CK(ctx, cudaStreamSynchronize(dv.back_stream));
        zsb_result *hr = (zsb_result *)((zsb_member *)dv.h_desc + k);
        for (size_t j = 0; j < k; j++) r[sh.idx[j]] = hr[j];
    }
    return 0;
}

// Which CUDA device owns a device pointer, answered from a cache of allocation ranges: members of a batch live
// in a few large allocations, so one runtime query per allocation serves a million members.  The range comes from
// the driver's cuMemGetAddressRange, reached through the runtime (cudaGetDriverEntryPoint): libzsb.so does not
// link libcuda, so it still loads where no driver is installed (the CPU-only test tier).
struct PtrOwner {
    struct Range { uintptr_t lo, hi; int device; };
    std::vector<Range> ranges;
    size_t last = 0;
    typedef int (*range_fn)(unsigned long long *, size_t *, unsigned long long);
    range_fn get_range = nullptr;
    bool looked = false;
    // -1: not a device pointer
    int device_of(const void *p) {
        const uintptr_t a = (uintptr_t)p;
        if (last < ranges.size() && a >= ranges[last].lo && a < ranges[last].hi) return ranges[last].device;
        for (size_t k = 0; k < ranges.size(); k++)
            if (a >= ranges[k].lo && a < ranges[k].hi) {
                last = k;
                return ranges[k].device;
            }
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
            cudaGetLastError();
            return -1;
        }
        if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged) return -1;
        if (!looked) {
            looked = true;
            void *fn = nullptr;
            cudaDriverEntryPointQueryResult qr;
            if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qr) == cudaSuccess && fn &&
                qr == cudaDriverEntryPointSuccess)
                get_range = (range_fn)fn;
            else
                cudaGetLastError();
        }
        Range r{a, a + 1, at.device};
        unsigned long long base = 0;
        size_t size = 0;
        if (get_range && get_range(&base, &size, (unsigned long long)a) == 0 && size) {
            r.lo = (uintptr_t)base;
            r.hi = (uintptr_t)base + size;
        }
        ranges.push_back(r);
        last = ranges.size() - 1;
        return r.device;
    }
};

// Members whose buffers are already in HBM, on ANY device of the context ("inputs are placed directly
// in the owning GPU's HBM; outputs stay there"): every member is decoded by the device that owns its buffers, all
// devices are enqueued before any is drained, results return in the caller's order.  A member on a device the
// context does not hold is an argument error.
static int batch_device_ptrs(zsb_ctx *ctx, size_t n, const zsb_member *m, zsb_result *r) {
    const int nd = (int)ctx->devs.size();
    PtrOwner owner;
    std::vector<int> dev_index_of;                   // CUDA ordinal -> index in ctx->devs
    for (int s = 0; s < nd; s++) {
        if ((int)dev_index_of.size() <= ctx->devs[s].device) dev_index_of.resize(ctx->devs[s].device + 1, -1);
        dev_index_of[ctx->devs[s].device] = s;
    }
    std::vector<std::vector<uint32_t>> part(nd);
    for (size_t i = 0; i < n; i++) {
        const void *p = m[i].in ? m[i].in : m[i].out;
        int s = 0;
        if (p) {
            const int d = owner.device_of(p);
            s = (d >= 0 && d < (int)dev_index_of.size()) ? dev_index_of[d] : -1;
            if (s < 0) {
                snprintf(ctx->err, sizeof ctx->err, "member %zu: its buffers are not in the memory of a device of this context", i);
                return ZSB_E_ARG;
            }
        }
        part[s].push_back((uint32_t)i);
    }
    ctx->launches = 0;
    // (results come back through the device's pinned descriptor buffer: a copy to pageable memory would block the
    // host until the device has finished, and the devices would run one after the other)
    struct Pending { std::vector<zsb_member> mm; zsb_result *hr = nullptr; };
    std::vector<Pending> pend(nd);
    for (int s = 0; s < nd; s++) {
        const size_t k = part[s].size();
        if (!k) continue;
        zsb_dev &dv = ctx->devs[s];
        CK(ctx, cudaSetDevice(dv.device));
        int rc;
        const size_t desc_bytes = k * (sizeof(zsb_member) + sizeof(zsb_result) + sizeof(uint32_t)) + 64;
        if ((rc = grow(ctx, dv.d_desc, dv.d_desc_cap, desc_bytes))) return rc;
        zsb_member *dm = (zsb_member *)dv.d_desc;
        zsb_result *dr = (zsb_result *)(dm + k);
        Pending &P = pend[s];
        P.mm.resize(k);
        if (k * sizeof(zsb_result) > dv.h_desc_cap) {
            if (dv.h_desc) CK(ctx, cudaFreeHost(dv.h_desc));
            dv.h_desc = nullptr;
            dv.h_desc_cap = 0;
            const size_t want = k * sizeof(zsb_result) + k * sizeof(zsb_result) / 8 + 4096;
            CK(ctx, cudaHostAlloc(&dv.h_desc, want, cudaHostAllocDefault));
            dv.h_desc_cap = want;
        }
        P.hr = (zsb_result *)dv.h_desc;
        bool raw_only = true;
        for (size_t j = 0; j < k; j++) {
            P.mm[j] = m[part[s][j]];
            raw_only &= P.mm[j].window_bits < 0 && P.mm[j].flags == 0;
            if (P.mm[j].in_len > 0xfffffff0ull || P.mm[j].out_cap > 0xfffffff0ull) P.mm[j].window_bits = 1000;   // -> Z_STREAM_ERROR
        }
        cudaStream_t st = dv.stream;
        CK(ctx, cudaMemcpyAsync(dm, P.mm.data(), k * sizeof(zsb_member), cudaMemcpyHostToDevice, st));
        // (no slot order of our own: run_device orders largest first and takes the large members out for the segmented path)
        if ((rc = run_device(ctx, dv, k, dm, dr, P.mm.data(), nullptr, raw_only ? ZSB_RAW_ONLY : ZSB_HAS_WRAPPED, st))) return rc;
        CK(ctx, cudaMemcpyAsync(P.hr, dr, k * sizeof(zsb_result), cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < nd; s++) {
        const size_t k = part[s].size();
        if (!k) continue;
        CK(ctx, cudaSetDevice(ctx->devs[s].device));
        CK(ctx, cudaStreamSynchronize(ctx->devs[s].stream));
        for (size_t j = 0; j < k; j++) r[part[s][j]] = pend[s].hr[j];
    }
    return 0;
}

// Host-pointer batches with LARGE members: the chunk pipeline of batch_host() decodes a member with one lane or one
// warp, so the members the segmented path would take (run_device: seg_min compressed bytes or more, shrinking) are
// staged whole on the devices, decoded as a device-resident batch (in segments) and copied back; the rest goes
// through the pipeline as before.  No overlap of copies and decode for the large part.
static int batch_host_split(zsb_ctx *ctx, size_t n, const zsb_member *m, zsb_result *r) {
    std::vector<uint32_t> L, S;
    if (ctx->segment && ctx->path == 0 && ctx->split_big != 2)
        for (size_t i = 0; i < n; i++) {
            const zsb_member &x = m[i];
            if (x.in && x.out && x.in_len >= ctx->seg_min && x.in_len <= 0xfffffff0ull && x.out_cap <= 0xfffffff0ull &&
                !(x.flags & ZSB_M_STORED) && wbits_valid(x.window_bits) && (uint64_t)x.in_len * 16 < (uint64_t)x.out_cap * 15)
                L.push_back((uint32_t)i);
        }
    if (L.empty()) return batch_host(ctx, n, m, r);
    int rc = 0;
    long launches = 0;
    if (L.size() < n) {                            // the rest, through the pipeline
        std::vector<char> isL(n, 0);
        for (uint32_t i : L) isL[i] = 1;
        for (size_t i = 0; i < n; i++)
            if (!isL[i]) S.push_back((uint32_t)i);
        std::vector<zsb_member> ms(S.size());
        std::vector<zsb_result> rs(S.size());
        for (size_t k = 0; k < S.size(); k++) ms[k] = m[S[k]];
        if ((rc = batch_host(ctx, S.size(), ms.data(), rs.data()))) return rc;
        for (size_t k = 0; k < S.size(); k++) r[S[k]] = rs[k];
        launches = ctx->launches;
    }
    const int nd = (int)ctx->devs.size();
    const size_t nl = L.size();
    std::vector<int32_t> shard_of(nl, 0);
    if (nd > 1) {
        std::vector<uint64_t> cost(nl);
        for (size_t k = 0; k < nl; k++) cost[k] = (uint64_t)m[L[k]].in_len + m[L[k]].out_cap;
        zsb_plan_shards(nl, cost.data(), nd, shard_of.data(), nullptr);
    }
    std::vector<size_t> in_off(nl), out_off(nl), in_total(nd, 0), out_total(nd, 0);
    for (size_t k = 0; k < nl; k++) {
        const int s = shard_of[k];
        in_off[k] = in_total[s];
        out_off[k] = out_total[s];
        in_total[s] += (m[L[k]].in_len + 19 + 255) & ~(size_t)255;
        out_total[s] += (m[L[k]].out_cap + 255) & ~(size_t)255;
    }
    for (int s = 0; s < nd; s++) {
        if (!in_total[s]) continue;
        zsb_dev &dv = ctx->devs[s];
        CK(ctx, cudaSetDevice(dv.device));
        if ((rc = grow(ctx, dv.d_in, dv.d_in_cap, in_total[s] + 256)) || (rc = grow(ctx, dv.d_out, dv.d_out_cap, out_total[s] + 256))) return rc;
    }
    std::vector<zsb_member> md(nl);
    std::vector<zsb_result> rd(nl);
    for (size_t k = 0; k < nl; k++) {
        zsb_dev &dv = ctx->devs[shard_of[k]];
        CK(ctx, cudaSetDevice(dv.device));
        md[k] = m[L[k]];
        md[k].in = dv.d_in + in_off[k];
        md[k].out = dv.d_out + out_off[k];
        CK(ctx, cudaMemcpyAsync(dv.d_in + in_off[k], m[L[k]].in, m[L[k]].in_len, cudaMemcpyHostToDevice, dv.copy_stream));
    }
    for (int s = 0; s < nd; s++)
        if (in_total[s]) {
            CK(ctx, cudaSetDevice(ctx->devs[s].device));
            CK(ctx, cudaStreamSynchronize(ctx->devs[s].copy_stream));
        }
    if ((rc = batch_device_ptrs(ctx, nl, md.data(), rd.data()))) return rc;
    for (size_t k = 0; k < nl; k++) {
        zsb_dev &dv = ctx->devs[shard_of[k]];
        CK(ctx, cudaSetDevice(dv.device));
        r[L[k]] = rd[k];
        const size_t nb = std::min<size_t>((size_t)rd[k].total_out, m[L[k]].out_cap);
        if (nb) CK(ctx, cudaMemcpyAsync(m[L[k]].out, dv.d_out + out_off[k], nb, cudaMemcpyDeviceToHost, dv.back_stream));
    }
    for (int s = 0; s < nd; s++)
        if (in_total[s]) {
            CK(ctx, cudaSetDevice(ctx->devs[s].device));
            CK(ctx, cudaStreamSynchronize(ctx->devs[s].back_stream));
        }
    ctx->launches += launches;
    return 0;
}

extern "C" int zsb_inflate_batch(zsb_ctx *ctx, size_t n, const zsb_member *m, zsb_result *r) {
    if (!ctx || (n && (!m || !r))) return ZSB_E_ARG;
    if (n == 0) return 0;
    int cur = 0;
    cudaGetDevice(&cur);
    // host or device pointers?  decided by the first member that has any buffer
    bool device_ptrs = false;
    for (size_t i = 0; i < n; i++) {
        const void *p = m[i].in ? m[i].in : m[i].out;
        if (!p) continue;
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) == cudaSuccess)
            device_ptrs = at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
        else
            cudaGetLastError();
        break;
    }
    int rc = device_ptrs ? batch_device_ptrs(ctx, n, m, r) : batch_host_split(ctx, n, m, r);
    cudaSetDevice(cur);
    return rc;
}

// the ZIP container walker (zsb_zip_scan / zsb_zip_members / zsb_zip_inflate)
#include "zsb_zip.cuh"

// ---- batched deflate (zsb_deflate_batch): blocks of 64 KiB, a warp each (zsb_deflate.cuh) ----------------------
#include "zsb_deflate.cuh"

extern "C" size_t zsb_deflate_bound(size_t in_len, int window_bits) {
    const size_t blocks = in_len / df::BLK + 1;
    return in_len + blocks * 16 + (window_bits < 0 ? 0 : window_bits > 15 ? 18 : 6) + 16;
}

extern "C" int zsb_deflate_batch(zsb_ctx *ctx, size_t n, const zsb_deflate_member *m, zsb_deflate_result *r) {
    static_assert(sizeof(df::Member) == sizeof(zsb_deflate_member) && sizeof(DfResult) == sizeof(zsb_deflate_result), "ABI");
    if (!ctx || (n && (!m || !r))) return ZSB_E_ARG;
    if (n == 0) return 0;
    if (n > 0x7fffffffull) return ZSB_E_ARG;
    int cur = 0;
    cudaGetDevice(&cur);
    zsb_dev &dv = ctx->devs[0];
    struct Restore { int cur; ~Restore() { cudaSetDevice(cur); } } restore{cur};
    CK(ctx, cudaSetDevice(dv.device));
    bool device_ptrs = false;
    for (size_t i = 0; i < n; i++) {
        const void *p = m[i].in ? m[i].in : m[i].out;
        if (!p) continue;
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, p) == cudaSuccess) device_ptrs = at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
        else cudaGetLastError();
        break;
    }
    auto valid = [](const zsb_deflate_member &x) {
        return (x.window_bits == -15 || x.window_bits == 15 || x.window_bits == 31) && x.level >= -1 && x.level <= 9 &&
               x.in_len <= 0xfffffff0ull && (x.in || !x.in_len) && (x.out || !x.out_cap);
    };
    // plan: members -> blocks, scratch offsets, staging offsets
    std::vector<df::Member> hm(n);
    std::vector<uint32_t> first(n + 1, 0);
    std::vector<df::Block> hb;
    std::vector<size_t> in_off(n, 0), out_off(n, 0);
    size_t scratch = 0, in_total = 0, out_total = 0;
    for (size_t i = 0; i < n; i++) {
        memcpy(&hm[i], &m[i], sizeof(df::Member));
        first[i] = (uint32_t)hb.size();
        const bool ok = valid(m[i]);
        if (!ok) {
            hm[i].in_len = 0;
            hm[i].out_cap = 0;
            hm[i].window_bits = -15;
            hm[i].level = 6;
        }
        in_off[i] = in_total;
        out_off[i] = out_total;
        in_total += (hm[i].in_len + 19) & ~(size_t)15;          // (3 bytes of slack behind every input: word reads)
        out_total += (hm[i].out_cap + 15) & ~(size_t)15;
        if (!ok) continue;
        const uint64_t len = hm[i].in_len;
        const uint64_t nblk = len ? (len + df::BLK - 1) / df::BLK : 1;
        if (hb.size() + nblk > 0x7fffffffull) return ZSB_E_ARG;
        for (uint64_t b = 0; b < nblk; b++) {
            df::Block k;
            k.member = (uint32_t)i;
            k.start = (uint32_t)(b * df::BLK);
            k.scratch_off = scratch;
            k.last = b + 1 == nblk;
            k.nbytes = 0;
            const uint64_t nb = std::min<uint64_t>(df::BLK, len - b * df::BLK);
            scratch += (nb + 64 + 15) & ~(size_t)15;
            hb.push_back(k);
        }
    }
    first[n] = (uint32_t)hb.size();
    const size_t nb = hb.size();
    static bool configured = false;
    if (!configured) {
        CK(ctx, cudaFuncSetAttribute(k_deflate, cudaFuncAttributeMaxDynamicSharedMemorySize, df::D_BYTES));
        configured = true;
    }
    int occ = 0;
    CK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_deflate, 32, df::D_BYTES));
    if (occ < 1) occ = 1;
    const unsigned grid = (unsigned)std::max<size_t>(1, std::min<size_t>(nb, (size_t)dv.sm_count * occ));
    // one device allocation, carved up
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o_mem = 0, o_blk = o_mem + up(n * sizeof(df::Member)), o_first = o_blk + up(nb * sizeof(df::Block)),
                 o_res = o_first + up((n + 1) * 4), o_cnt = o_res + up(n * sizeof(DfResult)), o_tok = o_cnt + 256,
                 o_scr = o_tok + up((size_t)grid * DF_TOK_WORDS * 4), o_in = o_scr + up(scratch + 64),
                 o_out = o_in + (device_ptrs ? 0 : up(in_total + 64)), need = o_out + (device_ptrs ? 0 : up(out_total + 64));
    int rc;
    if ((rc = grow(ctx, dv.d_df, dv.d_df_cap, need))) return rc;
    uint8_t *base = dv.d_df;
    cudaStream_t st = dv.stream;
    if (!device_ptrs) {
        for (size_t i = 0; i < n; i++) {
            if (hm[i].in_len) CK(ctx, cudaMemcpyAsync(base + o_in + in_off[i], hm[i].in, hm[i].in_len, cudaMemcpyHostToDevice, st));
            hm[i].in = base + o_in + in_off[i];
            hm[i].out = base + o_out + out_off[i];
        }
    }
    CK(ctx, cudaMemcpyAsync(base + o_mem, hm.data(), n * sizeof(df::Member), cudaMemcpyHostToDevice, st));
    if (nb) CK(ctx, cudaMemcpyAsync(base + o_blk, hb.data(), nb * sizeof(df::Block), cudaMemcpyHostToDevice, st));
    CK(ctx, cudaMemcpyAsync(base + o_first, first.data(), (n + 1) * 4, cudaMemcpyHostToDevice, st));
    CK(ctx, cudaMemsetAsync(base + o_cnt, 0, 256, st));
    ctx->launches = 0;
    if (nb) {
        k_deflate<<<grid, 32, df::D_BYTES, st>>>((const df::Member *)(base + o_mem), (df::Block *)(base + o_blk), (uint32_t)nb,
                                                 (uint32_t *)(base + o_cnt), (uint32_t *)(base + o_tok), base + o_scr);
        ctx->launches++;
        CK(ctx, cudaGetLastError());
    }
    k_dstitch<<<(unsigned)std::min<size_t>(n, (size_t)dv.sm_count * 8), CHK_THREADS, 0, st>>>(
        (const df::Member *)(base + o_mem), (const df::Block *)(base + o_blk), (const uint32_t *)(base + o_first), (uint32_t)n,
        (uint32_t *)(base + o_cnt) + 1, base + o_scr, (DfResult *)(base + o_res));
    ctx->launches++;
    CK(ctx, cudaGetLastError());
    CK(ctx, cudaMemcpyAsync(r, base + o_res, n * sizeof(DfResult), cudaMemcpyDeviceToHost, st));
    CK(ctx, cudaStreamSynchronize(st));
    for (size_t i = 0; i < n; i++) {
        if (!valid(m[i])) {
            memset(&r[i], 0, sizeof r[i]);
            r[i].ret = -2;
            continue;
        }
        if (!device_ptrs && r[i].ret == 1 && r[i].total_out)
            CK(ctx, cudaMemcpyAsync(m[i].out, base + o_out + out_off[i], r[i].total_out, cudaMemcpyDeviceToHost, st));
    }
    if (!device_ptrs) CK(ctx, cudaStreamSynchronize(st));
    return 0;
}

// ---- zsb_deflate_batch_static: the same two kernels with NO allocation anywhere -------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No
// counterpart in the reference.  Everything zsb_deflate_batch() carves
// out of its grown device allocation (descriptors, block table, results, counters, token scratch, block bytes)
// lives in one file-scope device array of a fixed size, so the batch is limited to ZSB_STATIC_MAX_MEMBERS members
// and ZSB_STATIC_MAX_BLOCKS 64 KiB blocks.  Every GPU that loads the library holds the array until exit; every
// call shares it (the host part is serialised by a mutex and the call synchronises its stream before it returns,
// so calls of one process do not overlap on the GPU; anything else that made them overlap would race).
constexpr size_t ST_MEMBERS = 64, ST_BLOCKS = 64, ST_GRID = 16;
constexpr size_t st_up(size_t v) { return (v + 255) & ~(size_t)255; }
constexpr size_t ST_O_MEM = 0, ST_O_BLK = ST_O_MEM + st_up(ST_MEMBERS * sizeof(df::Member)),
                 ST_O_FIRST = ST_O_BLK + st_up(ST_BLOCKS * sizeof(df::Block)), ST_O_RES = ST_O_FIRST + st_up((ST_MEMBERS + 1) * 4),
                 ST_O_CNT = ST_O_RES + st_up(ST_MEMBERS * sizeof(DfResult)), ST_O_TOK = ST_O_CNT + 256,
                 ST_O_SCR = ST_O_TOK + st_up(ST_GRID * DF_TOK_WORDS * 4),
                 ST_BYTES = ST_O_SCR + st_up(ST_BLOCKS * (df::BLK + 64 + 16) + 64);
static __device__ __align__(256) uint8_t g_static_scratch[ST_BYTES];       // 8.4 MB per GPU

static int st_fail(char *err, size_t cap, int code, const char *fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

extern "C" size_t zsb_deflate_static_limits(size_t *max_members, size_t *max_blocks) {
    if (max_members) *max_members = ST_MEMBERS;
    if (max_blocks) *max_blocks = ST_BLOCKS;
    return ST_BYTES;
}

extern "C" int zsb_deflate_batch_static(int device, size_t n, const zsb_deflate_member *m, zsb_deflate_result *r, void *stream,
                                        char *err, size_t err_cap) {
    static std::mutex mu;
    static uint8_t *base_of[64];            // device address of g_static_scratch, per device (0 = not asked yet)
    if (err && err_cap) err[0] = 0;
    if (n && (!m || !r)) return st_fail(err, err_cap, ZSB_E_ARG, "members / results is NULL");
    if (device < 0 || device >= 64) return st_fail(err, err_cap, ZSB_E_ARG, "device %d out of range", device);
    if (n == 0) return 0;
    if (n > ST_MEMBERS)
        return st_fail(err, err_cap, ZSB_E_LIMIT, "%zu members: the static scratch holds at most %zu per call", n, ST_MEMBERS);
    df::Member hm[ST_MEMBERS];
    df::Block hb[ST_BLOCKS];
    uint32_t first[ST_MEMBERS + 1];
    bool ok[ST_MEMBERS];
    size_t nb = 0, scratch = 0;
    for (size_t i = 0; i < n; i++) {
        const zsb_deflate_member &x = m[i];
        memcpy(&hm[i], &x, sizeof(df::Member));
        first[i] = (uint32_t)nb;
        ok[i] = (x.window_bits == -15 || x.window_bits == 15 || x.window_bits == 31) && x.level >= -1 && x.level <= 9 &&
                x.in_len <= 0xfffffff0ull && (x.in || !x.in_len) && (x.out || !x.out_cap);
        if (!ok[i]) {
            hm[i].in_len = 0;
            hm[i].out_cap = 0;
            hm[i].window_bits = -15;
            hm[i].level = 6;
            continue;
        }
        const uint64_t len = x.in_len, nblk = len ? (len + df::BLK - 1) / df::BLK : 1;
        if (nb + nblk > ST_BLOCKS)
            return st_fail(err, err_cap, ZSB_E_LIMIT, "more than %zu blocks of 64 KiB (member %zu of %zu bytes): beyond the static scratch",
                           ST_BLOCKS, i, (size_t)len);
        for (uint64_t b = 0; b < nblk; b++) {
            df::Block &k = hb[nb++];
            k.member = (uint32_t)i;
            k.start = (uint32_t)(b * df::BLK);
            k.scratch_off = scratch;
            k.last = b + 1 == nblk;
            k.nbytes = 0;
            scratch += (std::min<uint64_t>(df::BLK, len - b * df::BLK) + 64 + 15) & ~(size_t)15;
        }
    }
    first[n] = (uint32_t)nb;
    for (size_t i = 0; i < n; i++) {        // the caller's buffers must be device memory: there is nowhere to stage them
        const void *ps[2] = {hm[i].in_len ? (const void *)hm[i].in : nullptr, hm[i].out_cap ? (const void *)hm[i].out : nullptr};
        for (const void *p : ps) {
            if (!p) continue;
            cudaPointerAttributes at;
            if (cudaPointerGetAttributes(&at, p) != cudaSuccess || !(at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged)) {
                cudaGetLastError();
                return st_fail(err, err_cap, ZSB_E_ARG, "member %zu: in / out must be device pointers", i);
            }
        }
    }
    std::lock_guard<std::mutex> lk(mu);
    int cur = 0;
    cudaGetDevice(&cur);
    struct Restore { int cur; ~Restore() { cudaSetDevice(cur); } } restore{cur};
    cudaError_t e;
#define ST_CK(call)                                                                                         \
    if ((e = (call)) != cudaSuccess)                                                                        \
        return st_fail(err, err_cap, ZSB_E_CUDA, "%s: %s", #call, cudaGetErrorString(e))
    ST_CK(cudaSetDevice(device));
    if (!base_of[device]) {                 // once per device, with that device current
        uint8_t *p = nullptr;
        ST_CK(cudaGetSymbolAddress((void **)&p, g_static_scratch));
        base_of[device] = p;
    }
    uint8_t *base = base_of[device];
    cudaStream_t st = (cudaStream_t)stream;
    ST_CK(cudaMemcpyAsync(base + ST_O_MEM, hm, n * sizeof(df::Member), cudaMemcpyHostToDevice, st));
    if (nb) ST_CK(cudaMemcpyAsync(base + ST_O_BLK, hb, nb * sizeof(df::Block), cudaMemcpyHostToDevice, st));
    ST_CK(cudaMemcpyAsync(base + ST_O_FIRST, first, (n + 1) * 4, cudaMemcpyHostToDevice, st));
    ST_CK(cudaMemsetAsync(base + ST_O_CNT, 0, 256, st));             // the two work counters start at zero every call
    if (nb) {
        k_deflate<<<(unsigned)std::min(nb, ST_GRID), 32, df::D_BYTES, st>>>((const df::Member *)(base + ST_O_MEM), (df::Block *)(base + ST_O_BLK),
                                                                            (uint32_t)nb, (uint32_t *)(base + ST_O_CNT),
                                                                            (uint32_t *)(base + ST_O_TOK), base + ST_O_SCR);
        ST_CK(cudaGetLastError());
    }
    k_dstitch<<<(unsigned)n, CHK_THREADS, 0, st>>>((const df::Member *)(base + ST_O_MEM), (const df::Block *)(base + ST_O_BLK),
                                                   (const uint32_t *)(base + ST_O_FIRST), (uint32_t)n, (uint32_t *)(base + ST_O_CNT) + 1,
                                                   base + ST_O_SCR, (DfResult *)(base + ST_O_RES));
    ST_CK(cudaGetLastError());
    ST_CK(cudaMemcpyAsync(r, base + ST_O_RES, n * sizeof(DfResult), cudaMemcpyDeviceToHost, st));
    ST_CK(cudaStreamSynchronize(st));
#undef ST_CK
    for (size_t i = 0; i < n; i++)
        if (!ok[i]) {
            memset(&r[i], 0, sizeof r[i]);
            r[i].ret = -2;
        }
    return 0;
}

// the z_stream surface (inflateInit2_ / inflate / inflateEnd): same translation unit, it shares the kernels above
#include "zsb_zstream.cuh"
