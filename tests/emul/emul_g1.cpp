// tests/emul/emul_g1.cpp — TEST TOOL ONLY, never part of libzsb.so.
// Compiles the *device* source of the decoder (zlib-streams_b200/csrc/zsb_device.cuh) for the host
// with the CUDA keywords and the few intrinsics it uses shimmed, instantiated for a group width of
// one lane.  It lets the CPU-only test tier (-m "not gpu") run the decoder's control logic (headers,
// table build, symbol loop, stop bookkeeping) against the oracle where no GPU exists.  The product
// has no CPU path: libzsb.so is built from the same header by nvcc and runs only on the GPU.
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <vector>
#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__
static inline uint32_t __brev(uint32_t x) {
    x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
    x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
    x = ((x >> 4) & 0x0f0f0f0fu) | ((x & 0x0f0f0f0fu) << 4);
    x = ((x >> 8) & 0x00ff00ffu) | ((x & 0x00ff00ffu) << 8);
    return (x >> 16) | (x << 16);
}
template <class T> static inline T __ldg(const T *p) { return *p; }
static inline uint32_t __shfl_sync(unsigned, uint32_t v, int, int = 32) { return v; }
static inline void __syncwarp(unsigned = 0xffffffffu) {}
static inline bool __any_sync(unsigned, bool p) { return p; }
static inline bool __all_sync(unsigned, bool p) { return p; }
#define ZSB_FETCH_SLOT(counter) ((*(counter))++)
static inline uint32_t __shfl_up_sync(unsigned, uint32_t v, int, int = 32) { return v; }
static inline uint32_t __shfl_down_sync(unsigned, uint32_t v, int, int = 32) { return v; }
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return v; }
static inline unsigned __ballot_sync(unsigned, bool p) { return p; }
static inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
static inline unsigned __match_any_sync(unsigned, uint32_t) { return 1u; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
#define ZSB_SMEM_OR(p, v) (*(p) |= (v))
#define ZSB_CP4(sdst, gsrc) (*(sdst) = *(const uint32_t *)(gsrc))
#define ZSB_CP_COMMIT() ((void)0)
#define ZSB_CP_WAIT(n) ((void)0)
#define ZSB_FETCH_ADD(p, v) ((*(p) += (v)) - (v))
#define ZSB_SMEM_ADD(p, v) (*(p) += (v))
#include "../../zlib-streams_b200/csrc/zsb_device.cuh"
#include "../../zlib-streams_b200/csrc/zsb_lz.cuh"

extern "C" int emul_decode(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits,
                           zsb_result *res, uint32_t *aux4) {
    static uint16_t fixed[zsb::STREAM_U16];
    static uint32_t lut[zsb::LUT_WORDS];
    static bool init = false;
    if (!init) {
        zsb::init_lut(lut, 0, 1);
        zsb::build_fixed(fixed);
        init = true;
    }
    alignas(16) uint16_t T[zsb::STREAM_U16];
    memset(T, 0xA5, sizeof T);
    zsb_member m;
    m.in = in; m.in_len = in_len; m.out = out; m.out_cap = out_cap; m.window_bits = wbits; m.flags = 0;
    zsb::Aux a;
    uint32_t counter = 0;
    zsb::warp_inflate<1>(&m, res, &a, nullptr, 1, &counter, T, lut, fixed, 0, 1u);
    aux4[0] = a.kind; aux4[1] = a.trailer_pos; aux4[2] = a.need_trailer; aux4[3] = a.init_check;
    return res->ret;
}

static int g_big = 0;
extern "C" void emul_set_big(int on) { g_big = on; }

// the lane path with a warp of one lane: k_lane's decoder, then k_lz's resolver
extern "C" int emul_decode_lane(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits,
                                zsb_result *res, uint32_t *aux4, uint64_t seq_cap, uint32_t *nseq_out) {
    using namespace zsb;
    static uint8_t fixed[lane::SLOT];
    static uint32_t lut[LUT_WORDS];
    static bool init = false;
    alignas(16) static uint8_t scr[lane::WSCR];
    if (!init) {
        init_lut(lut, 0, 1);
        alignas(16) uint8_t T0[lane::SLOT];
        for (int i = 0; i < 288; i++) T0[lane::H_LENS + i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
        for (int i = 0; i < 32; i++) T0[lane::H_LENS + 288 + i] = 5;
        lane::build_tables<1>(T0, scr, 288, 32, false, 0);
        memcpy(fixed, T0, sizeof fixed);
        init = true;
    }
    alignas(16) uint8_t T[lane::SLOT];
    memset(T, 0xA5, sizeof T);
    zsb_member m;
    m.in = in; m.in_len = in_len; m.out = out; m.out_cap = out_cap; m.window_bits = wbits; m.flags = 0;
    Aux a;
    uint32_t counter = 0, order = 0;
    uint64_t seq_off = 0;
    lane::seq_t *seq = (lane::seq_t *)malloc((seq_cap + 1) * sizeof(lane::seq_t));
    lane::Batch B;
    B.members = &m; B.results = res; B.aux = &a; B.order = &order; B.seq_off = &seq_off; B.seq_base = seq;
    B.hist = nullptr; B.n = 1; B.counter = &counter; B.fixed_tab = fixed;
    uint32_t ring[lane::RING_WORDS] = {0, 0, 0, 0};
    if (g_big) {                                     // k_big's decoder: lane 0 with the direct tables
        alignas(16) static uint32_t big_lut[lane::BIG_LUT_BYTES / 4];
        memset(big_lut, 0xA5, sizeof big_lut);
        alignas(16) static lane::seq_t big_q[lane::BIG_QCAP];
        lane::warp_lanes<1, true>(B, T, scr, ring, lut, 0, big_lut, big_q);
    } else {
        lane::warp_lanes<1>(B, T, scr, ring, lut, 0);
    }
    if (nseq_out) *nseq_out = a.nseq;
    int rc = res->ret;
    if (a.nseq > seq_cap) rc = -1000;               // the host's scratch bound was too small
    else if (a.nseq) lz::resolve_member<1>(out, in, seq, a.nseq, 0);
    free(seq);
    aux4[0] = a.kind; aux4[1] = a.trailer_pos; aux4[2] = a.need_trailer; aux4[3] = a.init_check;
    return rc;
}

// the segmented path of large members (zsb_seg.cuh) with a warp of one lane: find the block starts near multiples of
// segb compressed bytes, decode the segments, chain them, resolve; info[0] = starts found, info[1] = 1 when the chain
// held (0: the caller sees the serial fallback's result, as on the device), info[2] = segments, info[3] = live ones
#include "../../zlib-streams_b200/csrc/zsb_seg.cuh"
extern "C" int emul_decode_seg(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits, uint32_t segb,
                               uint32_t seg_q, zsb_result *res, uint32_t *aux4, uint32_t *info) {
    using namespace zsb;
    alignas(16) static uint32_t fixLL[lane::BIG_LUT_BYTES / 4];
    static uint16_t ktab[seg::KTAB_ENTRIES];
    static uint8_t fixed[lane::SLOT];
    static uint32_t lut[LUT_WORDS];
    static bool init = false;
    alignas(16) static uint8_t scr[lane::WSCR];
    if (!init) {
        init_lut(lut, 0, 1);
        alignas(16) uint8_t T0[lane::SLOT];
        for (int i = 0; i < 288; i++) T0[lane::H_LENS + i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
        for (int i = 0; i < 32; i++) T0[lane::H_LENS + 288 + i] = 5;
        lane::build_tables<1>(T0, scr, 288, 32, false, 0);
        memcpy(fixed, T0, sizeof fixed);
        lane::big_build<1>(fixLL, fixed, lut, 0u, 0);
        for (uint32_t i = 0; i < (uint32_t)seg::KTAB_ENTRIES; i++) ktab[i] = seg::ktab_entry(i);
        init = true;
    }
    alignas(16) uint8_t T[lane::SLOT];
    memset(T, 0xA5, sizeof T);
    zsb_member m;
    m.in = in; m.in_len = in_len; m.out = out; m.out_cap = out_cap; m.window_bits = wbits; m.flags = 0;
    const uint32_t nseg = (uint32_t)((in_len + segb - 1) / segb);
    std::vector<lane::Seg> segs(nseg ? nseg : 1);
    std::vector<uint64_t> seq_off(segs.size());
    std::vector<lane::seq_t> arena((size_t)segs.size() * seg_q + 64);
    alignas(16) static uint8_t fsm[seg::F_BYTES];
    alignas(16) static uint8_t ringbuf[lz::RING_BYTES];
    uint32_t found = 0;
    for (uint32_t k = 0; k < segs.size(); k++) {
        lane::Seg &g = segs[k];
        memset(&g, 0, sizeof g);
        g.member = 0; g.k = k; g.first = 0; g.nseg = (uint32_t)segs.size();
        g.start_bit = 0; g.target = lane::SEG_NONE; g.ev = lane::EV_ABSENT;
        seq_off[k] = (uint64_t)k * seg_q;
        if (k) g.start_bit = lane::SEG_NONE;
        if (k && wbits != -16 && (uint64_t)k * segb >= seg::first_start(false))
            g.start_bit = seg::find_sync_fixed<1>(true, in, (uint32_t)in_len, 8ull * k * segb, 8ull * k * segb + 16ull * segb, fixLL);
        if (k && g.start_bit != lane::SEG_NONE) found++;
        else if (k && (uint64_t)k * segb >= seg::first_start(wbits == -16)) {
            g.start_bit = seg::find_start<1>(in, (uint32_t)in_len, wbits == -16, 8ull * k * segb, 8ull * (k + 1) * segb, fsm, ktab, lut, 0);
            if (g.start_bit != lane::SEG_NONE) found++;
        }
    }
    seg::plan_member(segs.data(), segb, (uint32_t)in_len, wbits == -16);
    Aux a;
    memset(&a, 0, sizeof a);
    uint32_t counter = 0;
    lane::Batch B;
    B.segs = segs.data(); B.seg_q = seg_q;
    B.members = &m; B.results = res; B.aux = &a; B.order = nullptr; B.seq_off = seq_off.data(); B.seq_base = arena.data();
    B.hist = nullptr; B.n = (uint32_t)segs.size(); B.counter = &counter; B.fixed_tab = fixed;
    uint32_t ring[lane::RING_WORDS] = {0, 0, 0, 0};
    lane::warp_lanes<1, false, true>(B, T, scr, ring, lut, 0);
    const bool ok = seg::chain_member(B, 0);
    info[0] = found; info[1] = ok; info[2] = (uint32_t)segs.size(); info[3] = 0;
    if (getenv("ZSB_EMUL_TRACE"))
        for (uint32_t k = 0; k < segs.size(); k++)
            if (segs[k].start_bit != lane::SEG_NONE)
                fprintf(stderr, "seg %u start %llu target %llu ev %d end %llu pos %u nseq %u base %u live %u\n", k,
                        (unsigned long long)segs[k].start_bit, (unsigned long long)segs[k].target, segs[k].ev,
                        8ull * segs[k].next - segs[k].nbits, segs[k].pos, segs[k].nseq, segs[k].base, segs[k].live);
    if (ok) {
        for (uint32_t k = 0; k < segs.size(); k++)
            if (segs[k].live) {
                info[3]++;
                if (!segs[k].nseq) continue;
                if (wbits != -16) lz::resolve_range_ring<1>(out, in, arena.data() + seq_off[k], segs[k].nseq, 0, segs[k].base, ringbuf);
                else lz::resolve_range<1>(out, in, arena.data() + seq_off[k], segs[k].nseq, 0, segs[k].base);
            }
        aux4[0] = a.kind; aux4[1] = a.trailer_pos; aux4[2] = a.need_trailer; aux4[3] = a.init_check;
        return res->ret;
    }
    uint32_t nseq = 0;
    return emul_decode_lane(in, in_len, out, out_cap, wbits, res, aux4, (uint64_t)out_cap + 8 * in_len + 64, &nseq);
}

// second half of the pipeline: what k_check does once the checksum of the output is known
extern "C" void emul_finish(const uint8_t *in, size_t in_len, zsb_result *res, const uint32_t *aux4,
                            uint32_t check_of_output) {
    zsb::Aux a;
    a.kind = aux4[0]; a.trailer_pos = aux4[1]; a.need_trailer = aux4[2]; a.init_check = aux4[3];
    if (a.kind == 0) return;
    uint32_t E = (uint32_t)res->total_out;
    zsb::finish_member(*res, a, in, (uint32_t)in_len, E ? check_of_output : a.init_check, E);
}

// ---------------------------------------------------------------------------------------------------
// The z_stream surface's device code (zsb_stream.cuh) with a warp of one lane: open/step/close with the
// signatures of oracle_api.h, so the streaming traces can be compared call by call with the oracle.
// The host bookkeeping below restates inflate() of zsb_zstream.cuh for a single slice.
// ---------------------------------------------------------------------------------------------------
#include <vector>
#include "../../zlib-streams_b200/csrc/zsb_stream.cuh"
#include "../../oracle/oracle_api.h"

namespace {
struct EmulZs {
    zsb::zs::State S;
    std::vector<uint8_t> hist, carry;
    uint32_t carry_len = 0, phase = 0, last_dt = 0;
    int wrap = 0;
    uint64_t total_in = 0, total_out = 0;
    uint32_t adler = 0;
    int msg = 0, data_type = 0;
};
uint32_t emul_crc(uint32_t crc, const uint8_t *p, size_t n) {
    crc = ~crc;
    for (size_t i = 0; i < n; i++) {
        crc ^= p[i];
        for (int k = 0; k < 8; k++) crc = (crc >> 1) ^ (0xEDB88320u & (0u - (crc & 1u)));
    }
    return ~crc;
}
uint32_t emul_adler(uint32_t ad, const uint8_t *p, size_t n) {
    uint32_t a = ad & 0xffff, b = ad >> 16;
    for (size_t i = 0; i < n; i++) {
        a = (a + p[i]) % 65521u;
        b = (b + a) % 65521u;
    }
    return (b << 16) | a;
}
}  // namespace

extern "C" void *emulzs_open(int windowBits, int *init_ret) {
    int wrap, d64 = 0, wb = windowBits, ret = 0;
    if (wb < 0) {
        if (wb < -16) ret = -2;
        wrap = 0;
        d64 = wb == -16;
        wb = -wb;
    } else {
        wrap = (wb >> 4) + 5;
        if (wb < 48) wb &= 15;
    }
    if (!ret && wb && (wb < 8 || wb > (d64 ? 16 : 15))) ret = -2;
    if (init_ret) *init_ret = ret;
    if (ret) return nullptr;
    EmulZs *z = new EmulZs();
    memset(&z->S, 0, sizeof z->S);
    z->S.phase = zsb::zs::PH_HEAD;
    z->S.d64 = (uint32_t)d64;
    z->S.wbits_arg = windowBits;
    z->S.wbits = (uint32_t)wb;
    z->hist.assign(zsb::zs::HIST, 0);
    z->carry.assign(256 << 10, 0);
    z->wrap = wrap;
    if (wrap) z->adler = (uint32_t)(wrap & 1);
    return z;
}

extern "C" void emulzs_close(void *h) { delete (EmulZs *)h; }

extern "C" int emulzs_step(void *hh, const uint8_t *in, size_t in_len_new, uint8_t *out, size_t out_cap, int flush,
                           size_t *consumed, size_t *produced, orc_result *r) {
    using namespace zsb;
    EmulZs *z = (EmulZs *)hh;
    static uint8_t fixed[lane::SLOT];
    static uint32_t lut[LUT_WORDS];
    static bool init = false;
    alignas(16) static uint8_t scr[lane::WSCR];
    if (!init) {
        init_lut(lut, 0, 1);
        alignas(16) uint8_t T0[lane::SLOT];
        for (int i = 0; i < 288; i++) T0[lane::H_LENS + i] = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
        for (int i = 0; i < 32; i++) T0[lane::H_LENS + 288 + i] = 5;
        lane::build_tables<1>(T0, scr, 288, 32, false, 0);
        memcpy(fixed, T0, sizeof fixed);
        init = true;
    }
    size_t used = 0, made = 0;
    int ret = 0;
    auto fill = [&]() {
        r->ret = ret; r->msg_id = z->msg; r->total_in = z->total_in; r->total_out = z->total_out;
        r->adler = z->adler; r->data_type = z->data_type;
        if (consumed) *consumed = used;
        if (produced) *produced = made;
        return ret;
    };
    if (z->phase == zs::PH_NEEDDICT) { ret = 2; return fill(); }
    if (z->phase == zs::PH_DONE || z->phase == zs::PH_BAD) {
        z->data_type = (int)z->last_dt;
        ret = z->phase == zs::PH_DONE ? 1 : -3;
        return fill();
    }
    const uint32_t len = z->carry_len + (uint32_t)in_len_new;
    std::vector<uint32_t> ibuf(len / 4 + 4);
    uint8_t *ib = (uint8_t *)ibuf.data();
    memcpy(ib, z->carry.data(), z->carry_len);
    if (in_len_new) memcpy(ib + z->carry_len, in, in_len_new);
    std::vector<uint32_t> wbuf((zs::HIST + out_cap) / 4 + 4);
    uint8_t *w = (uint8_t *)wbuf.data();
    memcpy(w, z->hist.data(), zs::HIST);
    const uint64_t a = out_cap / 3 + (uint64_t)len / 2 + 6, b8 = 8ull * len + 4, c2 = (uint64_t)out_cap + 2;
    const uint64_t bound = a < b8 ? (a < c2 ? a : c2) : (b8 < c2 ? b8 : c2);
    std::vector<lane::seq_t> seq(bound + 2);
    uint32_t ring[lane::RING_WORDS] = {0, 0, 0, 0};
    alignas(16) static uint32_t big_lut[lane::BIG_LUT_BYTES / 4];
    zs::stream_decode<1>(z->S, ib, len, (uint32_t)out_cap, seq.data(), fixed, scr, ring, lut, 0, big_lut,
                         flush == 5 ? zs::FL_BLOCK : flush == 6 ? zs::FL_TREES : zs::FL_NONE);
    if (z->S.call.nseq > bound) abort();
    if (z->S.call.nseq) lz::resolve_member<1>(w + zs::HIST, ib, seq.data(), z->S.call.nseq, 0);
    const uint32_t prod = z->S.call.produced;
    uint32_t check = z->S.check;
    if (z->S.kind && prod) check = z->S.kind == 2 ? emul_crc(check, w + zs::HIST, prod) : emul_adler(check, w + zs::HIST, prod);
    zs::finish_call(z->S, ib, len, check, flush == 4);
    const zs::Call c = z->S.call;
    memcpy(z->hist.data(), w + c.wprod, zs::HIST);
    if (prod) memcpy(out, w + zs::HIST, prod);
    uint32_t cpos = c.cpos;
    if (c.stop == EV_NEEDDICT) cpos = len < 6 ? len : 6;
    used = cpos > z->carry_len ? cpos - z->carry_len : 0;
    made = prod;
    const uint32_t ncarry = c.cpos > c.resume_byte ? c.cpos - c.resume_byte : 0;
    memmove(z->carry.data(), ib + c.resume_byte, ncarry);
    z->carry_len = ncarry;
    if (c.stop == EV_NEEDDICT) {
        z->adler = c.adler;
        z->phase = zs::PH_NEEDDICT;
        ret = 2;
        return fill();
    }
    z->total_in += used;
    z->total_out += made;
    if (z->wrap) z->adler = c.adler;
    z->data_type = (int)c.data_type;
    if (c.stop == EV_DONE) { ret = 1; z->phase = zs::PH_DONE; }
    else if (c.stop == EV_ERR) { ret = -3; z->msg = c.msg; z->phase = zs::PH_BAD; }
    z->last_dt = c.data_type & 0x7fu;
    if (((used == 0 && made == 0) || flush == 4) && ret == 0) ret = -5;
    return fill();
}

// what inflateMark() / inflateCodesUsed() / inflateSyncPoint() report after the last emulzs_step (zs::Call)
extern "C" void emulzs_aux(void *hh, long *out3) {
    EmulZs *z = (EmulZs *)hh;
    const bool fresh = z->total_in == 0 && z->total_out == 0 && z->S.phase == zsb::zs::PH_HEAD && z->carry_len == 0 && z->S.call.cpos == 0;
    out3[0] = fresh ? -65536 : (long)z->S.call.mark;
    out3[1] = (long)z->S.call.codes_used;
    out3[2] = (long)z->S.call.sync_point;
}

// ---------------------------------------------------------------------------------------------------
// The deflate (compression) device code (zsb_deflate.cuh) with a warp of one lane: blocks of 64 KiB, then the
// wrapper exactly as k_dstitch writes it.  Returns the stream length, or -1 when out_cap is too small.
// ---------------------------------------------------------------------------------------------------
#include "../../zlib-streams_b200/csrc/zsb_deflate.cuh"
extern "C" long emul_deflate(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits, int level) {
    using namespace zsb;
    std::vector<uint8_t> sm(df::D_BYTES + 64);
    std::vector<uint32_t> tok(df::BLK + 64);
    std::vector<uint8_t> inbuf(in_len + 16, 0);            // (word reads look up to 3 bytes past the end)
    memcpy(inbuf.data(), in, in_len);
    const uint32_t kind = wbits < 0 ? 0u : wbits > 15 ? 2u : 1u;
    std::vector<uint8_t> o;
    if (kind == 1) {
        const uint32_t fl = level < 0 ? 2u : level < 2 ? 0u : level < 6 ? 1u : level == 6 ? 2u : 3u;
        uint32_t h16 = (0x78u << 8) | (fl << 6);
        h16 += 31 - h16 % 31;
        o.push_back((uint8_t)(h16 >> 8));
        o.push_back((uint8_t)h16);
    } else if (kind == 2) {
        const uint8_t g[10] = {0x1f, 0x8b, 8, 0, 0, 0, 0, 0, (uint8_t)(level == 9 ? 2 : level == 1 ? 4 : 0), 3};
        o.insert(o.end(), g, g + 10);
    }
    const size_t nblk = in_len ? (in_len + df::BLK - 1) / df::BLK : 1;
    std::vector<uint32_t> scratch((df::BLK + 128) / 4);
    for (size_t b = 0; b < nblk; b++) {
        const uint32_t start = (uint32_t)(b * df::BLK);
        const uint32_t n = (uint32_t)(in_len - start < df::BLK ? in_len - start : df::BLK);
        memset(scratch.data(), 0xEE, scratch.size() * 4);
        const uint32_t nb = df::deflate_block<1>(sm.data(), inbuf.data(), in_len, start, n, b + 1 == nblk, level, tok.data(),
                                                 (uint8_t *)scratch.data(), 0);
        o.insert(o.end(), (uint8_t *)scratch.data(), (uint8_t *)scratch.data() + nb);
    }
    if (kind == 1) {
        const uint32_t a = emul_adler(1, in, in_len);
        for (int i = 3; i >= 0; i--) o.push_back((uint8_t)(a >> (8 * i)));
    } else if (kind == 2) {
        const uint32_t c = emul_crc(0, in, in_len);
        for (int i = 0; i < 4; i++) o.push_back((uint8_t)(c >> (8 * i)));
        for (int i = 0; i < 4; i++) o.push_back((uint8_t)((uint32_t)in_len >> (8 * i)));
    }
    if (o.size() > out_cap) return -1;
    memcpy(out, o.data(), o.size());
    return (long)o.size();
}

// the code-length builder alone (property tests of the length limits)
extern "C" void emul_build_lengths(const uint32_t *freq, uint32_t n, uint32_t maxbits, uint8_t *lens) {
    using namespace zsb;
    std::vector<uint8_t> sm(df::D_BYTES + 64);
    uint32_t *f = (uint32_t *)(sm.data() + df::D_LFREQ);
    for (uint32_t i = 0; i < n; i++) f[i] = freq[i];
    df::build_lengths<1>(sm.data() + df::D_HASH, f, n, maxbits, sm.data() + df::D_HASH + df::B_LLEN, 0);
    memcpy(lens, sm.data() + df::D_HASH + df::B_LLEN, n);
}
