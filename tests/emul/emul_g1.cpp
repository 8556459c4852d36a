// tests/emul/emul_g1.cpp — TEST TOOL ONLY, never part of libzsb.so.
// Compiles the *device* source of the decoder (zlib-streams_b200/csrc/zsb_device.cuh) for the host
// with the CUDA keywords and the few intrinsics it uses shimmed, instantiated for a group width of
// one lane.  It lets the CPU-only test tier (-m "not gpu") run the decoder's control logic (headers,
// table build, symbol loop, stop bookkeeping) against the oracle where no GPU exists.  The product
// has no CPU path: libzsb.so is built from the same header by nvcc and runs only on the GPU.
#include <cstdint>
#include <cstring>
#include <cstdlib>
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
static inline unsigned __ballot_sync(unsigned, bool p) { return p; }
static inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
static inline unsigned __match_any_sync(unsigned, uint32_t) { return 1u; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
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
    m.in = in; m.in_len = in_len; m.out = out; m.out_cap = out_cap; m.window_bits = wbits; m.reserved = 0;
    zsb::Aux a;
    uint32_t counter = 0;
    zsb::warp_inflate<1>(&m, res, &a, nullptr, 1, &counter, T, lut, fixed, 0, 1u);
    aux4[0] = a.kind; aux4[1] = a.trailer_pos; aux4[2] = a.need_trailer; aux4[3] = a.init_check;
    return res->ret;
}

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
    m.in = in; m.in_len = in_len; m.out = out; m.out_cap = out_cap; m.window_bits = wbits; m.reserved = 0;
    Aux a;
    uint32_t counter = 0, order = 0;
    uint64_t seq_off = 0;
    lane::seq_t *seq = (lane::seq_t *)malloc((seq_cap + 1) * sizeof(lane::seq_t));
    lane::Batch B;
    B.members = &m; B.results = res; B.aux = &a; B.order = &order; B.seq_off = &seq_off; B.seq_base = seq;
    B.hist = nullptr; B.n = 1; B.counter = &counter; B.fixed_tab = fixed;
    lane::warp_lanes<1>(B, T, scr, lut, 0);
    if (nseq_out) *nseq_out = a.nseq;
    int rc = res->ret;
    if (a.nseq > seq_cap) rc = -1000;               // the host's scratch bound was too small
    else if (a.nseq) lz::resolve_member<1>(out, in, seq, a.nseq, 0);
    free(seq);
    aux4[0] = a.kind; aux4[1] = a.trailer_pos; aux4[2] = a.need_trailer; aux4[3] = a.init_check;
    return rc;
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
