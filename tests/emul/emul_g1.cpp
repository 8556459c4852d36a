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
static inline uint32_t __shfl_up_sync(unsigned, uint32_t v, int, int) { return v; }
static inline unsigned __ballot_sync(unsigned, bool p) { return p; }
static inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
#include "../../zlib-streams_b200/csrc/zsb_device.cuh"

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

// second half of the pipeline: what k_check does once the checksum of the output is known
extern "C" void emul_finish(const uint8_t *in, size_t in_len, zsb_result *res, const uint32_t *aux4,
                            uint32_t check_of_output) {
    zsb::Aux a;
    a.kind = aux4[0]; a.trailer_pos = aux4[1]; a.need_trailer = aux4[2]; a.init_check = aux4[3];
    if (a.kind == 0) return;
    uint32_t E = (uint32_t)res->total_out;
    zsb::finish_member(*res, a, in, (uint32_t)in_len, E ? check_of_output : a.init_check, E);
}
