// zsb_zstream.cuh — the zlib z_stream surface of libzsb.so: inflateInit_, inflateInit2_, inflate,
// inflateEnd, inflateReset, inflateReset2, inflateResetKeep with zlib's signatures (system <zlib.h>),
// replacing reference src/inflate.c:112-229, 597-1287.  They are what src/wasm/inflate_stream_wasm.c:
// 26,33,41,49-53, src/wasm/inflate9_stream_wasm.c:19,27-30, src/wasm/wasm_stream_common.c:58 and
// test/payload_decompress.c:124,156,192 of the reference call.
//
// Every inflate() call runs on the GPU (k_zs_decode -> k_zs_lz -> k_zs_finish, see zsb_stream.cuh);
// there is no CPU decode path.  Included at the end of zsb_host.cu (one translation unit).
#pragma once
#include <zlib.h>
#include <mutex>
#include "zsb_stream.cuh"

namespace zsb {

constexpr uint32_t ZS_CARRY_LIMIT = 128u << 10;     // bytes of an unfinished unit a stream can carry to its next call

__global__ void __launch_bounds__(32) k_zs_decode(zs::State *S, const uint8_t *in, uint32_t in_len, uint32_t out_cap,
                                                  lane::seq_t *seq, const uint8_t *fixed_tab, uint32_t flush_mode) {
    __shared__ uint32_t lut[LUT_WORDS];
    __shared__ __align__(16) uint8_t scr[lane::WSCR];
    __shared__ uint32_t ring[32 * lane::RING_WORDS];
    __shared__ uint32_t big_lut[lane::BIG_LUT_BYTES / 4];
    init_lut(lut, threadIdx.x, 32);
    __syncwarp();
    zs::stream_decode<32>(*S, in, in_len, out_cap, seq, fixed_tab, scr, ring + threadIdx.x, lut, threadIdx.x, big_lut, flush_mode);
}

__global__ void __launch_bounds__(32) k_zs_lz(const zs::State *S, uint8_t *out, const uint8_t *in, const lane::seq_t *seq) {
    const uint32_t nseq = S->call.nseq;
    if (nseq) lz::resolve_member<32>(out, in, seq, nseq, threadIdx.x);
}

// Running check value, trailer (reference src/inflate.c:1194-1232), window accounting (:375-419,
// 1255-1260), the history for the next call and the carried input bytes.
// w = [HIST bytes of history | this call's output], hist = the stream's own history buffer.
__global__ void __launch_bounds__(CHK_THREADS) k_zs_finish(zs::State *S, uint8_t *in, uint32_t in_len, const uint8_t *w,
                                                           uint8_t *hist, uint8_t *carry, int finish) {
    __shared__ ChkShared sh;
    __shared__ uint32_t s_resume, s_cpos;
    const int tid = threadIdx.x;
    chk_init(sh);
    const uint32_t produced = S->call.produced, kind = S->kind;
    uint32_t check = S->check;
    if (kind != 0 && produced) check = cta_check(sh, kind, w + zs::HIST, produced, check);
    if (tid == 0) {
        zs::finish_call(*S, in, in_len, check, finish != 0);
        const zs::Call &c = S->call;
        s_resume = c.resume_byte;
        s_cpos = c.cpos;
    }
    __syncthreads();
    for (uint32_t i = tid; i < zs::HIST; i += CHK_THREADS) hist[i] = w[produced + i];
    uint32_t rb = s_resume, n = s_cpos > s_resume ? s_cpos - s_resume : 0u;
    if (n > ZS_CARRY_LIMIT) n = ZS_CARRY_LIMIT;           // the host answers Z_MEM_ERROR (a gzip header that long)
    for (uint32_t i = tid; i < n; i += CHK_THREADS) carry[i] = in[rb + i];
}

// ---- many z_streams per launch (zsb_inflate_streams): one CTA per stream ----
struct ZsJob {
    zs::State *S;
    uint8_t *in;           // [carry bytes | new bytes]: the new bytes are at in + carry_len
    uint8_t *w;            // [HIST bytes of history | this round's output]
    lane::seq_t *seq;
    uint8_t *hist, *carry;
    uint32_t carry_len, in_len, out_cap;
    int32_t finish;
    uint32_t flush_mode;   // zs::FL_*
};

__global__ void __launch_bounds__(CHK_THREADS) k_zsm_prep(const ZsJob *jobs) {
    const ZsJob j = jobs[blockIdx.x];
    for (uint32_t i = threadIdx.x; i < j.carry_len; i += CHK_THREADS) j.in[i] = j.carry[i];
    const uint4 *h4 = (const uint4 *)j.hist;
    uint4 *w4 = (uint4 *)j.w;
    for (uint32_t i = threadIdx.x; i < zs::HIST / 16; i += CHK_THREADS) w4[i] = h4[i];
}

__global__ void __launch_bounds__(32) k_zsm_decode(const ZsJob *jobs, const uint8_t *fixed_tab) {
    __shared__ uint32_t lut[LUT_WORDS];
    __shared__ __align__(16) uint8_t scr[lane::WSCR];
    __shared__ uint32_t ring[32 * lane::RING_WORDS];
    __shared__ uint32_t big_lut[lane::BIG_LUT_BYTES / 4];
    const ZsJob j = jobs[blockIdx.x];
    init_lut(lut, threadIdx.x, 32);
    __syncwarp();
    zs::stream_decode<32>(*j.S, j.in, j.in_len, j.out_cap, j.seq, fixed_tab, scr, ring + threadIdx.x, lut, threadIdx.x, big_lut, j.flush_mode);
}

__global__ void __launch_bounds__(32) k_zsm_lz(const ZsJob *jobs) {
    const ZsJob j = jobs[blockIdx.x];
    const uint32_t nseq = j.S->call.nseq;
    if (nseq) lz::resolve_member<32>(j.w + zs::HIST, j.in, j.seq, nseq, threadIdx.x);
}

__global__ void __launch_bounds__(CHK_THREADS) k_zsm_finish(const ZsJob *jobs, zs::Call *calls) {
    __shared__ ChkShared sh;
    __shared__ uint32_t s_resume, s_cpos;
    const ZsJob j = jobs[blockIdx.x];
    zs::State *S = j.S;
    const int tid = threadIdx.x;
    chk_init(sh);
    const uint32_t produced = S->call.produced, kind = S->kind;
    uint32_t check = S->check;
    if (kind != 0 && produced) check = cta_check(sh, kind, j.w + zs::HIST, produced, check);
    if (tid == 0) {
        zs::finish_call(*S, j.in, j.in_len, check, j.finish != 0);
        const zs::Call &c = S->call;
        s_resume = c.resume_byte;
        s_cpos = c.cpos;
        calls[blockIdx.x] = c;
    }
    __syncthreads();
    for (uint32_t i = tid; i < zs::HIST; i += CHK_THREADS) j.hist[i] = j.w[produced + i];
    uint32_t rb = s_resume, n = s_cpos > s_resume ? s_cpos - s_resume : 0u;
    if (n > ZS_CARRY_LIMIT) n = ZS_CARRY_LIMIT;           // the host answers Z_MEM_ERROR (a gzip header that long)
    for (uint32_t i = tid; i < n; i += CHK_THREADS) j.carry[i] = j.in[rb + i];
}

}  // namespace zsb

// ---------------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr uint32_t ZS_IN_SLICE = 1u << 20, ZS_OUT_SLICE = 1u << 20, ZS_CARRY_CAP = zsb::ZS_CARRY_LIMIT;
constexpr uint32_t ZS_MAGIC = 0x5a534221u;

// device buffers shared by all z_streams of the process (one inflate() runs at a time)
struct ZsShared {
    std::mutex mu;
    zsb_ctx *ctx = nullptr;
    cudaStream_t st = nullptr;
    uint8_t *d_in = nullptr, *d_w = nullptr;
    zsb::lane::seq_t *d_seq = nullptr;
    zsb::zs::Call *h_call = nullptr;       // pinned
    bool failed = false;
    // zsb_inflate_streams: staging for m_cap streams per round
    size_t m_cap = 0;
    uint8_t *m_in = nullptr, *m_w = nullptr;
    zsb::lane::seq_t *m_seq = nullptr;
    zsb::ZsJob *m_jobs = nullptr, *mh_jobs = nullptr;          // device / pinned
    zsb::zs::Call *m_calls = nullptr, *mh_calls = nullptr;     // device / pinned
    uint8_t *mh_in = nullptr, *mh_out = nullptr;               // pinned, compact: ZSM_IN / ZSM_OUT bytes per stream
};
ZsShared g_zs;

struct ZStream {                           // host mirror of one stream
    zsb::zs::State *d_state = nullptr;
    uint8_t *d_hist = nullptr, *d_carry = nullptr;
    uint32_t carry_len = 0;
    uint32_t phase = zsb::zs::PH_HEAD;
    uint32_t last_dt = 0;
    int wrap = 0, wbits_arg = 0;
};

// what strm->state points to; allocated with the caller's zalloc (reference src/inflate.c:210-214)
struct ZsHandle {
    z_streamp strm;
    uint32_t magic;
    ZStream *zs;
};

voidpf zs_calloc(voidpf, uInt items, uInt size) { return calloc(items, size); }
void zs_cfree(voidpf, voidpf p) { free(p); }

bool zs_shared_init() {
    if (g_zs.ctx) return true;
    if (g_zs.failed) return false;
    g_zs.ctx = zsb_create(nullptr, 0);
    if (!g_zs.ctx) {
        g_zs.failed = true;
        return false;
    }
    const size_t seqs = zsb_seq_bound_sizes(ZS_IN_SLICE + ZS_CARRY_CAP, ZS_OUT_SLICE) + 8;
    bool ok = cudaStreamCreateWithFlags(&g_zs.st, cudaStreamNonBlocking) == cudaSuccess &&
              cudaMalloc(&g_zs.d_in, ZS_IN_SLICE + ZS_CARRY_CAP + 256) == cudaSuccess &&
              cudaMalloc(&g_zs.d_w, zsb::zs::HIST + ZS_OUT_SLICE + 256) == cudaSuccess &&
              cudaMalloc(&g_zs.d_seq, seqs * sizeof(zsb::lane::seq_t)) == cudaSuccess &&
              cudaHostAlloc(&g_zs.h_call, sizeof(zsb::zs::Call), cudaHostAllocDefault) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        g_zs.failed = true;
        return false;
    }
    return true;
}

ZsHandle *zs_handle(z_streamp strm) {                // inflateStateCheck, reference src/inflate.c:100-110
    if (strm == Z_NULL || strm->zalloc == (alloc_func)0 || strm->zfree == (free_func)0) return nullptr;
    ZsHandle *h = (ZsHandle *)strm->state;
    if (h == nullptr || h->strm != strm || h->magic != ZS_MAGIC || h->zs == nullptr) return nullptr;
    return h;
}

// decode windowBits as reference src/inflate.c:147-184; false -> Z_STREAM_ERROR
bool zs_decode_wbits(int windowBits, int &wrap, int &d64, int &wbits) {
    d64 = 0;
    if (windowBits < 0) {
        if (windowBits < -16) return false;
        wrap = 0;
        d64 = windowBits == -16;
        windowBits = -windowBits;
    } else {
        wrap = (windowBits >> 4) + 5;
        if (windowBits < 48) windowBits &= 15;
    }
    if (windowBits && (windowBits < 8 || windowBits > (d64 ? 16 : 15))) return false;
    wbits = windowBits;
    return true;
}

int zs_reset_device(ZStream *z, int window_bits_arg, int d64, int wbits) {
    zsb::zs::State s0;
    memset(&s0, 0, offsetof(zsb::zs::State, tables));
    s0.phase = zsb::zs::PH_HEAD;
    s0.d64 = (uint32_t)d64;
    s0.wbits_arg = window_bits_arg;
    s0.wbits = (uint32_t)wbits;
    std::lock_guard<std::mutex> lk(g_zs.mu);
    if (cudaMemcpyAsync(z->d_state, &s0, offsetof(zsb::zs::State, tables), cudaMemcpyHostToDevice, g_zs.st) != cudaSuccess ||
        cudaStreamSynchronize(g_zs.st) != cudaSuccess) {
        cudaGetLastError();
        return Z_MEM_ERROR;
    }
    z->carry_len = 0;
    z->phase = zsb::zs::PH_HEAD;
    z->last_dt = 0;
    return Z_OK;
}

}  // namespace

extern "C" {

ZSB_API int inflateResetKeep(z_streamp strm) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    int wrap, d64, wbits;
    zs_decode_wbits(h->zs->wbits_arg, wrap, d64, wbits);
    strm->total_in = strm->total_out = 0;
    strm->msg = Z_NULL;
    if (wrap) strm->adler = (uLong)(wrap & 1);
    return zs_reset_device(h->zs, h->zs->wbits_arg, d64, wbits);
}

ZSB_API int inflateReset(z_streamp strm) { return inflateResetKeep(strm); }

ZSB_API int inflateReset2(z_streamp strm, int windowBits) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    int wrap, d64, wbits;
    if (!zs_decode_wbits(windowBits, wrap, d64, wbits)) return Z_STREAM_ERROR;
    h->zs->wrap = wrap;
    h->zs->wbits_arg = windowBits;
    return inflateReset(strm);
}

ZSB_API int inflateInit2_(z_streamp strm, int windowBits, const char *version, int stream_size) {
    if (version == Z_NULL || version[0] != ZLIB_VERSION[0] || stream_size != (int)sizeof(z_stream))
        return Z_VERSION_ERROR;
    if (strm == Z_NULL) return Z_STREAM_ERROR;
    strm->msg = Z_NULL;
    if (strm->zalloc == (alloc_func)0) {
        strm->zalloc = zs_calloc;
        strm->opaque = (voidpf)0;
    }
    if (strm->zfree == (free_func)0) strm->zfree = zs_cfree;
    ZsHandle *h = (ZsHandle *)strm->zalloc(strm->opaque, 1, sizeof(ZsHandle));
    if (h == nullptr) return Z_MEM_ERROR;
    h->strm = strm;
    h->magic = ZS_MAGIC;
    h->zs = nullptr;
    strm->state = (struct internal_state *)h;
    int wrap, d64, wbits;
    int ret = zs_decode_wbits(windowBits, wrap, d64, wbits) ? Z_OK : Z_STREAM_ERROR;
    if (ret == Z_OK) {
        bool ok;
        {
            std::lock_guard<std::mutex> lk(g_zs.mu);
            ok = zs_shared_init();
        }
        ZStream *z = ok ? new (std::nothrow) ZStream() : nullptr;
        if (z) {
            ok = cudaMalloc(&z->d_state, sizeof(zsb::zs::State)) == cudaSuccess &&
                 cudaMalloc(&z->d_hist, zsb::zs::HIST) == cudaSuccess &&
                 cudaMalloc(&z->d_carry, ZS_CARRY_CAP) == cudaSuccess;
            if (!ok) {
                cudaGetLastError();
                cudaFree(z->d_state);
                cudaFree(z->d_hist);
                cudaFree(z->d_carry);
                delete z;
                z = nullptr;
            }
        }
        if (!z) ret = Z_MEM_ERROR;      // no usable CUDA device or out of device memory: there is no CPU path
        else {
            h->zs = z;
            ret = inflateReset2(strm, windowBits);
        }
    }
    if (ret != Z_OK) {
        if (h->zs) {
            cudaFree(h->zs->d_state);
            cudaFree(h->zs->d_hist);
            cudaFree(h->zs->d_carry);
            delete h->zs;
        }
        strm->zfree(strm->opaque, h);
        strm->state = Z_NULL;
    }
    return ret;
}

ZSB_API int inflateInit_(z_streamp strm, const char *version, int stream_size) {
    return inflateInit2_(strm, MAX_WBITS, version, stream_size);
}

ZSB_API int inflateEnd(z_streamp strm) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    {
        std::lock_guard<std::mutex> lk(g_zs.mu);
        cudaFree(h->zs->d_state);
        cudaFree(h->zs->d_hist);
        cudaFree(h->zs->d_carry);
    }
    delete h->zs;
    h->magic = 0;
    strm->zfree(strm->opaque, h);
    strm->state = Z_NULL;
    return Z_OK;
}

}  // extern "C"

namespace {
// what inflate() leaves on the z_stream once its rounds are over (reference src/inflate.c:1253-1274); c is the
// last round's record
int zs_exit_bookkeeping(z_streamp strm, ZStream *z, const zsb::zs::Call &c, uInt in0, uInt out0, int flush) {
    using namespace zsb;
    int ret = Z_OK;
    const uInt in_used = in0 - strm->avail_in, out_used = out0 - strm->avail_out;
    if (c.stop == EV_NEEDDICT) {                                // returns before the exit bookkeeping
        strm->adler = c.adler;
        z->phase = zs::PH_NEEDDICT;
        return Z_NEED_DICT;
    }
    strm->total_in += in_used;
    strm->total_out += out_used;
    if (z->wrap) strm->adler = c.adler;
    strm->data_type = (int)c.data_type;
    if (c.stop == EV_DONE) {
        ret = Z_STREAM_END;
        z->phase = zs::PH_DONE;
    } else if (c.stop == EV_ERR || c.stop == EV_INITERR) {
        ret = Z_DATA_ERROR;
        strm->msg = (char *)zsb_message(c.msg);
        z->phase = zs::PH_BAD;
    }
    z->last_dt = c.data_type & 0x7fu;
    if (((in_used == 0 && out_used == 0) || flush == Z_FINISH) && ret == Z_OK) ret = Z_BUF_ERROR;
    return ret;
}
}  // namespace

// inflate(): reference src/inflate.c:597-1275, including the early returns of Z_BLOCK / Z_TREES
// (:832,854-857,883,1032) and the bits they add to data_type (:1269-1271).
extern "C" ZSB_API int inflate(z_streamp strm, int flush) {
    using namespace zsb;
    ZsHandle *h = zs_handle(strm);
    if (!h || strm->next_out == Z_NULL || (strm->next_in == Z_NULL && strm->avail_in != 0)) return Z_STREAM_ERROR;
    ZStream *z = h->zs;
    if (z->phase == zs::PH_NEEDDICT) return Z_NEED_DICT;        // src/inflate.c:824-827: no bookkeeping
    if (z->phase == zs::PH_DONE || z->phase == zs::PH_BAD) {    // sticky states, src/inflate.c:1233-1240
        strm->data_type = (int)z->last_dt;
        return z->phase == zs::PH_DONE ? Z_STREAM_END : Z_DATA_ERROR;
    }
    std::lock_guard<std::mutex> lk(g_zs.mu);
    int cur = 0;
    cudaGetDevice(&cur);
    const int dev = g_zs.ctx->devs[0].device;
    if (cur != dev) cudaSetDevice(dev);
    cudaStream_t st = g_zs.st;
    const uInt in0 = strm->avail_in, out0 = strm->avail_out;
    int ret = Z_OK;
    zs::Call c;
    memset(&c, 0, sizeof c);
    bool cuda_ok = true;
    for (;;) {
        const uint32_t ni = strm->avail_in < ZS_IN_SLICE ? strm->avail_in : ZS_IN_SLICE;
        const uint32_t no = strm->avail_out < ZS_OUT_SLICE ? strm->avail_out : ZS_OUT_SLICE;
        if (z->carry_len > ZS_CARRY_CAP) {
            ret = Z_MEM_ERROR;                                  // a wrapper header longer than the carry buffer
            break;
        }
        const uint32_t in_len = z->carry_len + ni;
        cuda_ok = cuda_ok && (z->carry_len == 0 ||
                  cudaMemcpyAsync(g_zs.d_in, z->d_carry, z->carry_len, cudaMemcpyDeviceToDevice, st) == cudaSuccess);
        cuda_ok = cuda_ok && (ni == 0 ||
                  cudaMemcpyAsync(g_zs.d_in + z->carry_len, strm->next_in, ni, cudaMemcpyHostToDevice, st) == cudaSuccess);
        cuda_ok = cuda_ok && cudaMemcpyAsync(g_zs.d_w, z->d_hist, zs::HIST, cudaMemcpyDeviceToDevice, st) == cudaSuccess;
        if (!cuda_ok) break;
        k_zs_decode<<<1, 32, 0, st>>>(z->d_state, g_zs.d_in, in_len, no, g_zs.d_seq, g_zs.ctx->devs[0].fixed_lane,
                                      flush == Z_BLOCK ? zs::FL_BLOCK : flush == Z_TREES ? zs::FL_TREES : zs::FL_NONE);
        k_zs_lz<<<1, 32, 0, st>>>(z->d_state, g_zs.d_w + zs::HIST, g_zs.d_in, g_zs.d_seq);
        k_zs_finish<<<1, CHK_THREADS, 0, st>>>(z->d_state, g_zs.d_in, in_len, g_zs.d_w, z->d_hist, z->d_carry,
                                               flush == Z_FINISH);
        cuda_ok = cudaGetLastError() == cudaSuccess &&
                  cudaMemcpyAsync(g_zs.h_call, &z->d_state->call, sizeof(zs::Call), cudaMemcpyDeviceToHost, st) == cudaSuccess &&
                  cudaStreamSynchronize(st) == cudaSuccess;
        if (!cuda_ok) break;
        c = *g_zs.h_call;
        if (c.produced) {
            cuda_ok = cudaMemcpyAsync(strm->next_out, g_zs.d_w + zs::HIST, c.produced, cudaMemcpyDeviceToHost, st) == cudaSuccess &&
                      cudaStreamSynchronize(st) == cudaSuccess;
            if (!cuda_ok) break;
        }
        uint32_t cpos = c.cpos;
        if (c.stop == EV_NEEDDICT) cpos = in_len < 6 ? in_len : 6;   // header + dictionary id were read
        const uint32_t used = cpos > z->carry_len ? cpos - z->carry_len : 0;
        strm->next_in += used;
        strm->avail_in -= used;
        strm->next_out += c.produced;
        strm->avail_out -= c.produced;
        z->carry_len = c.cpos > c.resume_byte ? c.cpos - c.resume_byte : 0;
        if (c.stop == EV_STARVE && strm->avail_in > 0) continue;
        if (c.stop == EV_OUTFULL && strm->avail_out > 0) continue;
        break;
    }
    if (cur != dev) cudaSetDevice(cur);
    if (!cuda_ok) {
        cudaGetLastError();
        return Z_MEM_ERROR;
    }
    if (ret == Z_MEM_ERROR) return ret;
    return zs_exit_bookkeeping(strm, z, c, in0, out0, flush);
}

// ---------------------------------------------------------------------------------------------------
// zsb_inflate_streams(): inflate(strms[i], flush) for n streams, with every round of GPU work (stage, decode,
// resolve, finish) done for all of them by ONE set of launches, one CTA per stream.  Per stream the result is
// exactly what inflate() gives: same rounds, same bookkeeping (zs_exit_bookkeeping).  This is the shape of
// the reference's real caller — many TransformStreams each feeding <= 32 KiB in / 64 KiB out per call
// (src/wasm/api/zlib-streams.js:64-138) — where a launch and a PCIe round trip per stream and call would
// dominate.  
// ---------------------------------------------------------------------------------------------------
namespace {
constexpr uint32_t ZSM_IN = 64u << 10, ZSM_OUT = 128u << 10;                 // per stream and round
constexpr size_t ZSM_IN_STRIDE = ZS_CARRY_CAP + ZSM_IN + 256, ZSM_W_STRIDE = zsb::zs::HIST + ZSM_OUT + 256;

bool zsm_reserve(size_t n) {
    if (n <= g_zs.m_cap) return true;
    const size_t cap = n + n / 4 + 16;
    const size_t seqs = zsb_seq_bound_sizes(ZSM_IN + ZS_CARRY_CAP, ZSM_OUT) + 8;
    cudaFree(g_zs.m_in); cudaFree(g_zs.m_w); cudaFree(g_zs.m_seq); cudaFree(g_zs.m_jobs); cudaFree(g_zs.m_calls);
    cudaFreeHost(g_zs.mh_jobs); cudaFreeHost(g_zs.mh_calls); cudaFreeHost(g_zs.mh_in); cudaFreeHost(g_zs.mh_out);
    g_zs.m_cap = 0;
    g_zs.m_in = g_zs.m_w = nullptr; g_zs.m_seq = nullptr; g_zs.m_jobs = g_zs.mh_jobs = nullptr;
    g_zs.m_calls = g_zs.mh_calls = nullptr; g_zs.mh_in = g_zs.mh_out = nullptr;
    const bool ok = cudaMalloc(&g_zs.m_in, cap * ZSM_IN_STRIDE) == cudaSuccess &&
                    cudaMalloc(&g_zs.m_w, cap * ZSM_W_STRIDE) == cudaSuccess &&
                    cudaMalloc(&g_zs.m_seq, cap * seqs * sizeof(zsb::lane::seq_t)) == cudaSuccess &&
                    cudaMalloc(&g_zs.m_jobs, cap * sizeof(zsb::ZsJob)) == cudaSuccess &&
                    cudaMalloc(&g_zs.m_calls, cap * sizeof(zsb::zs::Call)) == cudaSuccess &&
                    cudaHostAlloc(&g_zs.mh_jobs, cap * sizeof(zsb::ZsJob), cudaHostAllocDefault) == cudaSuccess &&
                    cudaHostAlloc(&g_zs.mh_calls, cap * sizeof(zsb::zs::Call), cudaHostAllocDefault) == cudaSuccess &&
                    cudaHostAlloc(&g_zs.mh_in, cap * (size_t)ZSM_IN, cudaHostAllocDefault) == cudaSuccess &&
                    cudaHostAlloc(&g_zs.mh_out, cap * (size_t)ZSM_OUT, cudaHostAllocDefault) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        return false;
    }
    g_zs.m_cap = cap;
    return true;
}
}  // namespace

extern "C" ZSB_API int zsb_inflate_streams(size_t n, void **strms_v, int flush, int *rets) {
    using namespace zsb;
    z_streamp *strms = (z_streamp *)strms_v;
    if (n && (!strms || !rets)) return ZSB_E_ARG;
    struct Live {
        size_t i;
        ZStream *z;
        uInt in0, out0;
        zs::Call c;
    };
    std::vector<Live> live;
    for (size_t i = 0; i < n; i++) {
        z_streamp strm = strms[i];
        ZsHandle *h = zs_handle(strm);
        if (!h || strm->next_out == Z_NULL || (strm->next_in == Z_NULL && strm->avail_in != 0)) {
            rets[i] = Z_STREAM_ERROR;
            continue;
        }
        ZStream *z = h->zs;
        if (z->phase == zs::PH_NEEDDICT) {
            rets[i] = Z_NEED_DICT;
            continue;
        }
        if (z->phase == zs::PH_DONE || z->phase == zs::PH_BAD) {
            strm->data_type = (int)z->last_dt;
            rets[i] = z->phase == zs::PH_DONE ? Z_STREAM_END : Z_DATA_ERROR;
            continue;
        }
        if (z->carry_len > ZS_CARRY_CAP) {                      // a wrapper header longer than the carry buffer
            rets[i] = Z_MEM_ERROR;
            continue;
        }
        Live l;
        l.i = i; l.z = z; l.in0 = strm->avail_in; l.out0 = strm->avail_out;
        memset(&l.c, 0, sizeof l.c);
        live.push_back(l);
        rets[i] = Z_OK;
    }
    if (live.empty()) return 0;
    std::lock_guard<std::mutex> lk(g_zs.mu);
    if (!zsm_reserve(live.size())) {
        for (const Live &l : live) rets[l.i] = Z_MEM_ERROR;
        return 0;
    }
    int cur = 0;
    cudaGetDevice(&cur);
    const int dev = g_zs.ctx->devs[0].device;
    if (cur != dev) cudaSetDevice(dev);
    cudaStream_t st = g_zs.st;
    const size_t seqs = zsb_seq_bound_sizes(ZSM_IN + ZS_CARRY_CAP, ZSM_OUT) + 8;
    std::vector<size_t> round(live.size());           // indices into live of the streams of this round
    for (size_t k = 0; k < live.size(); k++) round[k] = k;
    bool cuda_ok = true;
    while (!round.empty() && cuda_ok) {
        const size_t m = round.size();
        std::vector<uint32_t> nis(m);
        for (size_t k = 0; k < m; k++) {
            Live &l = live[round[k]];
            z_streamp strm = strms[l.i];
            const uint32_t ni = strm->avail_in < ZSM_IN ? strm->avail_in : ZSM_IN;
            const uint32_t no = strm->avail_out < ZSM_OUT ? strm->avail_out : ZSM_OUT;
            nis[k] = ni;
            if (ni) memcpy(g_zs.mh_in + k * (size_t)ZSM_IN, strm->next_in, ni);
            ZsJob &j = g_zs.mh_jobs[k];
            j.S = l.z->d_state;
            j.in = g_zs.m_in + k * ZSM_IN_STRIDE + ZS_CARRY_CAP - l.z->carry_len;   // the carry sits right before the new bytes
            j.w = g_zs.m_w + k * ZSM_W_STRIDE;
            j.seq = g_zs.m_seq + k * seqs;
            j.hist = l.z->d_hist;
            j.carry = l.z->d_carry;
            j.carry_len = l.z->carry_len;
            j.in_len = l.z->carry_len + ni;
            j.out_cap = no;
            j.finish = flush == Z_FINISH;
            j.flush_mode = flush == Z_BLOCK ? zs::FL_BLOCK : flush == Z_TREES ? zs::FL_TREES : zs::FL_NONE;
        }
        cuda_ok = cudaMemcpyAsync(g_zs.m_jobs, g_zs.mh_jobs, m * sizeof(ZsJob), cudaMemcpyHostToDevice, st) == cudaSuccess &&
                  cudaMemcpy2DAsync(g_zs.m_in + ZS_CARRY_CAP, ZSM_IN_STRIDE, g_zs.mh_in, ZSM_IN, ZSM_IN, m,
                                    cudaMemcpyHostToDevice, st) == cudaSuccess;
        if (!cuda_ok) break;
        k_zsm_prep<<<(unsigned)m, CHK_THREADS, 0, st>>>(g_zs.m_jobs);
        k_zsm_decode<<<(unsigned)m, 32, 0, st>>>(g_zs.m_jobs, g_zs.ctx->devs[0].fixed_lane);
        k_zsm_lz<<<(unsigned)m, 32, 0, st>>>(g_zs.m_jobs);
        k_zsm_finish<<<(unsigned)m, CHK_THREADS, 0, st>>>(g_zs.m_jobs, g_zs.m_calls);
        cuda_ok = cudaGetLastError() == cudaSuccess &&
                  cudaMemcpyAsync(g_zs.mh_calls, g_zs.m_calls, m * sizeof(zs::Call), cudaMemcpyDeviceToHost, st) == cudaSuccess &&
                  cudaStreamSynchronize(st) == cudaSuccess;
        if (!cuda_ok) break;
        uint32_t widest = 0;
        for (size_t k = 0; k < m; k++) widest = std::max(widest, g_zs.mh_calls[k].produced);
        if (widest) {
            cuda_ok = cudaMemcpy2DAsync(g_zs.mh_out, ZSM_OUT, g_zs.m_w + zs::HIST, ZSM_W_STRIDE, widest, m,
                                        cudaMemcpyDeviceToHost, st) == cudaSuccess &&
                      cudaStreamSynchronize(st) == cudaSuccess;
            if (!cuda_ok) break;
        }
        std::vector<size_t> next;
        for (size_t k = 0; k < m; k++) {
            Live &l = live[round[k]];
            z_streamp strm = strms[l.i];
            ZStream *z = l.z;
            const zs::Call c = g_zs.mh_calls[k];
            l.c = c;
            if (c.produced) memcpy(strm->next_out, g_zs.mh_out + k * (size_t)ZSM_OUT, c.produced);
            const uint32_t in_len = z->carry_len + nis[k];
            uint32_t cpos = c.cpos;
            if (c.stop == EV_NEEDDICT) cpos = in_len < 6 ? in_len : 6;   // header + dictionary id were read
            const uint32_t used = cpos > z->carry_len ? cpos - z->carry_len : 0;
            strm->next_in += used;
            strm->avail_in -= used;
            strm->next_out += c.produced;
            strm->avail_out -= c.produced;
            z->carry_len = c.cpos > c.resume_byte ? c.cpos - c.resume_byte : 0;
            if ((c.stop == EV_STARVE && strm->avail_in > 0) || (c.stop == EV_OUTFULL && strm->avail_out > 0)) {
                if (z->carry_len > ZS_CARRY_CAP) rets[l.i] = Z_MEM_ERROR;     // a wrapper header longer than the carry buffer
                else next.push_back(round[k]);
            }
        }
        round.swap(next);
    }
    if (cur != dev) cudaSetDevice(cur);
    if (!cuda_ok) cudaGetLastError();
    for (Live &l : live) {
        if (!cuda_ok) rets[l.i] = Z_MEM_ERROR;
        else if (rets[l.i] != Z_MEM_ERROR) rets[l.i] = zs_exit_bookkeeping(strms[l.i], l.z, l.c, l.in0, l.out0, flush);
    }
    return 0;
}
