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

__global__ void __launch_bounds__(32) k_zs_decode(zs::State *S, const uint8_t *in, uint32_t in_len, uint32_t out_cap,
                                                  lane::seq_t *seq, const uint8_t *fixed_tab) {
    __shared__ uint32_t lut[LUT_WORDS];
    __shared__ __align__(16) uint8_t scr[lane::WSCR];
    __shared__ uint32_t ring[32 * lane::RING_WORDS];
    init_lut(lut, threadIdx.x, 32);
    __syncwarp();
    zs::stream_decode<32>(*S, in, in_len, out_cap, seq, fixed_tab, scr, ring + threadIdx.x, lut, threadIdx.x);
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
    const uint32_t rb = s_resume, n = s_cpos > s_resume ? s_cpos - s_resume : 0u;
    for (uint32_t i = tid; i < n; i += CHK_THREADS) carry[i] = in[rb + i];
}

}  // namespace zsb

// ---------------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr uint32_t ZS_IN_SLICE = 1u << 20, ZS_OUT_SLICE = 1u << 20, ZS_CARRY_CAP = 128u << 10;
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

// inflate(): reference src/inflate.c:597-1275.  Z_BLOCK / Z_TREES are accepted and treated as
// Z_NO_FLUSH (they only add early returns the reference's callers never ask for).
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
        if (z->carry_len + ni > ZS_IN_SLICE + ZS_CARRY_CAP) {
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
        k_zs_decode<<<1, 32, 0, st>>>(z->d_state, g_zs.d_in, in_len, no, g_zs.d_seq, g_zs.ctx->devs[0].fixed_lane);
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
