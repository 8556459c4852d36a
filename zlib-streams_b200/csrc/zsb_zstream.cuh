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
    const uint32_t wprod = S->call.wprod;        // (written by thread 0 before the barrier above)
    for (uint32_t i = tid; i < zs::HIST; i += CHK_THREADS) hist[i] = w[wprod + i];
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
    const uint32_t wprod = S->call.wprod;
    for (uint32_t i = tid; i < zs::HIST; i += CHK_THREADS) j.hist[i] = j.w[wprod + i];
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
    // ---- what the auxiliary entry points need (reference src/inflate.c:231-258, 1289-1537) ----
    std::vector<uint8_t> carry_shadow;     // host copy of the carried bytes (the device's d_carry): inflateSync looks into
                                           // the held bytes, inflateGetHeader at the header, inflatePrime appends to them
    uint32_t bit0 = 0;                     // bit offset of the resume point inside carry_shadow[0]
    uint32_t held_bits = 0;                // the reference's state->bits at the last return
    bool header_seen = false;              // state->flags != -1
    uint32_t kind = 0;                     // 0 raw / unknown yet, 1 zlib, 2 gzip
    gz_headerp head = nullptr;             // inflateGetHeader()
    uint32_t dictid = 0;                   // Adler-32 the stream asked for with Z_NEED_DICT
    long mark = -65536;                    // inflateMark()
    unsigned long codes_used = 0;          // inflateCodesUsed()
    uint32_t sync_point = 0, whave = 0;
    bool in_sync = false;                  // mode SYNC: inflateSync() has started a search
    uint32_t sync_have = 0;
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

// keep_window: inflateResetKeep() leaves the sliding window alone (src/inflate.c:112-133); inflateReset() empties it
int zs_reset_device(ZStream *z, int window_bits_arg, int d64, int wbits, bool keep_window = false) {
    zsb::zs::State s0;
    memset(&s0, 0, offsetof(zsb::zs::State, tables));
    s0.phase = zsb::zs::PH_HEAD;
    s0.d64 = (uint32_t)d64;
    s0.wbits_arg = window_bits_arg;
    s0.wbits = (uint32_t)wbits;
    s0.back = -1;
    const uint32_t whave = keep_window ? z->whave : 0u;
    s0.whave = whave;
    s0.wsize = whave ? 1u << (wbits ? wbits : 15) : 0u;
    s0.noverify = (z->wrap && !(z->wrap & 4)) ? 1u : 0u;     // inflateValidate(strm, 0) survives inflateReset (state->wrap does)
    std::lock_guard<std::mutex> lk(g_zs.mu);
    if (cudaMemcpyAsync(z->d_state, &s0, offsetof(zsb::zs::State, tables), cudaMemcpyHostToDevice, g_zs.st) != cudaSuccess ||
        cudaStreamSynchronize(g_zs.st) != cudaSuccess) {
        cudaGetLastError();
        return Z_MEM_ERROR;
    }
    z->carry_len = 0;
    z->phase = zsb::zs::PH_HEAD;
    z->last_dt = 0;
    z->carry_shadow.clear();
    z->head = nullptr;                     // inflateResetKeep drops the registered gz_header (src/inflate.c:125)
    z->bit0 = 0; z->held_bits = 0; z->header_seen = false; z->kind = 0; z->dictid = 0; z->mark = -65536;
    z->codes_used = 0; z->sync_point = 0; z->whave = whave; z->in_sync = false; z->sync_have = 0;
    return Z_OK;
}

// Adler-32 (RFC 1950) of a dictionary on the host: inflateSetDictionary()'s identifier check
uint32_t zs_adler32(const uint8_t *p, size_t n) {
    uint32_t a = 1, b = 0;
    while (n) {
        size_t k = n < 5552 ? n : 5552;
        n -= k;
        while (k--) {
            a += *p++;
            b += a;
        }
        a %= 65521u;
        b %= 65521u;
    }
    return (b << 16) | a;
}

// After one round of GPU work: advance the caller's buffers, keep the host copy of the carried bytes, fill the
// registered gz_header when the header was completed, mirror what the auxiliary entry points report.
// ni = new input bytes offered this round (from strm->next_in).
void zs_after_round(z_streamp strm, ZStream *z, const zsb::zs::Call &c, uint32_t ni) {
    using namespace zsb;
    const uint32_t old_carry = z->carry_len, in_len = old_carry + ni;
    uint32_t cpos = c.cpos;
    if (c.stop == EV_NEEDDICT) cpos = in_len < 6 ? in_len : 6;      // header + dictionary id were read
    auto byte_at = [&](uint32_t i) -> uint8_t { return i < old_carry ? z->carry_shadow[i] : strm->next_in[i - old_carry]; };
    if (!z->header_seen && z->wrap && in_len >= 2 && !(((z->wrap & 2) && byte_at(0) == 0x1f && byte_at(1) == 0x8b))) {
        // src/inflate.c:660-664: two bytes that are not the gzip magic are taken for a zlib header, right or wrong
        z->header_seen = true;
        z->kind = 1;
        if (z->head) z->head->done = -1;
    }
    if (c.hdr_len) {
        z->header_seen = true;
        z->kind = c.hdr_kind;
        if (z->head) {
            if (c.hdr_kind == 2) {
                std::vector<uint8_t> hb(c.hdr_len);
                for (uint32_t i = 0; i < c.hdr_len; i++) hb[i] = byte_at(i);
                zsb_gz_header g;
                if (zsb_gzip_header(hb.data(), hb.size(), &g) == 0) {          // reference src/inflate.c:678-816
                    gz_headerp hd = z->head;
                    hd->text = (int)g.text;
                    hd->time = g.time;
                    hd->xflags = (int)g.xflags;
                    hd->os = (int)g.os;
                    const uint32_t flg = hb[3];
                    if (flg & 4) hd->extra_len = g.extra_len;
                    if ((flg & 4) && hd->extra != Z_NULL && g.extra_len)
                        memcpy(hd->extra, hb.data() + g.extra_off, std::min<uInt>(g.extra_len, hd->extra_max));
                    auto text_field = [&](Bytef *dst, uInt dst_max, uint32_t off, uint32_t len, bool present) {
                        if (dst == Z_NULL) return;
                        if (!present) return;
                        // the field with its terminating NUL, cut to dst_max (src/inflate.c:757-783)
                        const uInt n = std::min<uInt>(len + 1, dst_max);
                        if (n) memcpy(dst, hb.data() + (len ? off : 0), len ? std::min<uInt>(len, n) : 0);
                        if (len < n) dst[len] = 0;
                    };
                    if (flg & 8) text_field(hd->name, hd->name_max, g.name_off, g.name_len, true);
                    else hd->name = Z_NULL;
                    if (flg & 16) text_field(hd->comment, hd->comm_max, g.comment_off, g.comment_len, true);
                    else hd->comment = Z_NULL;
                    if (!(flg & 4)) hd->extra = Z_NULL;
                    hd->hcrc = (int)((flg >> 1) & 1);
                    hd->done = 1;
                }
            }
        }
    }
    const uint32_t used = cpos > old_carry ? cpos - old_carry : 0;
    const uint32_t ncarry = c.cpos > c.resume_byte ? c.cpos - c.resume_byte : 0;
    if (ncarry <= ZS_CARRY_CAP) {
        std::vector<uint8_t> nc(ncarry);
        for (uint32_t i = 0; i < ncarry; i++) nc[i] = byte_at(c.resume_byte + i);
        z->carry_shadow.swap(nc);
    }
    strm->next_in += used;
    strm->avail_in -= used;
    strm->next_out += c.produced;
    strm->avail_out -= c.produced;
    z->carry_len = ncarry;
    z->bit0 = c.bit0;
    z->held_bits = c.data_type & 63u;
    z->mark = (long)c.mark;
    z->codes_used = c.codes_used;
    z->sync_point = c.sync_point;
    z->whave = c.whave;
    z->in_sync = false;
    if (c.stop == EV_NEEDDICT) z->dictid = c.adler;
}

}  // namespace

extern "C" {

ZSB_API int inflateResetKeep(z_streamp strm) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    int wrap, d64, wbits;
    zs_decode_wbits(h->zs->wbits_arg, wrap, d64, wbits);
    wrap = h->zs->wrap;                                 // (inflateValidate / inflateSync may have changed it)
    strm->total_in = strm->total_out = 0;
    strm->msg = Z_NULL;
    if (wrap) strm->adler = (uLong)(wrap & 1);
    return zs_reset_device(h->zs, h->zs->wbits_arg, d64, wbits, true);
}

ZSB_API int inflateReset(z_streamp strm) {              // src/inflate.c:135-145: the window is emptied first
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    h->zs->whave = 0;
    return inflateResetKeep(strm);
}

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
        zs_after_round(strm, z, c, ni);
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
            zs_after_round(strm, z, c, nis[k]);
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


// ---------------------------------------------------------------------------------------------------
// The remaining zlib entry points of the inflate side (reference src/inflate.c:231-258,
// 1289-1537).  The stream lives on the device as {phase, resume bit, carried bytes, history}; these functions
// edit that state (small kernels / copies on the z_stream surface's own CUDA stream) and the host mirror.
// Where the reference's bit-granular state cannot be expressed (inflatePrime in the middle of a block header
// or symbol) the call answers Z_STREAM_ERROR; include/zsb.h lists those cases.
// ---------------------------------------------------------------------------------------------------
namespace zsb {
// history <- (history ++ dict)[-HIST:], and the window accounting of updatewindow() (src/inflate.c:375-419)
__global__ void __launch_bounds__(CHK_THREADS) k_zs_setdict(zs::State *S, uint8_t *hist, const uint8_t *dict, uint32_t n, int leave_dict_mode) {
    const int tid = threadIdx.x;
    if (n >= zs::HIST) {
        for (uint32_t i = tid; i < zs::HIST; i += CHK_THREADS) hist[i] = dict[n - zs::HIST + i];
    } else {
        // shift down by n in chunks (a chunk is read completely before it is written)
        for (uint32_t c0 = 0; c0 < zs::HIST - n; c0 += CHK_THREADS) {
            const uint32_t i = c0 + tid;
            uint8_t v = 0;
            if (i < zs::HIST - n) v = hist[i + n];
            __syncthreads();
            if (i < zs::HIST - n) hist[i] = v;
            __syncthreads();
        }
        for (uint32_t i = tid; i < n; i += CHK_THREADS) hist[zs::HIST - n + i] = dict[i];
    }
    if (tid == 0) {
        S->wsize = 1u << (S->wbits ? S->wbits : 15u);
        const uint64_t h = (uint64_t)S->whave + n;
        S->whave = h > S->wsize ? S->wsize : (uint32_t)h;
        S->call.whave = S->whave;
        if (leave_dict_mode) {               // mode DICT -> the next inflate() enters TYPE (src/inflate.c:822-830)
            S->phase = zs::PH_DICTSET;
            S->kind = 1;
            S->check = 1;                    // adler32(0L, Z_NULL, 0)
            S->bit0 = 0;
        }
    }
}
}  // namespace zsb

namespace {
// one word of the device State
template <typename T> bool zs_poke(ZStream *z, size_t offset, const T &v) {
    return cudaMemcpyAsync((uint8_t *)z->d_state + offset, &v, sizeof(T), cudaMemcpyHostToDevice, g_zs.st) == cudaSuccess &&
           cudaStreamSynchronize(g_zs.st) == cudaSuccess;
}
struct ZsDeviceGuard {                     // the z_stream surface runs on the context's first device
    int cur = 0, dev = 0;
    ZsDeviceGuard() {
        cudaGetDevice(&cur);
        dev = g_zs.ctx->devs[0].device;
        if (cur != dev) cudaSetDevice(dev);
    }
    ~ZsDeviceGuard() {
        if (cur != dev) cudaSetDevice(cur);
    }
};
}  // namespace

extern "C" {

// reference src/inflate.c:1309-1339
ZSB_API int inflateSetDictionary(z_streamp strm, const Bytef *dictionary, uInt dictLength) {
    using namespace zsb;
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    ZStream *z = h->zs;
    if (z->wrap != 0 && z->phase != zs::PH_NEEDDICT) return Z_STREAM_ERROR;
    if (z->phase == zs::PH_NEEDDICT && zs_adler32(dictionary, dictLength) != z->dictid) return Z_DATA_ERROR;
    std::lock_guard<std::mutex> lk(g_zs.mu);
    ZsDeviceGuard dg;
    // only the last 64 KiB can ever be referenced
    const uint32_t n = dictLength > zs::HIST ? zs::HIST : dictLength;
    const Bytef *src = dictionary + (dictLength - n);
    bool ok = n == 0 || cudaMemcpyAsync(g_zs.d_in, src, n, cudaMemcpyHostToDevice, g_zs.st) == cudaSuccess;
    if (ok) {
        // (dictLength, not n, feeds the accounting: whave saturates at the window size either way)
        k_zs_setdict<<<1, CHK_THREADS, 0, g_zs.st>>>(z->d_state, z->d_hist, g_zs.d_in, n, z->phase == zs::PH_NEEDDICT);
        ok = cudaGetLastError() == cudaSuccess &&
             cudaMemcpyAsync(g_zs.h_call, &z->d_state->call, sizeof(zs::Call), cudaMemcpyDeviceToHost, g_zs.st) == cudaSuccess &&
             cudaStreamSynchronize(g_zs.st) == cudaSuccess;
    }
    if (!ok) {
        cudaGetLastError();
        z->phase = zs::PH_BAD;
        return Z_MEM_ERROR;
    }
    z->whave = g_zs.h_call->whave;
    if (z->phase == zs::PH_NEEDDICT) {
        z->phase = zs::PH_DICTSET;
        z->kind = 1;
        z->header_seen = true;
    }
    return Z_OK;
}

// reference src/inflate.c:1289-1307
ZSB_API int inflateGetDictionary(z_streamp strm, Bytef *dictionary, uInt *dictLength) {
    using namespace zsb;
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    ZStream *z = h->zs;
    if (z->whave && dictionary != Z_NULL) {
        std::lock_guard<std::mutex> lk(g_zs.mu);
        ZsDeviceGuard dg;
        if (cudaMemcpyAsync(dictionary, z->d_hist + (zs::HIST - z->whave), z->whave, cudaMemcpyDeviceToHost, g_zs.st) != cudaSuccess ||
            cudaStreamSynchronize(g_zs.st) != cudaSuccess) {
            cudaGetLastError();
            return Z_MEM_ERROR;
        }
    }
    if (dictLength != Z_NULL) *dictLength = z->whave;
    return Z_OK;
}

// reference src/inflate.c:1341-1354
ZSB_API int inflateGetHeader(z_streamp strm, gz_headerp head) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    if ((h->zs->wrap & 2) == 0) return Z_STREAM_ERROR;
    h->zs->head = head;
    head->done = 0;
    return Z_OK;
}

// reference src/inflate.c:1386-1434 (syncsearch :1368-1384)
ZSB_API int inflateSync(z_streamp strm) {
    using namespace zsb;
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    ZStream *z = h->zs;
    if (strm->avail_in == 0 && z->held_bits < 8) return Z_BUF_ERROR;
    auto search = [&](const uint8_t *buf, unsigned len) {
        unsigned got = z->sync_have, next = 0;
        while (next < len && got < 4) {
            if (buf[next] == (got < 2 ? 0 : 0xff)) got++;
            else if (buf[next]) got = 0;
            else got = 4 - got;
            next++;
        }
        z->sync_have = got;
        return next;
    };
    if (!z->in_sync) {                       // first time: the whole bytes still in the bit buffer are searched first
        z->in_sync = true;
        z->sync_have = 0;
        const uint32_t nb = z->held_bits >> 3;
        if (nb && z->carry_shadow.size() >= nb) search(z->carry_shadow.data() + (z->carry_shadow.size() - nb), nb);
        z->held_bits = 0;
    }
    const unsigned len = search(strm->next_in, strm->avail_in);
    strm->avail_in -= len;
    strm->next_in += len;
    strm->total_in += len;
    if (z->sync_have != 4) return Z_DATA_ERROR;
    // found: restart at a block boundary; no header yet -> raw, else the wrapper stays but its check is not computed
    const bool header_seen = z->header_seen;
    const uint32_t kind = header_seen ? z->kind : 0;
    if (!header_seen) z->wrap = 0;
    else z->wrap &= ~4;
    const uLong in = strm->total_in, out = strm->total_out;
    int wrap_unused, d64, wbits;
    zs_decode_wbits(z->wbits_arg, wrap_unused, d64, wbits);
    strm->total_in = strm->total_out = 0;
    strm->msg = Z_NULL;
    if (z->wrap) strm->adler = (uLong)(z->wrap & 1);
    const int saved_wrap = z->wrap;
    if (zs_reset_device(z, z->wbits_arg, d64, wbits) != Z_OK) return Z_MEM_ERROR;
    z->wrap = saved_wrap;
    z->header_seen = header_seen;
    z->kind = kind;
    strm->total_in = in;
    strm->total_out = out;
    {
        std::lock_guard<std::mutex> lk(g_zs.mu);
        ZsDeviceGuard dg;
        struct { uint32_t phase, bit0, last, d64; int32_t wbits_arg; uint32_t kind; } head6 = {zs::PH_BLOCK, 0u, 0u, (uint32_t)d64, z->wbits_arg, kind};
        bool ok = cudaMemcpyAsync(z->d_state, &head6, sizeof head6, cudaMemcpyHostToDevice, g_zs.st) == cudaSuccess;
        const uint32_t one = 1, chk = (uint32_t)(z->wrap & 1);
        ok = ok && zs_poke(z, offsetof(zs::State, noverify), one) && zs_poke(z, offsetof(zs::State, check), chk);
        if (!ok) {
            cudaGetLastError();
            return Z_MEM_ERROR;
        }
    }
    z->phase = zs::PH_BLOCK;
    return Z_OK;
}

// reference src/inflate.c:1444-1450
ZSB_API int inflateSyncPoint(z_streamp strm) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    return (int)h->zs->sync_point;
}

// reference src/inflate.c:1452-1492
ZSB_API int inflateCopy(z_streamp dest, z_streamp source) {
    using namespace zsb;
    ZsHandle *hs = zs_handle(source);
    if (!hs || dest == Z_NULL) return Z_STREAM_ERROR;
    ZsHandle *hd = (ZsHandle *)source->zalloc(source->opaque, 1, sizeof(ZsHandle));
    if (hd == nullptr) return Z_MEM_ERROR;
    ZStream *z = new (std::nothrow) ZStream(*hs->zs);
    bool ok = z != nullptr;
    if (ok) {
        z->d_state = nullptr; z->d_hist = nullptr; z->d_carry = nullptr;
        std::lock_guard<std::mutex> lk(g_zs.mu);
        ZsDeviceGuard dg;
        ok = cudaMalloc(&z->d_state, sizeof(zs::State)) == cudaSuccess && cudaMalloc(&z->d_hist, zs::HIST) == cudaSuccess &&
             cudaMalloc(&z->d_carry, ZS_CARRY_CAP) == cudaSuccess &&
             cudaMemcpyAsync(z->d_state, hs->zs->d_state, sizeof(zs::State), cudaMemcpyDeviceToDevice, g_zs.st) == cudaSuccess &&
             cudaMemcpyAsync(z->d_hist, hs->zs->d_hist, zs::HIST, cudaMemcpyDeviceToDevice, g_zs.st) == cudaSuccess &&
             cudaMemcpyAsync(z->d_carry, hs->zs->d_carry, ZS_CARRY_CAP, cudaMemcpyDeviceToDevice, g_zs.st) == cudaSuccess &&
             cudaStreamSynchronize(g_zs.st) == cudaSuccess;
        if (!ok) {
            cudaGetLastError();
            cudaFree(z->d_state); cudaFree(z->d_hist); cudaFree(z->d_carry);
        }
    }
    if (!ok) {
        delete z;
        source->zfree(source->opaque, hd);
        return Z_MEM_ERROR;
    }
    memcpy((void *)dest, (const void *)source, sizeof(z_stream));
    hd->strm = dest;
    hd->magic = ZS_MAGIC;
    hd->zs = z;
    dest->state = (struct internal_state *)hd;
    return Z_OK;
}

// reference src/inflate.c:231-248.  Bits are inserted in front of the input still to come; the carried bytes are
// rewritten so that they end with the reference's bit buffer followed by the new bits.  That needs the carried bytes
// to BE the bit buffer (a stream at a block boundary, at its start, or primed before): in the middle of a block
// header or symbol the unit is decoded again from its first bit here, and the call answers Z_STREAM_ERROR.
ZSB_API int inflatePrime(z_streamp strm, int bits, int value) {
    using namespace zsb;
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    if (bits == 0) return Z_OK;
    ZStream *z = h->zs;
    if (z->wrap && z->phase == zs::PH_HEAD && !z->header_seen) return Z_STREAM_ERROR;   // (a wrapper header is parsed bytewise here)
    const uint32_t held = z->held_bits;
    const bool exact = (uint64_t)z->carry_len * 8 - z->bit0 == held && z->carry_shadow.size() == z->carry_len;
    uint64_t hold = 0;
    uint32_t nbits = held;
    if (bits < 0) {
        if (!exact && held) return Z_STREAM_ERROR;
        nbits = 0;
    } else {
        if (bits > 16 || held + (uInt)bits > 32) return Z_STREAM_ERROR;
        if (!exact) return Z_STREAM_ERROR;
        for (uint32_t i = 0; i < z->carry_len; i++) hold |= (uint64_t)z->carry_shadow[i] << (8 * i);
        hold >>= z->bit0;
        hold |= (uint64_t)((unsigned)value & ((1u << bits) - 1u)) << held;
        nbits = held + (uint32_t)bits;
    }
    // the new carried bytes: nbits bits ending at a byte boundary
    const uint32_t nbytes = (nbits + 7) / 8, bit0 = nbytes * 8 - nbits;
    std::vector<uint8_t> nc(nbytes);
    const uint64_t v = hold << bit0;
    for (uint32_t i = 0; i < nbytes; i++) nc[i] = (uint8_t)(v >> (8 * i));
    std::lock_guard<std::mutex> lk(g_zs.mu);
    ZsDeviceGuard dg;
    bool ok = nbytes == 0 || cudaMemcpyAsync(z->d_carry, nc.data(), nbytes, cudaMemcpyHostToDevice, g_zs.st) == cudaSuccess;
    ok = ok && zs_poke(z, offsetof(zs::State, bit0), bit0);
    if (!ok) {
        cudaGetLastError();
        return Z_MEM_ERROR;
    }
    z->carry_shadow.swap(nc);
    z->carry_len = nbytes;
    z->bit0 = bit0;
    z->held_bits = nbits;
    return Z_OK;
}

// reference src/inflate.c:1494-1507: built without INFLATE_ALLOW_INVALID_DISTANCE_TOOFAR_ARRR, as the reference is
ZSB_API int inflateUndermine(z_streamp strm, int subvert) {
    (void)subvert;
    return zs_handle(strm) ? Z_DATA_ERROR : Z_STREAM_ERROR;
}

// reference src/inflate.c:1509-1519
ZSB_API int inflateValidate(z_streamp strm, int check) {
    using namespace zsb;
    ZsHandle *h = zs_handle(strm);
    if (!h) return Z_STREAM_ERROR;
    ZStream *z = h->zs;
    if (check && z->wrap) z->wrap |= 4;
    else z->wrap &= ~4;
    const uint32_t nv = (z->wrap & 4) ? 0u : 1u;
    std::lock_guard<std::mutex> lk(g_zs.mu);
    ZsDeviceGuard dg;
    if (!zs_poke(z, offsetof(zs::State, noverify), nv)) {
        cudaGetLastError();
        return Z_MEM_ERROR;
    }
    return Z_OK;
}

// reference src/inflate.c:1521-1530
ZSB_API long inflateMark(z_streamp strm) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return -(1L << 16);
    return h->zs->mark;
}

// reference src/inflate.c:1532-1537
ZSB_API unsigned long inflateCodesUsed(z_streamp strm) {
    ZsHandle *h = zs_handle(strm);
    if (!h) return (unsigned long)-1;
    return h->zs->codes_used;
}

}  // extern "C"
