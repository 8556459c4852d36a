/* zsb.h — C ABI of the B200 batched inflate engine (libzsb.so).
 *
 * Two surfaces, both plain C (pointers and sizes only, no CUDA/torch types in any signature):
 *
 *  (1) The zlib z_stream surface the reference's callers bind:
 *        inflateInit_      replaces reference src/inflate.c:226-229
 *        inflateInit2_     replaces reference src/inflate.c:186-224   (windowBits -16 = Deflate64)
 *        inflate           replaces reference src/inflate.c:597-1275
 *        inflateEnd        replaces reference src/inflate.c:1277-1287
 *        inflateReset/inflateReset2/inflateResetKeep   reference src/inflate.c:112-184
 *      and the auxiliary entry points of the inflate side:
 *        inflatePrime          src/inflate.c:231    inflateGetDictionary  :1289    inflateSetDictionary :1309
 *        inflateGetHeader      :1341                inflateSync           :1386    inflateSyncPoint     :1444
 *        inflateCopy           :1452                inflateUndermine      :1494    inflateValidate      :1509
 *        inflateMark           :1521                inflateCodesUsed      :1532
 *      inflate() honours Z_BLOCK and Z_TREES (early returns + the 128 / 256 bits of data_type).
 *      Deviations, all of them in states the reference's callers never enter: (a) inflatePrime answers
 *      Z_STREAM_ERROR when the stream stopped inside a block header or a symbol (the unit is decoded again
 *      from its first bit here, so bits cannot be spliced into it) or before the header of a zlib / gzip
 *      stream; at a block boundary, at the start of a raw stream and after an earlier inflatePrime it works
 *      as in the reference; (b) the fields of a gz_header registered with inflateGetHeader are filled when the
 *      header is complete (done = 1), not field by field while it arrives; (c) inflateCodesUsed reports the
 *      code-length code's table only once the block header is complete.
 *      They are declared by the system <zlib.h> (the same header the reference compiles against) and
 *      are exactly what src/wasm/inflate_stream_wasm.c:26,33,41,49-53, src/wasm/inflate9_stream_wasm.c:
 *      19,27-30, src/wasm/wasm_stream_common.c:58 and test/payload_decompress.c:124,156,192 call.
 *      libzsb.so exports them with zlib's signatures; every call runs on the GPU (there is no CPU
 *      fallback: without a usable CUDA device they return Z_STREAM_ERROR / Z_MEM_ERROR).
 *
 *  (2) A thin batch call (new; the throughput surface).  Per member it is defined to be equivalent to
 *      the reference sequence
 *          inflateInit2(&s, window_bits); s.next_in/avail_in = in/in_len; s.next_out/avail_out = out/
 *          out_cap; ret = inflate(&s, Z_FINISH); inflateEnd(&s);
 *      (test/payload_decompress.c:124-192 with a buffer large enough for one call) and reports what
 *      that sequence leaves on the z_stream: ret, msg, total_in, total_out, adler, data_type.
 */
#ifndef ZSB_H
#define ZSB_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ZSB_VERSION "0.1.0"
#if defined(__GNUC__)
#define ZSB_API __attribute__((visibility("default")))
#else
#define ZSB_API
#endif

/* One independent stream ("member").  in/out may be host or device pointers (all members of one call
 * must be of the same kind; device pointers must belong to the context's device). */
typedef struct zsb_member {
    const void *in;      /* compressed bytes */
    size_t in_len;       /* < 2^32 (zlib's avail_in is a uInt) */
    void *out;           /* output buffer; bytes [total_out, out_cap) are unspecified after the call */
    size_t out_cap;      /* < 2^32 */
    int32_t window_bits; /* as inflateInit2: -8..-15 raw, -16 raw Deflate64, 8..15 zlib, +16 gzip, +32 auto */
    int32_t flags;       /* 0, or ZSB_M_* (extensions for container formats; the reference has no equivalent) */
} zsb_member;

/* zsb_member.flags */
#define ZSB_M_CRC32  1   /* raw member: report the CRC-32 of the output bytes in zsb_result.check (ZIP entries) */
#define ZSB_M_STORED 2   /* the member is not deflate data: copy in -> out (ZIP method 0); ret = Z_STREAM_END, or
                            Z_BUF_ERROR with out_cap bytes copied when out_cap < in_len; window_bits is ignored */

/* What the reference leaves on the z_stream after the one-shot sequence above (src/inflate.c:1253-1274). */
typedef struct zsb_result {
    int32_t ret;        /* Z_STREAM_END 1, Z_NEED_DICT 2, Z_STREAM_ERROR -2, Z_DATA_ERROR -3, Z_BUF_ERROR -5 */
    int32_t msg_id;     /* 0 = strm->msg == NULL, else index for zsb_message() */
    uint64_t total_in;
    uint64_t total_out;
    uint32_t check;     /* strm->adler: Adler-32 (zlib) / CRC-32 (gzip) of the output, 0 for raw */
    int32_t data_type;  /* strm->data_type */
} zsb_result;

typedef struct zsb_ctx zsb_ctx;

/* Create a context over `ndev` CUDA devices (device ordinals in `devices`; NULL/0 = current device).
 * Returns NULL when no CUDA device is usable (zsb_create_error() says why). */
ZSB_API zsb_ctx *zsb_create(const int *devices, int ndev);
ZSB_API void zsb_destroy(zsb_ctx *ctx);
ZSB_API const char *zsb_create_error(void);
ZSB_API const char *zsb_last_error(const zsb_ctx *ctx);
ZSB_API int zsb_num_devices(const zsb_ctx *ctx);

/* Decode n members.  Host pointers: the batch is sharded over the context's devices by
 * zsb_plan_shards(), staged through device arenas (host<->device copies included) and the call
 * returns when every result and output byte is in host memory.  Device pointers: decoded in place on
 * the first device, results copied to the host array `r`.  Returns 0, or a negative zsb error
 * (ZSB_E_*) for failures of the call itself (per-member outcomes are in r[i].ret). */
ZSB_API int zsb_inflate_batch(zsb_ctx *ctx, size_t n, const zsb_member *m, zsb_result *r);

/* Device-resident form (the kernel-only path): `d_members` and `d_results` are device arrays on device
 * `dev_index` of the context, all in/out pointers are device pointers on it.  `h_members` is a host
 * copy of the same descriptors (only in_len / out_cap / window_bits are read: the host sizes the
 * decoder's scratch from them without a device round trip).  `h_order` (optional, host, n x uint32)
 * is the order in which members are handed to the decoder (largest first balances best; NULL =
 * identity).  `flags` = ZSB_RAW_ONLY when every member has window_bits < 0, else ZSB_HAS_WRAPPED.
 * Work is enqueued on `cuda_stream` (a cudaStream_t, NULL = default stream); the call does not wait
 * for the GPU.  One call at a time per device of a context. */
ZSB_API int zsb_inflate_batch_device(zsb_ctx *ctx, int dev_index, size_t n, const zsb_member *d_members,
                             const zsb_member *h_members, zsb_result *d_results, const uint32_t *h_order,
                             unsigned flags, void *cuda_stream);

#define ZSB_HAS_WRAPPED 1u /* some member has window_bits >= 0 (header/trailer + checksum kernels needed) */
#define ZSB_HAS_LARGE   2u /* reserved */
#define ZSB_RAW_ONLY    4u /* every member has window_bits < 0: skip checksum/trailer kernels */

/* Tuning knobs: "path" 0 auto / 1 lane path (one member per lane, k_lane + k_lz) / 2 group path
 * (k_inflate<G>); "profile" 1 = time the decode kernels (zsb_profile_read); "lanes_per_member" G of the group path (1,2,4,8,16,32); "seq_arena_bytes" most
 * sequence scratch the lane path may hold at once (larger batches are decoded in slices; 0, the default = half
 * of the device's free memory, at least 32 GiB); "split_big" 1 (default) members far larger than the rest of
 * their batch are decoded a warp each (zsb_plan_big), 0 never; "chunks" pipeline depth of host-buffer batches
 * (0 = choose); "segment" 1 (default): members of "seg_min" (131072) compressed bytes or more, unless they do not
 * shrink, are decoded in parallel from block starts found near every multiple of "seg_bytes" (0 = chosen per batch, 8-64 KiB) of their input
 * and verified to fit together (device-resident batches), 0 never. */
ZSB_API int zsb_set_option(zsb_ctx *ctx, const char *name, long value);
ZSB_API long zsb_get_option(const zsb_ctx *ctx, const char *name);

/* With option "profile" = 1 every k_lane / k_lz launch is bracketed by CUDA events on the launching
 * stream.  zsb_profile_read() waits for them, returns the number of launch pairs recorded on device
 * `dev_index` since the last read and their summed durations in milliseconds. */
ZSB_API long zsb_profile_read(zsb_ctx *ctx, int dev_index, double *lane_ms, double *lz_ms);
/* ... and of the k_check launches (Adler-32 / CRC-32 + trailer checks). */
ZSB_API long zsb_profile_read_check(zsb_ctx *ctx, int dev_index, double *check_ms);
/* ... and of the segmented parts (large members decoded in segments, option "segment"): search, decode, chain, resolve. */
ZSB_API long zsb_profile_read_seg(zsb_ctx *ctx, int dev_index, double *seg_ms);

/* Number of kernels the last zsb_inflate_batch*() call launched (for accounting). */
ZSB_API long zsb_last_launches(const zsb_ctx *ctx);

/* Length-aware sharding (pure host code, no GPU needed): assigns each of n members with cost
 * estimate cost[i] to one of nshards shards, longest-processing-time first.  shard_of[i] receives the
 * shard index; load[s] (optional) the summed cost per shard.  Deterministic. */
ZSB_API int zsb_plan_shards(size_t n, const uint64_t *cost, int nshards, int32_t *shard_of, uint64_t *load);

/* Length-aware split inside one GPU (pure host code): of n members given LARGEST FIRST, how many are so much
 * larger than the rest that they are decoded a warp each (k_big) beside the others (one lane each, k_lane).
 * 0 for a uniform batch.  The cost model is measured on B200; sm_count = SMs of the device. */
ZSB_API size_t zsb_plan_big(size_t n, const uint64_t *in_len, const uint64_t *out_cap, int sm_count);

/* Pinned host memory for fast host<->device staging. */
ZSB_API void *zsb_host_alloc(size_t bytes);
ZSB_API void zsb_host_free(void *p);

/* strm->msg strings (reference src/inflate.c:652...1223, src/inffast.c:172,268,283); NULL for 0. */
#define ZSB_NUM_MESSAGES 19
ZSB_API const char *zsb_message(int msg_id);

/* ---- ZIP container walker: archive bytes -> batch descriptors -> verified entries ----
 * The reference decodes the Deflate64 (method 9) and deflate (method 8) entries zip.js hands it one at a
 * time; its generator reads the same header fields (src/generator/create_deflate64.cpp:51-93: local header
 * signature PK\3\4, method @+8, sizes @+18/+22, name/extra lengths @+26/+28).  Here a whole archive becomes
 * one batch.  The walker is host code over the archive's directory only; entry data is touched by the GPU. */
typedef struct zsb_zip_entry {
    uint64_t header_off;   /* local file header */
    uint64_t data_off;     /* first compressed byte (header_off + 30 + name + extra of the LOCAL header) */
    uint64_t comp_size;
    uint64_t uncomp_size;
    uint32_t crc32;        /* from the central directory */
    uint16_t method;       /* 0 stored, 8 deflate, 9 Deflate64; anything else is reported, not decoded */
    uint16_t gp_flags;     /* general purpose bits (bit 0 encrypted, bit 3 sizes also in a data descriptor) */
    uint32_t name_off;     /* file name bytes inside the archive (central directory copy) */
    uint16_t name_len;
    uint16_t pad;
} zsb_zip_entry;

/* Parses the end-of-central-directory record (ZIP64 included) and the central directory of the archive at
 * host address `zip`.  Writes at most max_entries entries, *count receives the number of entries in the
 * archive.  Returns 0, or ZSB_E_ZIP when the directory is malformed or points outside the archive. */
ZSB_API int zsb_zip_scan(const void *zip, size_t len, zsb_zip_entry *entries, size_t max_entries, size_t *count);

/* Batch descriptors for n entries: in = in_base + data_off, out = out_base + out_off[i], out_cap =
 * uncomp_size, window_bits -15 / -16 by method, flags ZSB_M_CRC32 (| ZSB_M_STORED for method 0).
 * in_base / out_base may be host or device addresses of the archive and of the output area.  Entries that
 * cannot be decoded (encrypted, unknown method, sizes >= 4 GiB) get window_bits 1000, i.e. Z_STREAM_ERROR. */
ZSB_API int zsb_zip_members(const void *in_base, const zsb_zip_entry *entries, size_t n, void *out_base,
                            const uint64_t *out_off, zsb_member *members);

/* zsb_zip_members + zsb_inflate_batch + the directory checks, per entry: ret = Z_STREAM_END when the entry
 * decoded to exactly uncomp_size bytes whose CRC-32 (computed on the GPU) equals the directory's; Z_DATA_ERROR
 * with msg "incorrect data check" / "incorrect length check" (the reference's gzip trailer messages,
 * src/inflate.c:1209,1223) when CRC or size differ; otherwise the decoder's own ret / msg. */
ZSB_API int zsb_zip_inflate(zsb_ctx *ctx, const void *zip, const zsb_zip_entry *entries, size_t n, void *out_base,
                            const uint64_t *out_off, zsb_result *results);

/* ---- gzip files of several members ----
 * The reference's inflate() stops after one member (src/inflate.c:1231-1235: mode DONE, trailing bytes stay
 * unread, total_in = member length); real .gz files are often concatenations.  Every file i is a buffer of
 * in_len[i] bytes holding one or more gzip members (window_bits 31; 47 also accepts zlib members) back to
 * back; its members are decoded one after the other into out[i] (capacity out_cap[i]).  Each round decodes
 * the next member of every unfinished file as one batch.  Per file, r[i] reports the LAST member decoded:
 * ret = Z_STREAM_END when the whole buffer was consumed by complete members, else that member's ret / msg;
 * total_in / total_out are summed over the file's members; check is the last member's.  members[i]
 * (optional) receives the number of complete members.  Host or device pointers, as zsb_inflate_batch. */
ZSB_API int zsb_gunzip_files(zsb_ctx *ctx, size_t n, const void *const *in, const size_t *in_len, void *const *out,
                             const size_t *out_cap, int window_bits, zsb_result *r, uint32_t *members);

/* ---- gzip header fields ----
 * What inflateGetHeader() surfaces while inflate() walks a gzip header (reference src/inflate.c:678-816 fills
 * gz_header; :1341-1356 registers it): a host-side parse of the member that starts at `in`, the data itself is
 * not touched.  Offsets are relative to `in`; a field that is absent has length 0 and offset 0; name and comment
 * lengths do not count their terminating NUL.  Returns 0 with h filled when the header is complete (the deflate
 * data starts at in + h->header_len), Z_BUF_ERROR (-5) when the buffer ends inside the header, Z_DATA_ERROR (-3)
 * with h->msg_id = the zsb_message() index inflate() would report there ("incorrect header check", "unknown
 * compression method", "unknown header flags set", "header crc mismatch"), or ZSB_E_ARG. */
typedef struct zsb_gz_header {
    uint32_t text;        /* FTEXT: the compressor believed the data is text */
    uint32_t time;        /* MTIME */
    uint32_t xflags;      /* XFL */
    uint32_t os;          /* OS */
    uint32_t extra_off, extra_len;
    uint32_t name_off, name_len;
    uint32_t comment_off, comment_len;
    uint32_t hcrc;        /* 1 = a header CRC is present (and was verified) */
    uint32_t header_len;
    int32_t msg_id;       /* 0 unless the return value is Z_DATA_ERROR */
    uint32_t reserved;
} zsb_gz_header;
ZSB_API int zsb_gzip_header(const void *in, size_t len, zsb_gz_header *h);

/* ---- many z_streams per call ----
 * rets[i] = inflate(strms[i], flush) for n streams initialised with inflateInit2_() of this library, with
 * every round of GPU work done for all of them by one set of launches (one CTA per stream) and one pair of
 * PCIe copies: the shape of the reference's real caller, many TransformStreams feeding <= 32 KiB in / 64 KiB
 * out per call (src/wasm/api/zlib-streams.js:64-138).  strms is an array of z_streamp (declared void * here so
 * that this header does not need <zlib.h>); each z_stream is updated exactly as inflate() would.  Returns 0, or
 * ZSB_E_ARG. */
ZSB_API int zsb_inflate_streams(size_t n, void **strms, int flush, int *rets);

/* ---- batched deflate (compression) on the GPU ----
 * The compress half of the reference's product: what src/wasm/deflate_stream_wasm.c:22-69 obtains from
 * deflateInit / deflateInit2(level, Z_DEFLATED, -15 | 15 | 31, 8, Z_DEFAULT_STRATEGY) + deflate(Z_FINISH) +
 * deflateEnd per stream, for n independent members in one call.  window_bits: -15 raw deflate, 15 zlib, 31 gzip
 * (the three the reference's glue uses; the window is always 32 KiB); level: 0 = stored blocks, 1..9 or -1 = LZ77 +
 * Huffman coding (one parameter set; the level only goes into the zlib / gzip header fields).  The output is a
 * valid stream of the requested format — it decodes to the input with the reference's inflate() — but not the
 * byte string zlib would produce (compressed streams are not unique).  ret: Z_STREAM_END (1), Z_BUF_ERROR (-5)
 * when out_cap is too small (zsb_deflate_bound() is always enough), Z_STREAM_ERROR (-2) for parameters outside
 * the above.  check = Adler-32 (zlib) / CRC-32 (gzip) of the input.  Host or device pointers, as
 * zsb_inflate_batch; runs on the first device of the context. */
typedef struct zsb_deflate_member {
    const void *in;
    size_t in_len;       /* < 2^32 */
    void *out;
    size_t out_cap;
    int32_t window_bits;
    int32_t level;
} zsb_deflate_member;
typedef struct zsb_deflate_result {
    int32_t ret;
    int32_t blocks;      /* 64 KiB input blocks the member was cut into (each compressed by one warp) */
    uint64_t total_in;
    uint64_t total_out;
    uint32_t check;
    uint32_t reserved;
} zsb_deflate_result;
ZSB_API int zsb_deflate_batch(zsb_ctx *ctx, size_t n, const zsb_deflate_member *members, zsb_deflate_result *results);
ZSB_API size_t zsb_deflate_bound(size_t in_len, int window_bits);

/* ---- batched deflate with no allocation ----
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The same kernels and the same
 * output bytes as zsb_deflate_batch(), but every piece of device scratch (descriptors, block table, token and block
 * buffers, results) lives in ONE file-scope device array of zsb_deflate_static_limits() bytes (about 8.4 MB) that
 * every GPU which loads the library holds until exit; the call makes no cudaMalloc / cudaFree / allocating library
 * call and needs no zsb_ctx.  members[i].in / .out must be DEVICE pointers on `device` (there is nowhere to stage
 * host buffers); members and results themselves are host arrays.  Limits: at most max_members (64) members and
 * max_blocks (64) blocks of 64 KiB over the whole call (a member of L bytes takes max(1, ceil(L / 65536)) blocks);
 * beyond them the call returns ZSB_E_LIMIT and says which limit in err.  The kernels run on `stream` (a
 * cudaStream_t; NULL = the default stream) and the call synchronises that stream before it returns.  All calls
 * share the array, which keeps the last call's contents (the call clears what must start at zero); calls in one
 * process are serialised by a mutex — work that overlapped on the GPU with another user of the array would race.
 * err (optional, err_cap bytes) receives a message when the return value is not 0: ZSB_E_ARG, ZSB_E_LIMIT, ZSB_E_CUDA. */
ZSB_API int zsb_deflate_batch_static(int device, size_t n, const zsb_deflate_member *members, zsb_deflate_result *results,
                                     void *stream, char *err, size_t err_cap);
/* bytes of the static array; the two limits through the pointers (either may be NULL) */
ZSB_API size_t zsb_deflate_static_limits(size_t *max_members, size_t *max_blocks);

#define ZSB_E_LIMIT (-104)
#define ZSB_E_ZIP   (-103)
#define ZSB_E_ARG   (-100)
#define ZSB_E_CUDA  (-101)
#define ZSB_E_NOMEM (-102)

#ifdef __cplusplus
}
#endif
#endif
