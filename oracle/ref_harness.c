/* oracle/ref_harness.c — drives the UNMODIFIED reference inflate() through its public zlib API
 * (inflateInit2_/inflate/inflateEnd: reference src/inflate.c:186,597,1277), the way the
 * reference's own harness does (test/payload_decompress.c:124-192) and the way its WASM glue does
 * (src/wasm/wasm_stream_common.c:37-63).  TEST INFRASTRUCTURE ONLY; built into oracle/_ref/. */
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <zlib.h>
#include "oracle_api.h"

static int msg_to_id(const char *m) {
    if (!m) return 0;
    for (int i = 1; i < ORC_NUM_MESSAGES; i++)
        if (strcmp(m, orc_messages[i]) == 0) return i;
    return -1;
}

static void fill_result(z_stream *s, int ret, orc_result *r) {
    r->ret = ret;
    r->msg_id = msg_to_id(s->msg);
    r->total_in = s->total_in;
    r->total_out = s->total_out;
    r->adler = (uint32_t)s->adler;
    r->data_type = s->data_type;
}

int ref_oneshot(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits,
                orc_result *r) {
    z_stream s;
    static uint8_t dummy;
    memset(&s, 0, sizeof s);
    int ret = inflateInit2_(&s, wbits, ZLIB_VERSION, (int)sizeof(z_stream));
    if (ret != Z_OK) {
        memset(r, 0, sizeof *r);
        r->ret = ret;
        return ret;
    }
    s.next_in = (Bytef *)in;
    s.avail_in = (uInt)in_len;
    s.next_out = out ? out : &dummy; /* next_out == NULL is Z_STREAM_ERROR (inflate.c:617) */
    s.avail_out = (uInt)out_cap;
    ret = inflate(&s, Z_FINISH);
    fill_result(&s, ret, r);
    inflateEnd(&s);
    return ret;
}

/* inflateCodesUsed() (reference src/inflate.c:1532) after a one-shot decode: the number of table entries the
 * reference's inflate_table() allocated for the last block's codes.  Test hook for the table-space limits
 * (src/inftrees.c:211-228,295-297). */
long ref_codes_used(const uint8_t *in, size_t in_len, int wbits, orc_result *r) {
    z_stream s;
    static uint8_t sink[1024];
    memset(&s, 0, sizeof s);
    int ret = inflateInit2_(&s, wbits, ZLIB_VERSION, (int)sizeof(z_stream));
    if (ret != Z_OK) return -1;
    s.next_in = (Bytef *)in;
    s.avail_in = (uInt)in_len;
    s.next_out = sink;
    s.avail_out = sizeof sink;
    ret = inflate(&s, Z_FINISH);
    if (r) fill_result(&s, ret, r);
    long used = (long)inflateCodesUsed(&s);
    inflateEnd(&s);
    return used;
}

void *ref_open(int wbits, int *init_ret) {
    z_stream *s = (z_stream *)calloc(1, sizeof(z_stream));
    int ret = inflateInit2_(s, wbits, ZLIB_VERSION, (int)sizeof(z_stream));
    if (init_ret) *init_ret = ret;
    if (ret != Z_OK) {
        free(s);
        return 0;
    }
    return s;
}

int ref_step(void *h, const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int flush,
             size_t *consumed, size_t *produced, orc_result *r) {
    z_stream *s = (z_stream *)h;
    static uint8_t dummy;
    s->next_in = (Bytef *)in;
    s->avail_in = (uInt)in_len;
    s->next_out = out ? out : &dummy;
    s->avail_out = (uInt)out_cap;
    int ret = inflate(s, flush);
    if (consumed) *consumed = in_len - s->avail_in;
    if (produced) *produced = out_cap - s->avail_out;
    fill_result(s, ret, r);
    return ret;
}

void ref_close(void *h) {
    if (!h) return;
    inflateEnd((z_stream *)h);
    free(h);
}

/* ---- CPU baseline: one stream per core ---- */
typedef struct {
    size_t n;
    const uint8_t *const *in;
    const size_t *in_len;
    const size_t *out_cap;
    const int *wbits;
    int tid, threads;
    uint64_t sum_out, bad;
    uint8_t *buf;
    size_t buf_cap;
} bench_arg;

static void *bench_worker(void *p) {
    bench_arg *a = (bench_arg *)p;
    a->sum_out = a->bad = 0;
    for (size_t i = (size_t)a->tid; i < a->n; i += (size_t)a->threads) {
        orc_result r;
        size_t cap = a->out_cap[i] <= a->buf_cap ? a->out_cap[i] : a->buf_cap;
        ref_oneshot(a->in[i], a->in_len[i], a->buf, cap, a->wbits[i], &r);
        a->sum_out += r.total_out;
        if (r.ret != Z_STREAM_END) a->bad++;
    }
    return 0;
}

double ref_bench(size_t n, const uint8_t *const *in, const size_t *in_len, const size_t *out_cap,
                 const int *wbits, int threads, int reps, uint64_t *sum_out, uint64_t *bad) {
    /* fixedtables() builds its static tables lazily and is not thread-safe
     * (reference src/inflate.c:256-263): warm both modes on this thread first. */
    static const uint8_t fixed_empty[2] = {0x03, 0x00};
    uint8_t tmp[4];
    orc_result r;
    ref_oneshot(fixed_empty, 2, tmp, sizeof tmp, -15, &r);
    ref_oneshot(fixed_empty, 2, tmp, sizeof tmp, -16, &r);

    size_t maxcap = 1;
    for (size_t i = 0; i < n; i++)
        if (out_cap[i] > maxcap) maxcap = out_cap[i];
    if (threads < 1) threads = 1;
    bench_arg *args = (bench_arg *)calloc((size_t)threads, sizeof(bench_arg));
    pthread_t *th = (pthread_t *)calloc((size_t)threads, sizeof(pthread_t));
    for (int t = 0; t < threads; t++) {
        args[t].n = n; args[t].in = in; args[t].in_len = in_len; args[t].out_cap = out_cap;
        args[t].wbits = wbits; args[t].tid = t; args[t].threads = threads;
        args[t].buf = (uint8_t *)malloc(maxcap); args[t].buf_cap = maxcap;
    }
    double best = 1e30;
    for (int rep = 0; rep < reps; rep++) {
        struct timespec t0, t1;
        clock_gettime(CLOCK_MONOTONIC, &t0);
        for (int t = 0; t < threads; t++) pthread_create(&th[t], 0, bench_worker, &args[t]);
        for (int t = 0; t < threads; t++) pthread_join(th[t], 0);
        clock_gettime(CLOCK_MONOTONIC, &t1);
        double dt = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
        if (dt < best) best = dt;
    }
    uint64_t so = 0, b = 0;
    for (int t = 0; t < threads; t++) { so += args[t].sum_out; b += args[t].bad; free(args[t].buf); }
    if (sum_out) *sum_out = so;
    if (bad) *bad = b;
    free(args); free(th);
    return best;
}
