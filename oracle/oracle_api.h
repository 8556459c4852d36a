/* oracle/oracle_api.h — common C surface of the two CPU checkers.  TEST INFRASTRUCTURE ONLY.
 *
 *   ref_*  : the UNMODIFIED reference (src/inflate.c, src/inffast.c, src/inftrees.c compiled in
 *            place from /root/reference by oracle/Makefile into oracle/_ref/libzsb_ref.so).
 *   orc_*  : this repo's plain-C restatement of the same algorithm (oracle/orc_inflate.c).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load these libraries.  The product (libzsb.so) never links or calls them.
 */
#ifndef ZSB_ORACLE_API_H
#define ZSB_ORACLE_API_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Observable per-stream result, i.e. what a caller of the reference can see on the z_stream
 * after inflate() returns (reference src/inflate.c:1253-1274). */
typedef struct orc_result {
    int32_t ret;        /* return code of the last inflate() call */
    int32_t msg_id;     /* index into orc_messages[] (0 = strm->msg == NULL) */
    uint64_t total_in;  /* strm->total_in */
    uint64_t total_out; /* strm->total_out */
    uint32_t adler;     /* strm->adler (0 was stored before init so raw streams report 0) */
    int32_t data_type;  /* strm->data_type */
} orc_result;

/* The fixed set of strm->msg strings inflate() can set; index = msg_id.
 * (reference src/inflate.c:652,657,666,687,803,865,874,916,940,964,987,1001,1016,1026,1080,
 *  1118,1150,1209,1223 and src/inffast.c:172,268,283) */
#define ORC_NUM_MESSAGES 19
extern const char *const orc_messages[ORC_NUM_MESSAGES];

#define ORC_DECLARE(P)                                                                          \
    /* inflateInit2(wbits); inflate(Z_FINISH) once; inflateEnd.  Returns r->ret. */             \
    int P##_oneshot(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits,  \
                    orc_result *r);                                                             \
    /* Streaming: open/step/close.  step = one inflate(flush) call with the given buffers. */   \
    void *P##_open(int wbits, int *init_ret);                                                   \
    int P##_step(void *h, const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap,       \
                 int flush, size_t *consumed, size_t *produced, orc_result *r);                 \
    void P##_close(void *h);                                                                    \
    /* CPU baseline: decode n members, one member per thread at a time (static round robin),  \
     * each = Init2 + inflate(Z_FINISH) into a private buffer + End; returns best-of-reps      \
     * seconds; *sum_out = total decompressed bytes of one pass; *bad = members whose ret != 1 */ \
    double P##_bench(size_t n, const uint8_t *const *in, const size_t *in_len,                  \
                     const size_t *out_cap, const int *wbits, int threads, int reps,            \
                     uint64_t *sum_out, uint64_t *bad);

ORC_DECLARE(ref)
ORC_DECLARE(orc)

#ifdef __cplusplus
}
#endif
#endif
