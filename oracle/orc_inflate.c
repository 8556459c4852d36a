/* oracle/orc_inflate.c — plain-C restatement of the reference's inflate / Deflate64 decoder.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the CPU checker ("oracle") the CUDA path is compared
 * against.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may build, load
 * or call it; the product library (libzsb.so) never links it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement against
 *   (1) the reference's own golden fixtures test/ref-data/ *.deflate64 (sha256 table),
 *   (2) the unmodified reference compiled in place (oracle/_ref/libzsb_ref.so) on the fixtures,
 *       on zlib/gzip/raw/Deflate64 streams, on corrupted streams and on random chunkings,
 *       comparing ret, msg, total_in, total_out, adler, data_type and every output byte.
 *
 * What is restated (reference file:line):
 *   windowBits decoding, reset            src/inflate.c:112-224
 *   zlib / gzip header states              src/inflate.c:629-830
 *   block dispatch                         src/inflate.c:831-869
 *   stored blocks                          src/inflate.c:870-904
 *   dynamic header                         src/inflate.c:905-1033
 *   symbol loop (slow path; the fast path  src/inflate.c:1034-1193, src/inffast.c:50-304
 *     is externally invisible)
 *   trailer                                src/inflate.c:1194-1235
 *   exit bookkeeping                       src/inflate.c:1253-1274
 *   code-set accept/reject + table space   src/inftrees.c:121-153, 211-228, 275-297
 *   length/distance base+extra tables      src/inftrees.c:59-87
 *   fixed code                             src/inflate.c:260-298
 *   sliding window                         src/inflate.c:375-419
 *   CRC-32 / Adler-32: madler/zlib @5a82f71 (crc32.c, adler32.c; submodule absent from the
 *     reference tree) — restated from their published definitions (CRC-32/ISO-HDLC, RFC 1950).
 *
 * The restatement keeps its own structure: a canonical (count/first/sorted-symbol) Huffman
 * decoder instead of the reference's two-level lookup tables, explicit helper functions instead of
 * the macro state machine, and a linear "history ring".  What it preserves exactly is the
 * observable behaviour: which bytes are pulled from the input before each decision, the order of
 * the checks, and the values left on the stream when inflate() returns.
 */
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "oracle_api.h"

enum { Z_OK_ = 0, Z_STREAM_END_ = 1, Z_NEED_DICT_ = 2, Z_STREAM_ERROR_ = -2, Z_DATA_ERROR_ = -3,
       Z_MEM_ERROR_ = -4, Z_BUF_ERROR_ = -5 };
enum { FL_NO = 0, FL_FINISH = 4, FL_BLOCK = 5, FL_TREES = 6 };

/* phases; the order matters for the exit rules (inflate.c:1255-1256 compares modes) */
enum {
    P_HEAD, P_GZ_FLAGS, P_GZ_TIME, P_GZ_OS, P_GZ_EXLEN, P_GZ_EXTRA, P_GZ_NAME, P_GZ_COMMENT,
    P_GZ_HCRC, P_DICTID, P_DICT,
    P_BLOCK, P_BLOCKDO, P_STORED, P_COPY_ARMED, P_COPY, P_DYN, P_DYN_CLENS, P_DYN_LENS,
    P_SYM_ARMED, P_SYM, P_LENEXT, P_DIST, P_DISTEXT, P_MATCH, P_LIT,
    P_CHECK, P_LENGTH, P_DONE, P_BAD
};

/* canonical Huffman code: count per length + symbols sorted by (length, symbol) */
typedef struct {
    uint16_t count[16];
    uint16_t sym[288];
    int kind; /* 0 = normal (complete), 1 = no codes at all, 2 = a single 1-bit code */
} huff_t;

typedef struct orc_stream {
    const uint8_t *in;  size_t in_avail;  uint64_t total_in;
    uint8_t *out;       size_t out_avail; uint64_t total_out;
    int msg; uint32_t adler; int data_type;

    int wrap;           /* bit0 zlib, bit1 gzip, bit2 verify (inflate.c:163-170) */
    int d64;            /* Deflate64 (inflate.c:160) */
    unsigned wbits;
    uint64_t acc; unsigned nacc;   /* bit accumulator (inflate.c "hold"/"bits") */
    int phase, last;
    int gzflags;        /* -1 before the header, 0 zlib, else gzip FLG word (inflate.c:124,671,680) */
    uint32_t check;     /* running header crc / data check */
    uint64_t total;     /* output counted for ISIZE */
    unsigned length, offset, extra;
    unsigned nlen, ndist, ncode, have;
    uint16_t lens[320];
    huff_t clen, lit, dist;
    /* history of earlier calls' output (inflate.c:375-419) */
    uint8_t *win; unsigned wsize, whave, wnext;
} orc_stream;

/* ---- checksums (published definitions) ---- */
static uint32_t crc_tab[256];
static pthread_once_t crc_once = PTHREAD_ONCE_INIT;
static void crc_init(void) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xEDB88320u & (0u - (c & 1)));
        crc_tab[i] = c;
    }
}
static uint32_t crc32_upd(uint32_t crc, const uint8_t *p, size_t n) {
    pthread_once(&crc_once, crc_init);
    crc = ~crc;
    while (n--) crc = crc_tab[(crc ^ *p++) & 0xff] ^ (crc >> 8);
    return ~crc;
}
static uint32_t adler32_upd(uint32_t ad, const uint8_t *p, size_t n) {
    uint32_t a = ad & 0xffff, b = ad >> 16;
    while (n) {
        size_t k = n < 5552 ? n : 5552;
        n -= k;
        while (k--) { a += *p++; b += a; }
        a %= 65521u; b %= 65521u;
    }
    return (b << 16) | a;
}
static uint32_t check_upd(const orc_stream *s, uint32_t c, const uint8_t *p, size_t n) {
    return s->gzflags ? crc32_upd(c, p, n) : adler32_upd(c, p, n); /* inflate.c:424-429 */
}

/* ---- bit input: one byte at a time, exactly as needed (inflate.c:481-513) ---- */
static int need(orc_stream *s, unsigned n) {
    while (s->nacc < n) {
        if (s->in_avail == 0) return 0;
        s->acc |= (uint64_t)(*s->in++) << s->nacc;
        s->in_avail--;
        s->nacc += 8;
    }
    return 1;
}
static unsigned peek(const orc_stream *s, unsigned n) {
    return (unsigned)(s->acc & ((1ull << n) - 1));
}
static void drop(orc_stream *s, unsigned n) { s->acc >>= n; s->nacc -= n; }
static void dropall(orc_stream *s) { s->acc = 0; s->nacc = 0; }

/* ---- code sets ---- */
enum { K_CODES, K_LENS, K_DISTS };

/* Table space the reference's builder would allocate (inftrees.c:222-228, 275-297): the root
 * table plus one sub-table per distinct root-bit prefix of the codes longer than root, each
 * sub-table sized from the remaining length counts. */
static unsigned table_space(const uint16_t *count, unsigned root, unsigned minl, unsigned maxl) {
    unsigned rem[16];
    for (unsigned l = 0; l < 16; l++) rem[l] = count[l];
    unsigned used = 1u << root;
    unsigned code = 0;               /* canonical code, MSB first */
    unsigned low = ~0u;
    for (unsigned len = minl; len <= maxl; len++) {
        for (unsigned k = count[len]; k; k--) {
            if (len > root) {
                unsigned prefix = code >> (len - root);
                if (prefix != low) {
                    unsigned curr = len - root;
                    int left = 1 << curr;
                    while (curr + root < maxl) {
                        left -= (int)rem[curr + root];
                        if (left <= 0) break;
                        curr++;
                        left <<= 1;
                    }
                    used += 1u << curr;
                    low = prefix;
                }
            }
            rem[len]--;
            code++;
        }
        code <<= 1;
    }
    return used;
}

/* Accept/reject a set of code lengths and prepare the canonical decoder.
 * Returns 0 if the reference's inflate_table() would return 0, nonzero otherwise. */
static int huff_build(huff_t *h, int kind, const uint16_t *lens, unsigned n, int d64) {
    unsigned offs[16];
    memset(h->count, 0, sizeof h->count);
    for (unsigned i = 0; i < n; i++) h->count[lens[i]]++;
    unsigned maxl = 15, minl = 1;
    while (maxl >= 1 && h->count[maxl] == 0) maxl--;
    if (maxl == 0) {                         /* inftrees.c:131-140 */
        if (d64) return -1;
        h->kind = 1;
        return 0;
    }
    while (minl < maxl && h->count[minl] == 0) minl++;
    int left = 1;                            /* inftrees.c:146-153 */
    for (unsigned l = 1; l <= 15; l++) {
        left = (left << 1) - (int)h->count[l];
        if (left < 0) return -1;
    }
    if (left > 0 && (kind == K_CODES || maxl != 1)) return -1;
    h->kind = left > 0 ? 2 : 0;
    offs[1] = 0;
    for (unsigned l = 1; l < 15; l++) offs[l + 1] = offs[l] + h->count[l];
    for (unsigned i = 0; i < n; i++)
        if (lens[i]) h->sym[offs[lens[i]]++] = (uint16_t)i;
    if (kind != K_CODES) {                   /* inftrees.c:211-212, 226-228, 295-297 */
        unsigned root = kind == K_LENS ? 9 : 6;
        if (root > maxl) root = maxl;
        if (root < minl) root = minl;
        unsigned limit = kind == K_LENS ? (d64 ? 851u : 852u) : (d64 ? 593u : 592u);
        if (table_space(h->count, root, minl, maxl) > limit) return 1;
    }
    return 0;
}

/* Decode one code without consuming it.  Pulls input a byte at a time until the code is resolved
 * (equivalent to the lookup loops at inflate.c:950-954, 1047-1059, 1099-1111).
 * Returns 1 and (*sym,*nb) for a symbol, -1 and *nb for an invalid code, 0 when starved. */
static int huff_peek(orc_stream *s, const huff_t *h, unsigned *sym, unsigned *nb) {
    if (h->kind == 1) {                      /* two-entry error table: 1 bit, val 0 */
        if (!need(s, 1)) return 0;
        *nb = 1; *sym = 0;
        return -1;
    }
    unsigned code = 0, first = 0, index = 0;
    for (unsigned len = 1; len <= 15; len++) {
        if (!need(s, len)) return 0;
        code |= (unsigned)(s->acc >> (len - 1)) & 1;
        unsigned cnt = h->count[len];
        if (code < first + cnt) {
            *sym = h->sym[index + (code - first)];
            *nb = len;
            return 1;
        }
        if (h->kind == 2) { *nb = 1; *sym = 0; return -1; } /* the one hole of a 1-bit code */
        index += cnt;
        first = (first + cnt) << 1;
        code <<= 1;
    }
    *nb = 15; *sym = 0;
    return -1; /* unreachable for complete codes */
}

static const uint16_t len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35,
                                      43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3,
                                      4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t dist_base[32] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193,
                                       257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145,
                                       8193, 12289, 16385, 24577, 32769, 49153};
static const uint8_t dist_extra[32] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8,
                                       9, 9, 10, 10, 11, 11, 12, 12, 13, 13, 14, 14};

static void fixed_codes(orc_stream *s) {     /* inflate.c:271-289 */
    uint16_t l[288];
    unsigned i = 0;
    while (i < 144) l[i++] = 8;
    while (i < 256) l[i++] = 9;
    while (i < 280) l[i++] = 7;
    while (i < 288) l[i++] = 8;
    huff_build(&s->lit, K_LENS, l, 288, s->d64);
    for (i = 0; i < 32; i++) l[i] = 5;
    huff_build(&s->dist, K_DISTS, l, 32, s->d64);
}

/* ---- lifecycle ---- */
static int orc_reset2(orc_stream *s, int wbits) { /* inflate.c:147-184 */
    int wrap;
    if (wbits < 0) {
        if (wbits < -16) return Z_STREAM_ERROR_;
        wrap = 0;
        s->d64 = wbits == -16;
        wbits = -wbits;
    } else {
        wrap = (wbits >> 4) + 5;
        s->d64 = 0;
        if (wbits < 48) wbits &= 15;
    }
    if (wbits && (wbits < 8 || wbits > (s->d64 ? 16 : 15))) return Z_STREAM_ERROR_;
    s->wrap = wrap;
    s->wbits = (unsigned)wbits;
    s->wsize = s->whave = s->wnext = 0;
    s->total_in = s->total_out = s->total = 0;
    s->msg = 0;
    if (wrap) s->adler = (uint32_t)(wrap & 1);
    s->phase = P_HEAD;
    s->last = 0;
    s->gzflags = -1;
    s->acc = 0; s->nacc = 0;
    return Z_OK_;
}

static void fail(orc_stream *s, int msg) { s->msg = msg; s->phase = P_BAD; }

/* append this call's output to the history (inflate.c:375-419) */
static int history_update(orc_stream *s, const uint8_t *end, size_t copy) {
    if (!s->win) {
        s->win = (uint8_t *)malloc((size_t)1 << s->wbits);
        if (!s->win) return 1;
    }
    if (s->wsize == 0) { s->wsize = 1u << s->wbits; s->wnext = 0; s->whave = 0; }
    if (copy >= s->wsize) {
        memcpy(s->win, end - s->wsize, s->wsize);
        s->wnext = 0; s->whave = s->wsize;
        return 0;
    }
    for (size_t i = 0; i < copy; i++) {       /* byte-wise ring append: same final contents */
        s->win[s->wnext++] = end[(ptrdiff_t)i - (ptrdiff_t)copy];
        if (s->wnext == s->wsize) s->wnext = 0;
        if (s->whave < s->wsize) s->whave++;
    }
    return 0;
}

static uint32_t swap32(uint32_t q) {
    return (q >> 24) | ((q >> 8) & 0xff00) | ((q & 0xff00) << 8) | (q << 24);
}
static void crc_word(orc_stream *s, unsigned nbytes) { /* inflate.c:433-448 */
    uint8_t b[4];
    for (unsigned i = 0; i < nbytes; i++) b[i] = (uint8_t)(s->acc >> (8 * i));
    s->check = crc32_upd(s->check, b, nbytes);
}

static int orc_inflate(orc_stream *s, int flush) {
    static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    if (!s->out || (!s->in && s->in_avail)) return Z_STREAM_ERROR_;
    if (s->phase == P_BLOCK) s->phase = P_BLOCKDO;          /* inflate.c:622 */
    const size_t in0 = s->in_avail;
    uint8_t *const out_begin = s->out;       /* first byte written by this call */
    uint8_t *counted = s->out;               /* output before this is already in total/check */
    int ret = Z_OK_;
    int hdrcrc;
    unsigned sym, nb, copy;

    for (;;) switch (s->phase) {
    case P_HEAD:
        if (s->wrap == 0) { s->phase = P_BLOCKDO; break; }
        if (!need(s, 16)) goto leave;
        if ((s->wrap & 2) && s->acc == 0x8b1f) {            /* gzip magic */
            if (s->wbits == 0) s->wbits = 15;
            s->check = 0;
            crc_word(s, 2);
            dropall(s);
            s->phase = P_GZ_FLAGS;
            break;
        }
        if (!(s->wrap & 1) || ((peek(s, 8) << 8) + (unsigned)(s->acc >> 8)) % 31) {
            fail(s, 1); break;
        }
        if (peek(s, 4) != 8) { fail(s, 2); break; }
        drop(s, 4);
        {
            unsigned len = peek(s, 4) + 8;
            if (s->wbits == 0) s->wbits = len;
            if (len > 15 || len > s->wbits) { fail(s, 3); break; }
        }
        s->gzflags = 0;
        s->adler = s->check = 1;
        s->phase = (s->acc & 0x200) ? P_DICTID : P_BLOCK;
        dropall(s);
        break;
    case P_GZ_FLAGS:
        if (!need(s, 16)) goto leave;
        s->gzflags = (int)s->acc;
        if ((s->gzflags & 0xff) != 8) { fail(s, 2); break; }
        if (s->gzflags & 0xe000) { fail(s, 4); break; }
        if ((s->gzflags & 0x0200) && (s->wrap & 4)) crc_word(s, 2);
        dropall(s);
        s->phase = P_GZ_TIME;
        /* fallthrough */
    case P_GZ_TIME:
        if (!need(s, 32)) goto leave;
        if ((s->gzflags & 0x0200) && (s->wrap & 4)) crc_word(s, 4);
        dropall(s);
        s->phase = P_GZ_OS;
        /* fallthrough */
    case P_GZ_OS:
        if (!need(s, 16)) goto leave;
        if ((s->gzflags & 0x0200) && (s->wrap & 4)) crc_word(s, 2);
        dropall(s);
        s->phase = P_GZ_EXLEN;
        /* fallthrough */
    case P_GZ_EXLEN:
        if (s->gzflags & 0x0400) {
            if (!need(s, 16)) goto leave;
            s->length = (unsigned)s->acc;
            if ((s->gzflags & 0x0200) && (s->wrap & 4)) crc_word(s, 2);
            dropall(s);
        }
        s->phase = P_GZ_EXTRA;
        /* fallthrough */
    case P_GZ_EXTRA:
        hdrcrc = (s->gzflags & 0x0200) && (s->wrap & 4);
        if (s->gzflags & 0x0400) {
            copy = s->length;
            if (copy > s->in_avail) copy = (unsigned)s->in_avail;
            if (copy) {
                if (hdrcrc) s->check = crc32_upd(s->check, s->in, copy);
                s->in += copy; s->in_avail -= copy; s->length -= copy;
            }
            if (s->length) goto leave;
        }
        s->length = 0;
        s->phase = P_GZ_NAME;
        /* fallthrough */
    case P_GZ_NAME:
    case P_GZ_COMMENT:
        /* zero-terminated strings: consume through the NUL or to the end of the input */
        while (s->phase == P_GZ_NAME || s->phase == P_GZ_COMMENT) {
            unsigned bit = s->phase == P_GZ_NAME ? 0x0800 : 0x1000;
            hdrcrc = (s->gzflags & 0x0200) && (s->wrap & 4);
            if (s->gzflags & bit) {
                if (s->in_avail == 0) goto leave;
                unsigned c, k = 0;
                do { c = s->in[k++]; } while (c && k < s->in_avail);
                if (hdrcrc) s->check = crc32_upd(s->check, s->in, k);
                s->in += k; s->in_avail -= k;
                if (c) goto leave;
            }
            s->length = 0;
            s->phase = s->phase == P_GZ_NAME ? P_GZ_COMMENT : P_GZ_HCRC;
        }
        /* fallthrough */
    case P_GZ_HCRC:
        if (s->gzflags & 0x0200) {
            if (!need(s, 16)) goto leave;
            if ((s->wrap & 4) && s->acc != (s->check & 0xffff)) { fail(s, 5); break; }
            dropall(s);
        }
        s->adler = s->check = 0;
        s->phase = P_BLOCK;
        break;
    case P_DICTID:
        if (!need(s, 32)) goto leave;
        s->adler = s->check = swap32((uint32_t)s->acc);
        dropall(s);
        s->phase = P_DICT;
        /* fallthrough */
    case P_DICT:
        /* no preset dictionary can be supplied through this surface (inflate.c:824-827):
         * returns without the exit bookkeeping */
        return Z_NEED_DICT_;
    case P_BLOCK:
        if (flush == FL_BLOCK || flush == FL_TREES) goto leave;
        /* fallthrough */
    case P_BLOCKDO:
        if (s->last) { drop(s, s->nacc & 7); s->phase = P_CHECK; break; }
        if (!need(s, 3)) goto leave;
        s->last = (int)peek(s, 1);
        drop(s, 1);
        switch (peek(s, 2)) {
        case 0: s->phase = P_STORED; break;
        case 1:
            fixed_codes(s);
            s->phase = P_SYM_ARMED;
            if (flush == FL_TREES) { drop(s, 2); goto leave; }
            break;
        case 2: s->phase = P_DYN; break;
        default: fail(s, 6);
        }
        drop(s, 2);
        break;
    case P_STORED:
        drop(s, s->nacc & 7);
        if (!need(s, 32)) goto leave;
        if ((s->acc & 0xffff) != ((s->acc >> 16) ^ 0xffff)) { fail(s, 7); break; }
        s->length = (unsigned)s->acc & 0xffff;
        dropall(s);
        s->phase = P_COPY_ARMED;
        if (flush == FL_TREES) goto leave;
        /* fallthrough */
    case P_COPY_ARMED:
        s->phase = P_COPY;
        /* fallthrough */
    case P_COPY:
        copy = s->length;
        if (copy) {
            if (copy > s->in_avail) copy = (unsigned)s->in_avail;
            if (copy > s->out_avail) copy = (unsigned)s->out_avail;
            if (copy == 0) goto leave;
            memcpy(s->out, s->in, copy);
            s->in += copy; s->in_avail -= copy;
            s->out += copy; s->out_avail -= copy;
            s->length -= copy;
            break;
        }
        s->phase = P_BLOCK;
        break;
    case P_DYN:
        if (!need(s, 14)) goto leave;
        s->nlen = peek(s, 5) + 257; drop(s, 5);
        s->ndist = peek(s, 5) + 1;  drop(s, 5);
        s->ncode = peek(s, 4) + 4;  drop(s, 4);
        if (s->nlen > 286 || (!s->d64 && s->ndist > 30)) { fail(s, 8); break; }
        s->have = 0;
        s->phase = P_DYN_CLENS;
        /* fallthrough */
    case P_DYN_CLENS:
        while (s->have < s->ncode) {
            if (!need(s, 3)) goto leave;
            s->lens[order[s->have++]] = (uint16_t)peek(s, 3);
            drop(s, 3);
        }
        while (s->have < 19) s->lens[order[s->have++]] = 0;
        if (huff_build(&s->clen, K_CODES, s->lens, 19, s->d64)) { fail(s, 9); break; }
        s->have = 0;
        s->phase = P_DYN_LENS;
        /* fallthrough */
    case P_DYN_LENS:
        while (s->have < s->nlen + s->ndist) {
            unsigned rep, val;
            if (!huff_peek(s, &s->clen, &sym, &nb)) goto leave;   /* op is not looked at here */
            if (sym < 16) {
                drop(s, nb);
                s->lens[s->have++] = (uint16_t)sym;
                continue;
            }
            if (sym == 16) {
                if (!need(s, nb + 2)) goto leave;
                drop(s, nb);
                if (s->have == 0) { fail(s, 10); break; }
                val = s->lens[s->have - 1];
                rep = 3 + peek(s, 2); drop(s, 2);
            } else if (sym == 17) {
                if (!need(s, nb + 3)) goto leave;
                drop(s, nb);
                val = 0; rep = 3 + peek(s, 3); drop(s, 3);
            } else {
                if (!need(s, nb + 7)) goto leave;
                drop(s, nb);
                val = 0; rep = 11 + peek(s, 7); drop(s, 7);
            }
            if (s->have + rep > s->nlen + s->ndist) { fail(s, 10); break; }
            while (rep--) s->lens[s->have++] = (uint16_t)val;
        }
        if (s->phase == P_BAD) break;
        if (s->lens[256] == 0) { fail(s, 11); break; }
        if (huff_build(&s->lit, K_LENS, s->lens, s->nlen, s->d64)) { fail(s, 12); break; }
        if (huff_build(&s->dist, K_DISTS, s->lens + s->nlen, s->ndist, s->d64)) { fail(s, 13); break; }
        s->phase = P_SYM_ARMED;
        if (flush == FL_TREES) goto leave;
        /* fallthrough */
    case P_SYM_ARMED:
        s->phase = P_SYM;
        /* fallthrough */
    case P_SYM: {
        int k = huff_peek(s, &s->lit, &sym, &nb);
        if (k == 0) goto leave;
        drop(s, nb);
        if (k < 0 || sym > 285) { fail(s, 14); break; }
        if (sym < 256) { s->length = sym; s->phase = P_LIT; break; }
        if (sym == 256) { s->phase = P_BLOCK; break; }
        sym -= 257;
        s->length = len_base[sym];
        s->extra = len_extra[sym];
        if (sym == 28 && s->d64) { s->length = 3; s->extra = 16; }   /* inftrees.c:73-79 */
        s->phase = P_LENEXT;
    }
        /* fallthrough */
    case P_LENEXT:
        if (s->extra) {
            if (!need(s, s->extra)) goto leave;
            s->length += peek(s, s->extra);
            drop(s, s->extra);
        }
        s->phase = P_DIST;
        /* fallthrough */
    case P_DIST: {
        int k = huff_peek(s, &s->dist, &sym, &nb);
        if (k == 0) goto leave;
        drop(s, nb);
        if (k < 0 || (sym >= 30 && !s->d64)) { fail(s, 15); break; }
        s->offset = dist_base[sym];
        s->extra = dist_extra[sym];
        s->phase = P_DISTEXT;
    }
        /* fallthrough */
    case P_DISTEXT:
        if (s->extra) {
            if (!need(s, s->extra)) goto leave;
            s->offset += peek(s, s->extra);
            drop(s, s->extra);
        }
        s->phase = P_MATCH;
        /* fallthrough */
    case P_MATCH: {
        if (s->out_avail == 0) goto leave;
        size_t made = (size_t)(s->out - out_begin);
        copy = s->length;
        if (copy > s->out_avail) copy = (unsigned)s->out_avail;
        if (s->offset > made) {                               /* reaches into earlier calls */
            unsigned back = s->offset - (unsigned)made;
            if (back > s->whave) { fail(s, 16); break; }
            if (copy > back) copy = back;                     /* then continue from this call */
            unsigned pos = (s->wnext + s->wsize - back) % s->wsize;
            for (unsigned i = 0; i < copy; i++) {
                *s->out++ = s->win[pos];
                if (++pos == s->wsize) pos = 0;
            }
        } else {
            const uint8_t *from = s->out - s->offset;
            for (unsigned i = 0; i < copy; i++) *s->out++ = *from++;
        }
        s->out_avail -= copy;
        s->length -= copy;
        if (s->length == 0) s->phase = P_SYM;
        break;
    }
    case P_LIT:
        if (s->out_avail == 0) goto leave;
        *s->out++ = (uint8_t)s->length;
        s->out_avail--;
        s->phase = P_SYM;
        break;
    case P_CHECK:
        if (s->wrap) {
            if (!need(s, 32)) goto leave;
            size_t n = (size_t)(s->out - counted);
            s->total_out += n; s->total += n;
            if ((s->wrap & 4) && n) s->adler = s->check = check_upd(s, s->check, counted, n);
            counted = s->out;
            if ((s->wrap & 4) &&
                (s->gzflags ? (uint32_t)s->acc : swap32((uint32_t)s->acc)) != s->check) {
                fail(s, 17); break;
            }
            dropall(s);
        }
        s->phase = P_LENGTH;
        /* fallthrough */
    case P_LENGTH:
        if (s->wrap && s->gzflags) {
            if (!need(s, 32)) goto leave;
            if ((s->wrap & 4) && (uint32_t)s->acc != (uint32_t)(s->total & 0xffffffffu)) {
                fail(s, 18); break;
            }
            dropall(s);
        }
        s->phase = P_DONE;
        /* fallthrough */
    case P_DONE:
        ret = Z_STREAM_END_;
        goto leave;
    case P_BAD:
        ret = Z_DATA_ERROR_;
        goto leave;
    default:
        return Z_STREAM_ERROR_;
    }

leave: {                                                     /* inflate.c:1253-1274 */
    size_t n = (size_t)(s->out - counted);
    if (s->wsize || (n && s->phase < P_BAD && (s->phase < P_CHECK || flush != FL_FINISH)))
        if (history_update(s, s->out, n)) return Z_MEM_ERROR_;
    size_t in_used = in0 - s->in_avail;
    s->total_in += in_used;
    s->total_out += n; s->total += n;
    if ((s->wrap & 4) && n) s->adler = s->check = check_upd(s, s->check, counted, n);
    s->data_type = (int)s->nacc + (s->last ? 64 : 0) + (s->phase == P_BLOCK ? 128 : 0) +
                   (s->phase == P_SYM_ARMED || s->phase == P_COPY_ARMED ? 256 : 0);
    if (((in_used == 0 && n == 0) || flush == FL_FINISH) && ret == Z_OK_)
        ret = Z_BUF_ERROR_;
    return ret;
}
}

/* ---- harness surface (oracle_api.h) ---- */
static void fill(const orc_stream *s, int ret, orc_result *r) {
    r->ret = ret; r->msg_id = s->msg; r->total_in = s->total_in; r->total_out = s->total_out;
    r->adler = s->adler; r->data_type = s->data_type;
}

void *orc_open(int wbits, int *init_ret) {
    orc_stream *s = (orc_stream *)calloc(1, sizeof *s);
    int ret = s ? orc_reset2(s, wbits) : Z_MEM_ERROR_;
    if (init_ret) *init_ret = ret;
    if (ret != Z_OK_) { free(s); return 0; }
    return s;
}

int orc_step(void *h, const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int flush,
             size_t *consumed, size_t *produced, orc_result *r) {
    orc_stream *s = (orc_stream *)h;
    static uint8_t dummy;
    s->in = in; s->in_avail = in_len;
    s->out = out ? out : &dummy; s->out_avail = out_cap;
    int ret = orc_inflate(s, flush);
    if (consumed) *consumed = in_len - s->in_avail;
    if (produced) *produced = out_cap - s->out_avail;
    fill(s, ret, r);
    return ret;
}

void orc_close(void *h) {
    orc_stream *s = (orc_stream *)h;
    if (!s) return;
    free(s->win);
    free(s);
}

int orc_oneshot(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_cap, int wbits,
                orc_result *r) {
    int ret;
    void *h = orc_open(wbits, &ret);
    if (!h) { memset(r, 0, sizeof *r); r->ret = ret; return ret; }
    ret = orc_step(h, in, in_len, out, out_cap, FL_FINISH, 0, 0, r);
    orc_close(h);
    return ret;
}

typedef struct {
    size_t n; const uint8_t *const *in; const size_t *in_len; const size_t *out_cap;
    const int *wbits; int tid, threads; uint64_t sum_out, bad; uint8_t *buf; size_t buf_cap;
} bench_arg;

static void *bench_worker(void *p) {
    bench_arg *a = (bench_arg *)p;
    a->sum_out = a->bad = 0;
    for (size_t i = (size_t)a->tid; i < a->n; i += (size_t)a->threads) {
        orc_result r;
        size_t cap = a->out_cap[i] <= a->buf_cap ? a->out_cap[i] : a->buf_cap;
        orc_oneshot(a->in[i], a->in_len[i], a->buf, cap, a->wbits[i], &r);
        a->sum_out += r.total_out;
        if (r.ret != Z_STREAM_END_) a->bad++;
    }
    return 0;
}

double orc_bench(size_t n, const uint8_t *const *in, const size_t *in_len, const size_t *out_cap,
                 const int *wbits, int threads, int reps, uint64_t *sum_out, uint64_t *bad) {
    size_t maxcap = 1;
    for (size_t i = 0; i < n; i++) if (out_cap[i] > maxcap) maxcap = out_cap[i];
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
