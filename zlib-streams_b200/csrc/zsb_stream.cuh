// zsb_stream.cuh — the resumable single-stream decoder behind the z_stream surface (inflateInit2_ /
// inflate / inflateEnd of include/zsb.h, replacing reference src/inflate.c:186-224, 597-1287).
//
// One inflate() call = one launch of k_zs_decode (this file: symbols -> LZ sequences, the lane path's
// decoder with one active lane), one of k_lz's resolver over those sequences, one of k_zs_finish
// (running check value, trailer, sliding history).  Between calls the stream lives in device memory:
// a State (resume point + Huffman tables), up to 64 KiB of history and a few carried input bytes.
//
// Resuming never re-enters the middle of a unit.  A unit is a wrapper header, a block header, one
// symbol or one trailer word; when the input ends inside a unit the call reports what the reference
// reports (all input consumed, the same bit count in data_type) but keeps the unit's bytes as CARRY
// and decodes the unit again from its first bit when more input arrives.  What the reference holds
// in its `hold`/`bits`/`mode` (src/inflate.h:83-128) is therefore {phase, bit0, carry}.
#pragma once
#include "zsb_lz.cuh"

namespace zsb {
namespace zs {

enum : uint32_t { PH_HEAD = 0, PH_BLOCK = 1, PH_STORED = 2, PH_SYM = 3, PH_PENDLIT = 4, PH_PENDMATCH = 5,
                  PH_CHECK = 6, PH_LENGTH = 7, PH_DONE = 8, PH_BAD = 9, PH_NEEDDICT = 10,
                  PH_DICTSET = 11 };      // inflateSetDictionary() answered Z_NEED_DICT: mode TYPE is entered in the next call

constexpr uint32_t HIST = 65536;          // bytes of history kept in front of every call's output

// result of one launch sequence, read back by the host
struct Call {
    int32_t stop;          // EV_STARVE, EV_OUTFULL, EV_DONE, EV_ERR, EV_NEEDDICT
    int32_t msg;           // message id for EV_ERR
    uint32_t cpos;         // bytes of the (carry + new) input the reference has consumed
    uint32_t resume_byte;  // first byte that has to be presented again (carry = [resume_byte, cpos))
    uint32_t produced;
    uint32_t data_type;    // strm->data_type
    uint32_t adler;        // strm->adler
    uint32_t nseq;
    uint32_t trailer_pos;  // when need_trailer: byte offset of the trailer
    uint32_t need_trailer;
    int32_t mark;          // inflateMark()'s value at this stop (src/inflate.c:1521-1530)
    uint32_t codes_used;   // inflateCodesUsed()'s value (src/inflate.c:1532)
    uint32_t sync_point;   // inflateSyncPoint(): waiting for a stored block's length bytes with no bits held
    uint32_t whave;        // history bytes the reference's window would hold (inflateGetDictionary)
    uint32_t hdr_len;      // nonzero in the call that completed the zlib / gzip header: its length in bytes
    uint32_t hdr_kind;     // 1 zlib, 2 gzip (with hdr_len)
    uint32_t bit0;         // bit offset of the resume point inside the first carried byte
    uint32_t wprod;        // bytes of this call's output that enter the sliding window (see finish_call)
    uint32_t pad[2];
};

struct State {
    uint32_t phase, bit0, last, d64;
    int32_t wbits_arg;
    uint32_t kind;         // 0 raw, 1 zlib, 2 gzip (known once the header is read)
    uint32_t rem;          // stored bytes / match bytes still to produce
    uint32_t mdist, plit;
    uint32_t whave, wsize, wbits;   // history accounting as src/inflate.c:375-419 does it
    uint32_t check;        // running Adler-32 / CRC-32 of all output
    uint32_t total_lo;     // output so far, modulo 2^32 (gzip ISIZE)
    uint32_t last_dt;      // data_type low bits of the last stop (for calls after the end)
    uint32_t noverify;     // inflateValidate(strm, 0) / inflateSync(): wrap & 4 cleared — no check value is computed or compared
    uint32_t was;          // full length of the match being copied (inflateMark)
    int32_t back;          // bits of the current symbol's codes consumed so far, -1 outside symbols (inflateMark)
    uint32_t codes_used;
    uint32_t pad;
    Call call;
    alignas(16) uint8_t tables[lane::SLOT];
};

using lane::Lane;
using lane::seq_t;

struct Run {               // lane 0's bookkeeping during one call
    uint32_t phase;
    uint64_t unit;         // bit position where the current unit started
    int stop;
    uint32_t last0;        // BFINAL as it was when the current block header unit started
    bool hdr_stop;
    uint32_t hdr_tin, hdr_bits;
    const uint32_t *big_lut;   // direct tables of the current block (lane::big_build), or NULL
    uint32_t sync_point;
    uint64_t first_bit;        // bit position this call resumed at
    bool resumed_partial;      // ... inside a symbol whose first codes the reference has consumed already (state->back > 0)
    uint32_t hdr_len, hdr_kind;
    uint32_t flush_mode;       // 0, FL_BLOCK (inflate(Z_BLOCK)) or FL_TREES (inflate(Z_TREES))
};
// inflate()'s flush values that add early returns (reference src/inflate.c:832,854-857,883,1032)
enum : uint32_t { FL_NONE = 0, FL_BLOCK = 1, FL_TREES = 2 };

ZSB_DEV uint64_t bitpos(const BitIn &b) { return 8ull * b.next - b.nbits; }

// position the bit reader on an arbitrary bit
ZSB_DEV void seek_bit(BitIn &b, uint64_t bp) {
    bi_seek(b, (uint32_t)(bp >> 3));
    const uint32_t r = (uint32_t)(bp & 7);
    if (r) {
        bi_fill(b);
        bi_drop(b, r);
    }
}

// Runs lane 0 until the call stops (returns 0), a dynamic block's tables have to be built (1) or the
// fixed tables have to be copied (2).  R.stop is set when 0 is returned.
ZSB_DEV int run_serial(State &S, Lane &L, Run &R, uint8_t *T, const uint32_t *lut, bool &primed) {
    Dec &d = L.d;
    BitIn &b = d.b;
    for (;;) {
        switch (R.phase) {
        case PH_DICTSET:                           // src/inflate.c:822-830: DICT with the dictionary set -> TYPE
            L.mm.kind = 1;
            d.at_type = 1;
            R.phase = PH_BLOCK;
            break;
        case PH_HEAD: {
            HeadOut h = parse_wrapper(b.p, b.len, S.wbits_arg);
            L.mm.kind = h.kind;
            L.mm.adler0 = h.adler;
            if (h.ev != EV_NONE) {
                R.hdr_stop = true;
                d.msg = h.msg;
                R.hdr_tin = h.ev == EV_STARVE ? b.len : h.pulled;
                R.hdr_bits = h.bits;
                R.unit = 0;
                R.stop = h.ev;
                return 0;
            }
            S.kind = h.kind;
            S.check = h.adler;
            R.hdr_len = h.pos;
            R.hdr_kind = h.kind;
            if (S.wbits == 0) S.wbits = h.kind == 1 ? (uint32_t)(b.p[0] >> 4) + 8u : 15u;
            seek_bit(b, 8ull * h.pos + (h.kind == 0 ? S.bit0 : 0u));      // (a raw stream may have been primed: inflatePrime)
            d.at_type = h.kind != 0;
            R.phase = PH_BLOCK;
            break;
        }
        case PH_BLOCK: {
            R.unit = bitpos(b);
            R.last0 = d.last;
            // mode TYPE reached inside this call (header end, end of a block): Z_BLOCK / Z_TREES leave here
            // (src/inflate.c:831-832); a call that STARTS here goes on (TYPE -> TYPEDO, src/inflate.c:625)
            if (R.flush_mode && d.at_type) {
                R.stop = EV_BLOCKSTOP;
                return 0;
            }
            if (d.last) {
                bi_drop(b, b.nbits & 7);
                R.unit = bitpos(b);
                R.stop = EV_DONE;
                return 0;
            }
            if (!bi_need(b, 3)) {
                R.stop = EV_STARVE;
                return 0;
            }
            d.last = bi_peek(b, 1);
            bi_drop(b, 1);
            const uint32_t btype = bi_peek(b, 2);
            bi_drop(b, 2);
            d.at_type = 0;
            if (btype == 3) {
                d.msg = M_BTYPE;
                R.stop = EV_ERR;
                return 0;
            }
            if (btype == 2) {
                L.ndist = 0;
                const int ev = lane::lane_dynamic_header(L, T, lut);
                // inflateCodesUsed(): once the code-length code is accepted its table is what the reference holds
                // (src/inflate.c:933-945), until the two block codes replace it; a rejected one leaves none
                if (ev != EV_NONE && (L.ndist & 0x80000000u)) S.codes_used = L.ndist & 0xffffu;
                else if (ev == EV_ERR && d.msg == M_CLENS) S.codes_used = 0;
                if (ev != EV_NONE) {
                    R.stop = ev;
                    return 0;
                }
                return 1;
            }
            if (btype == 1) return 2;
            bi_drop(b, b.nbits & 7);
            const uint32_t bytepos = b.next - (b.nbits >> 3);
            if (b.len - bytepos < 4) {
                R.sync_point = b.len == bytepos;      // mode STORED with an empty bit buffer (inflateSyncPoint)
                b.nbits = 8 * (b.len - bytepos);
                b.next = b.len;
                R.stop = EV_STARVE;
                return 0;
            }
            const uint32_t v = le32(b.p + bytepos);
            if ((v & 0xffffu) != ((v >> 16) ^ 0xffffu)) {
                d.msg = M_STORED;
                d.extent = 32;
                R.stop = EV_ERR;
                return 0;
            }
            S.rem = v & 0xffffu;
            seek_bit(b, 8ull * (bytepos + 4));
            R.phase = PH_STORED;
            if (R.flush_mode == FL_TREES) {           // mode COPY_ (src/inflate.c:882-883)
                R.unit = bitpos(b);
                R.stop = EV_BLOCKSTOP;
                return 0;
            }
            break;
        }
        case PH_STORED: {
            const uint32_t bytepos = b.next - (b.nbits >> 3);
            uint32_t n = S.rem;
            if (n > b.len - bytepos) n = b.len - bytepos;
            if (n > d.cap - d.pos) n = d.cap - d.pos;
            lane::push_raw(L, bytepos, n);
            S.rem -= n;
            seek_bit(b, 8ull * (bytepos + n));
            R.unit = bitpos(b);
            if (S.rem) {                          // src/inflate.c:886-890: no input or no room left
                b.next = bytepos + n;
                b.nbits = 0;
                b.hold = 0;
                R.stop = bytepos + n >= b.len ? EV_STARVE : EV_OUTFULL;
                return 0;
            }
            d.at_type = 1;
            R.phase = PH_BLOCK;
            break;
        }
        case PH_PENDLIT:
            R.unit = bitpos(b);
            if (d.pos >= d.cap) {
                R.stop = EV_OUTFULL;
                return 0;
            }
            lane::push_literal(L, S.plit);
            R.phase = PH_SYM;
            break;
        case PH_PENDMATCH: {
            R.unit = bitpos(b);
            if (d.pos >= d.cap) {
                R.stop = EV_OUTFULL;
                return 0;
            }
            if (S.mdist > d.pos + L.hist) {       // src/inflate.c:1146-1153
                d.msg = M_TOOFAR;
                R.stop = EV_ERR;
                return 0;
            }
            uint32_t n = d.cap - d.pos;
            if (n > S.rem) n = S.rem;
            lane::push_match(L, n, S.mdist);
            S.rem -= n;
            if (S.rem) {
                R.stop = EV_OUTFULL;
                return 0;
            }
            R.phase = PH_SYM;
            break;
        }
        case PH_SYM: {
            // plain literals and whole matches through the direct tables (the loop of k_big); it stops at a
            // symbol boundary, and whatever stopped it is decoded by the checked path below
            // (a symbol the previous call was cut in goes through the checked path: the reference finishes it on its slow
            //  path, and what that leaves in state->back outlives the inflate_fast() run that follows)
            if (R.big_lut && !(R.resumed_partial && bitpos(b) == R.first_bit)) lane::big_run(L, T, lut, R.big_lut);
            R.unit = bitpos(b);
            // Would the reference decode this symbol inside inflate_fast() (src/inflate.c:1038-1044: at least 6 input
            // bytes and 258 bytes of room at the symbol's start, never in Deflate64 mode)?  Its bit buffer holds less
            // than a byte between symbols on both paths, so the input left is a function of the bit position.
            // (a symbol cut by the end of the previous call's input is finished on the slow path: modes LENEXT / DIST / DISTEXT)
            const bool fastsym = !S.d64 && b.len - (uint32_t)((R.unit + 7) >> 3) >= 6u && d.cap - d.pos >= 258u &&
                                 !(R.resumed_partial && R.unit == R.first_bit);
            const int ev = lane::careful_symbol(d, T, lut);
            // the bits of this symbol's codes consumed so far: what the reference keeps in state->back
            // (src/inflate.c:1046-1142); inflate_fast() does not touch it, whatever it runs into
            if (!fastsym) S.back = (int32_t)(bitpos(b) - R.unit);
            if (ev == EV_LIT) {
                S.plit = d.mdist;
                R.phase = PH_PENDLIT;
            } else if (ev == EV_MATCH) {
                S.rem = d.mlen;
                S.was = d.mlen;
                S.mdist = d.mdist;
                R.phase = PH_PENDMATCH;
            } else if (ev == EV_EOB) {
                d.at_type = 1;
                S.back = -1;
                R.phase = PH_BLOCK;
            } else {
                R.stop = ev;                      // EV_ERR or EV_STARVE (the unit is decoded again later)
                return 0;
            }
            break;
        }
        default:                                  // PH_CHECK, PH_LENGTH: the finish kernel reads the trailer
            R.unit = bitpos(b);
            R.stop = EV_DONE;
            return 0;
        }
    }
}

// One call's decode: all W lanes enter, lane 0 owns the stream.
template <int W>
ZSB_DEV void stream_decode(State &S, const uint8_t *in, uint32_t in_len, uint32_t out_cap, seq_t *seq,
                           const uint8_t *fixed_tab, uint8_t *scr, uint32_t *ring, const uint32_t *lut, int ln,
                           uint32_t *big_lut = nullptr, uint32_t flush_mode = FL_NONE) {
    Lane L;
    Run R;
    Dec &d = L.d;
    uint8_t *T = S.tables;
    bool primed = false;
    const bool active = ln == 0;
    d.b.p = in; d.b.len = in_len; d.b.next = 0; d.b.nbits = 0; d.b.hold = 0;
    d.out = nullptr; d.pos = 0; d.cap = out_cap; d.d64 = S.d64; d.last = S.last; d.at_type = 0; d.msg = 0;
    d.extent = 0; d.mlen = 0; d.mdist = 0; d.mtrunc = 0;
    L.mm.idx = 0; L.mm.kind = S.kind; L.mm.adler0 = S.check; L.mm.hdr_stop = false; L.mm.hdr_tin = 0; L.mm.hdr_bits = 0;
    L.q0 = nullptr; L.qend = seq + (1u << 30); L.qbase = 0;     // big_run's queue guard: the call's array is bounded by the host
    L.seq = L.seq0 = seq; L.lw = 0; L.nl = 0; L.ring = ring; L.hist = S.whave; L.nlen = 0; L.ndist = 0;
    R.phase = S.phase; R.unit = 0; R.stop = EV_NONE; R.last0 = S.last; R.hdr_stop = false; R.hdr_tin = 0; R.hdr_bits = 0;
    if (active && R.phase != PH_HEAD) seek_bit(d.b, S.bit0);
    R.first_bit = bitpos(d.b);
    R.resumed_partial = S.phase == PH_SYM && S.back > 0;
    R.big_lut = big_lut;
    R.sync_point = 0;
    R.hdr_len = 0;
    R.hdr_kind = 0;
    R.flush_mode = flush_mode;
    // a call that resumes inside a block: its tables are in S.tables, the direct tables are rebuilt from them
    if (big_lut && (S.phase == PH_SYM || S.phase == PH_PENDLIT || S.phase == PH_PENDMATCH))
        lane::big_build<W>(big_lut, T, lut, S.d64, ln);
    for (;;) {
        int want = 0;
        if (active) want = run_serial(S, L, R, T, lut, primed);
        lane::Wp<W>::sync();                  // lane 0's code lengths (in S.tables) visible to the warp
        want = (int)lane::Wp<W>::shfl((uint32_t)want, 0);
        if (want == 0) break;
        const uint32_t info = lane::Wp<W>::shfl(L.nlen | (L.ndist << 12), 0);
        int msg = 0;
        if (want == 2) {
            for (int i = ln; i < lane::SLOT / 2; i += W) ((uint16_t *)T)[i] = ((const uint16_t *)fixed_tab)[i];
            lane::Wp<W>::sync();
        } else {
            msg = lane::build_tables<W>(T, scr, info & 1023u, info >> 12, S.d64 != 0, ln);
        }
        if (big_lut && !msg) lane::big_build<W>(big_lut, T, lut, S.d64, ln);
        if (active) {
            // inflateCodesUsed(): entries of the tables the reference built last (src/inflate.c:992-1027); a rejected
            // literal/length set leaves none, a rejected distance set leaves the literal/length table's
            if (want == 1)
                S.codes_used = msg == M_LITLENS ? 0u : ((const uint32_t *)(scr + lane::W_LIM))[0] +
                               (msg == M_DISTS ? 0u : ((const uint32_t *)(scr + lane::W_OFF0))[0]);
            if (msg) {
                d.msg = msg;
                R.stop = EV_ERR;
            } else {
                R.phase = PH_SYM;
                S.back = -1;                              // a block starts: the reference left -1 at the last end of block
                if (flush_mode == FL_TREES) {             // mode LEN_ (src/inflate.c:854-857, 1031-1032)
                    R.unit = bitpos(d.b);
                    R.stop = EV_BLOCKSTOP;
                }
            }
        }
        if (lane::Wp<W>::shfl((uint32_t)msg, 0) || flush_mode == FL_TREES) break;
    }
    if (!active) return;
    // ---- what the reference leaves on the stream at this stop (write_result, zsb_device.cuh) ----
    lane::flush_literals(L);
    zsb_result r;
    Aux a;
    L.mm.hdr_stop = R.hdr_stop;
    L.mm.hdr_tin = R.hdr_tin;
    L.mm.hdr_bits = R.hdr_bits;
    write_result(d, L.mm, R.stop, &r, &a);
    Call &c = S.call;
    c.stop = R.stop;
    c.msg = r.msg_id;
    c.cpos = (uint32_t)r.total_in;
    c.produced = d.pos;
    c.data_type = (uint32_t)r.data_type;
    c.adler = r.check;
    c.nseq = (uint32_t)(L.seq - L.seq0);
    c.need_trailer = a.need_trailer;
    c.trailer_pos = a.trailer_pos;
    c.sync_point = R.sync_point;
    c.bit0 = 0;
    c.hdr_len = R.hdr_len;
    c.hdr_kind = R.hdr_kind;
    c.codes_used = S.codes_used;
    // inflateMark(): (back << 16) + bytes of the current stored block still to copy (mode COPY) / bytes of the current
    // match already copied (mode MATCH)
    {
        const bool in_symbol = R.phase == PH_SYM || R.phase == PH_PENDLIT || R.phase == PH_PENDMATCH;
        if (!in_symbol) S.back = -1;
        int32_t low = 0;
        if (R.phase == PH_STORED && R.stop != EV_BLOCKSTOP && R.stop != EV_ERR) low = (int32_t)S.rem;
        if (R.phase == PH_PENDMATCH && R.stop != EV_ERR) low = (int32_t)(S.was - S.rem);
        c.mark = (int32_t)((uint32_t)S.back << 16) + low;
    }
    if (R.phase == PH_CHECK || R.phase == PH_LENGTH) {     // resumed inside the trailer
        c.need_trailer = 1;
        c.trailer_pos = 0;
    }
    // ---- resume point ----
    S.last = d.last;
    if (R.stop == EV_STARVE && R.phase == PH_BLOCK) S.last = R.last0;   // the header is read again
    if (R.stop == EV_STARVE || R.stop == EV_OUTFULL || R.stop == EV_BLOCKSTOP) {
        uint64_t rb = R.unit;
        if (R.stop == EV_STARVE && R.phase == PH_STORED) rb = 8ull * in_len;
        c.resume_byte = (uint32_t)(rb >> 3);
        S.bit0 = (uint32_t)(rb & 7);
        c.bit0 = S.bit0;
        S.phase = R.phase;
        if (c.resume_byte > c.cpos) c.resume_byte = c.cpos;
    } else if (R.stop == EV_DONE) {
        c.resume_byte = c.cpos;
        S.bit0 = 0;
        S.phase = a.need_trailer || c.need_trailer ? (R.phase == PH_LENGTH ? PH_LENGTH : PH_CHECK) : PH_DONE;
    } else {
        c.resume_byte = c.cpos;
        S.bit0 = 0;
        S.phase = R.stop == EV_NEEDDICT ? PH_NEEDDICT : PH_BAD;
    }
}


// What k_zs_finish's first thread does once the check value of this call's output is known: running
// check and length, the trailer (reference src/inflate.c:1194-1232), the window accounting
// (src/inflate.c:1255-1260, 375-419).  `check` = S.check continued over the new output.
ZSB_DEV void finish_call(State &S, const uint8_t *in, uint32_t in_len, uint32_t check, bool finish) {
    Call c = S.call;
    const uint32_t produced = c.produced, kind = S.kind;
    S.total_lo += produced;
    if (kind != 0 && !S.noverify) {   // kind is only set once the wrapper header has been read
        S.check = check;
        c.adler = check;
    }
    // The reference zeroes its per-call output count once it has read the check word of a zlib / gzip trailer
    // (src/inflate.c:1196-1200: out -= left ... out = left), so what such a call produced before the trailer never
    // reaches updatewindow() (:1255-1260): neither whave nor the window contents change.
    bool window_skip = false;
    if (c.need_trailer && c.stop == EV_DONE) {
        uint32_t pos = c.trailer_pos, phase = S.phase;
        bool done = false;
        if (phase == PH_CHECK) {
            if (in_len - pos >= 4) window_skip = true;
            if (in_len - pos < 4) {
                c.stop = EV_STARVE; c.cpos = in_len; c.resume_byte = pos; c.data_type = 8 * (in_len - pos) + 64;
                done = true;
            } else {
                uint32_t v = le32(in + pos);
                if (kind == 1) v = (v >> 24) | ((v >> 8) & 0xff00u) | ((v & 0xff00u) << 8) | (v << 24);
                pos += 4;
                if (!S.noverify && v != check) {
                    c.stop = EV_ERR; c.msg = M_DATACHECK; c.cpos = pos; c.resume_byte = pos; c.data_type = 32 + 64;
                    S.phase = PH_BAD;
                    done = true;
                } else {
                    phase = PH_LENGTH;
                }
            }
        }
        if (!done && kind == 2) {
            if (in_len - pos < 4) {
                c.stop = EV_STARVE; c.cpos = in_len; c.resume_byte = pos; c.data_type = 8 * (in_len - pos) + 64;
                S.phase = PH_LENGTH;
                done = true;
            } else {
                const uint32_t isz = le32(in + pos);
                pos += 4;
                if (!S.noverify && isz != S.total_lo) {
                    c.stop = EV_ERR; c.msg = M_LENCHECK; c.cpos = pos; c.resume_byte = pos; c.data_type = 32 + 64;
                    S.phase = PH_BAD;
                    done = true;
                }
            }
        }
        if (!done) {
            c.stop = EV_DONE; c.cpos = pos; c.resume_byte = pos; c.data_type = 64;
            S.phase = PH_DONE;
        }
    }
    // history the reference would keep (src/inflate.c:1255-1260, 375-419)
    const uint32_t ph = S.phase;
    const uint32_t wprod = window_skip ? 0u : produced;
    c.wprod = 0;
    if (S.wsize || (wprod && ph < PH_BAD && (ph < PH_CHECK || !finish))) {
        S.wsize = 1u << S.wbits;
        const uint64_t h = (uint64_t)S.whave + wprod;
        S.whave = h > S.wsize ? S.wsize : (uint32_t)h;
        c.wprod = wprod;
    }
    c.whave = S.whave;
    S.call = c;
}

}  // namespace zs
}  // namespace zsb
