// zsb_zip.cuh — ZIP container walker: archive directory -> batch descriptors -> verified entries
// (SURVEY.md §8f rank 1).  Host code, included by zsb_host.cu; entry data is only touched by the GPU
// (decode kernels + the CRC-32 of k_check).  Layout facts are APPNOTE.TXT's; the reference's generator
// reads the same local-header fields (src/generator/create_deflate64.cpp:51-93).
#pragma once

namespace zsbzip {

static inline uint16_t rd16(const uint8_t *p) { return (uint16_t)(p[0] | (p[1] << 8)); }
static inline uint32_t rd32(const uint8_t *p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }
static inline uint64_t rd64(const uint8_t *p) { return (uint64_t)rd32(p) | ((uint64_t)rd32(p + 4) << 32); }

constexpr uint32_t SIG_LOCAL = 0x04034b50u, SIG_CENTRAL = 0x02014b50u, SIG_EOCD = 0x06054b50u;
constexpr uint32_t SIG_EOCD64 = 0x06064b50u, SIG_LOC64 = 0x07064b50u;

// [off, off + n) inside an archive of len bytes
static inline bool inside(uint64_t off, uint64_t n, uint64_t len) { return off <= len && n <= len - off; }

static int scan(const uint8_t *z, uint64_t len, zsb_zip_entry *out, size_t max_entries, size_t *count) {
    if (count) *count = 0;
    if (len < 22) return ZSB_E_ZIP;
    // end of central directory: the last signature within the final 22 + 65535 bytes whose comment fits
    uint64_t eocd = UINT64_MAX;
    const uint64_t lowest = len > 22 + 65535 ? len - 22 - 65535 : 0;
    for (uint64_t p = len - 22;; p--) {
        if (rd32(z + p) == SIG_EOCD && p + 22 + rd16(z + p + 20) <= len) {
            eocd = p;
            break;
        }
        if (p == lowest) break;
    }
    if (eocd == UINT64_MAX) return ZSB_E_ZIP;
    uint64_t total = rd16(z + eocd + 10), cd_size = rd32(z + eocd + 12), cd_off = rd32(z + eocd + 16);
    if (rd16(z + eocd + 4) != rd16(z + eocd + 6) && rd16(z + eocd + 4) != 0xffff) return ZSB_E_ZIP;   // spans disks
    if (total == 0xffff || cd_size == 0xffffffffu || cd_off == 0xffffffffu) {
        if (eocd < 20 || rd32(z + eocd - 20) != SIG_LOC64) return ZSB_E_ZIP;
        const uint64_t e64 = rd64(z + eocd - 20 + 8);
        if (!inside(e64, 56, len) || rd32(z + e64) != SIG_EOCD64) return ZSB_E_ZIP;
        total = rd64(z + e64 + 32);
        cd_size = rd64(z + e64 + 40);
        cd_off = rd64(z + e64 + 48);
    }
    if (!inside(cd_off, cd_size, len)) return ZSB_E_ZIP;
    if (total > cd_size / 46) return ZSB_E_ZIP;
    uint64_t p = cd_off;
    const uint64_t cd_end = cd_off + cd_size;
    for (uint64_t i = 0; i < total; i++) {
        if (!inside(p, 46, cd_end) || rd32(z + p) != SIG_CENTRAL) return ZSB_E_ZIP;
        const uint16_t nlen = rd16(z + p + 28), xlen = rd16(z + p + 30), clen = rd16(z + p + 32);
        if (!inside(p + 46, (uint64_t)nlen + xlen + clen, cd_end)) return ZSB_E_ZIP;
        zsb_zip_entry e;
        memset(&e, 0, sizeof e);
        e.gp_flags = rd16(z + p + 8);
        e.method = rd16(z + p + 10);
        e.crc32 = rd32(z + p + 16);
        e.comp_size = rd32(z + p + 20);
        e.uncomp_size = rd32(z + p + 24);
        e.header_off = rd32(z + p + 42);
        e.name_off = (uint32_t)(p + 46);
        e.name_len = nlen;
        if (p + 46 > 0xffffffffull) e.name_len = 0;          // name_off does not reach: the name is not reported
        // ZIP64 extended information (id 1): the fields whose 32-bit value is all ones, in this order
        const uint8_t *x = z + p + 46 + nlen, *xe = x + xlen;
        while (xe - x >= 4) {
            const uint16_t id = rd16(x), sz = rd16(x + 2);
            if (sz > xe - x - 4) return ZSB_E_ZIP;
            if (id == 1) {
                const uint8_t *q = x + 4, *qe = q + sz;
                if (e.uncomp_size == 0xffffffffu) { if (qe - q < 8) return ZSB_E_ZIP; e.uncomp_size = rd64(q); q += 8; }
                if (e.comp_size == 0xffffffffu) { if (qe - q < 8) return ZSB_E_ZIP; e.comp_size = rd64(q); q += 8; }
                if (e.header_off == 0xffffffffu) { if (qe - q < 8) return ZSB_E_ZIP; e.header_off = rd64(q); q += 8; }
            }
            x += 4 + sz;
        }
        // the local header decides where the data starts (its name / extra lengths may differ from the directory's)
        if (!inside(e.header_off, 30, len) || rd32(z + e.header_off) != SIG_LOCAL) return ZSB_E_ZIP;
        e.data_off = e.header_off + 30 + rd16(z + e.header_off + 26) + rd16(z + e.header_off + 28);
        if (!inside(e.data_off, e.comp_size, len)) return ZSB_E_ZIP;
        if (i < max_entries && out) out[i] = e;
        p += 46 + (uint64_t)nlen + xlen + clen;
    }
    if (count) *count = (size_t)total;
    return 0;
}

}  // namespace zsbzip

extern "C" int zsb_zip_scan(const void *zip, size_t len, zsb_zip_entry *entries, size_t max_entries, size_t *count) {
    if (!zip || (max_entries && !entries)) return ZSB_E_ARG;
    return zsbzip::scan((const uint8_t *)zip, len, entries, max_entries, count);
}

extern "C" int zsb_zip_members(const void *in_base, const zsb_zip_entry *e, size_t n, void *out_base,
                               const uint64_t *out_off, zsb_member *m) {
    if (n && (!e || !m || !out_off)) return ZSB_E_ARG;
    for (size_t i = 0; i < n; i++) {
        zsb_member &x = m[i];
        x.in = (const uint8_t *)in_base + e[i].data_off;
        x.in_len = (size_t)e[i].comp_size;
        x.out = (uint8_t *)out_base + out_off[i];
        x.out_cap = (size_t)e[i].uncomp_size;
        x.flags = ZSB_M_CRC32;
        x.window_bits = e[i].method == 9 ? -16 : -15;
        if (e[i].method == 0) x.flags |= ZSB_M_STORED;
        const bool ok = (e[i].method == 0 || e[i].method == 8 || e[i].method == 9) && !(e[i].gp_flags & 1) &&
                        e[i].comp_size <= 0xfffffff0ull && e[i].uncomp_size <= 0xfffffff0ull;
        if (!ok) {
            x.window_bits = 1000;       // Z_STREAM_ERROR
            x.flags = 0;
        }
    }
    return 0;
}

extern "C" int zsb_zip_inflate(zsb_ctx *ctx, const void *zip, const zsb_zip_entry *e, size_t n, void *out_base,
                               const uint64_t *out_off, zsb_result *r) {
    if (!ctx || (n && (!zip || !e || !r || !out_off))) return ZSB_E_ARG;
    std::vector<zsb_member> m(n);
    int rc = zsb_zip_members(zip, e, n, out_base, out_off, m.data());
    if (rc) return rc;
    if ((rc = zsb_inflate_batch(ctx, n, m.data(), r))) return rc;
    for (size_t i = 0; i < n; i++) {
        if (r[i].ret != 1) continue;                       // the decoder's own verdict stands
        if (r[i].total_in != e[i].comp_size || r[i].total_out != e[i].uncomp_size) {
            r[i].ret = -3;
            r[i].msg_id = 18;                              // "incorrect length check"
        } else if (r[i].check != e[i].crc32) {
            r[i].ret = -3;
            r[i].msg_id = 17;                              // "incorrect data check"
        }
    }
    return 0;
}

// gzip files of several members: rounds of one member per unfinished file (include/zsb.h)
extern "C" int zsb_gunzip_files(zsb_ctx *ctx, size_t n, const void *const *in, const size_t *in_len, void *const *out,
                                const size_t *out_cap, int window_bits, zsb_result *r, uint32_t *members) {
    if (!ctx || (n && (!in || !in_len || !out || !out_cap || !r))) return ZSB_E_ARG;
    std::vector<size_t> live(n), ipos(n, 0), opos(n, 0);
    for (size_t i = 0; i < n; i++) {
        live[i] = i;
        memset(&r[i], 0, sizeof r[i]);
        r[i].ret = -5;                                    // an empty buffer: Z_BUF_ERROR, as inflate() on no input
        if (members) members[i] = 0;
    }
    std::vector<zsb_member> m;
    std::vector<zsb_result> rr;
    while (!live.empty()) {
        m.resize(live.size());
        rr.resize(live.size());
        for (size_t k = 0; k < live.size(); k++) {
            const size_t i = live[k];
            m[k].in = (const uint8_t *)in[i] + ipos[i];
            m[k].in_len = in_len[i] - ipos[i];
            m[k].out = (uint8_t *)out[i] + opos[i];
            m[k].out_cap = out_cap[i] - opos[i];
            m[k].window_bits = window_bits;
            m[k].flags = 0;
        }
        const int rc = zsb_inflate_batch(ctx, live.size(), m.data(), rr.data());
        if (rc) return rc;
        std::vector<size_t> next;
        for (size_t k = 0; k < live.size(); k++) {
            const size_t i = live[k];
            const zsb_result &x = rr[k];
            r[i].ret = x.ret;
            r[i].msg_id = x.msg_id;
            r[i].check = x.check;
            r[i].data_type = x.data_type;
            r[i].total_in += x.total_in;
            r[i].total_out += x.total_out;
            ipos[i] += (size_t)x.total_in;
            opos[i] += (size_t)x.total_out;
            if (x.ret == 1) {
                if (members) members[i]++;
                if (ipos[i] < in_len[i]) next.push_back(i);       // another member follows
            }
        }
        live.swap(next);
    }
    return 0;
}

// gzip member header -> fields (include/zsb.h; the states HEAD..HCRC of the reference's inflate(), src/inflate.c:678-816)
namespace zsbgz {
static uint32_t crc32_bytes(const uint8_t *p, size_t n) {          // a header is tens of bytes: bit at a time
    uint32_t c = 0xffffffffu;
    for (size_t i = 0; i < n; i++) {
        c ^= p[i];
        for (int k = 0; k < 8; k++) c = (c >> 1) ^ (0xedb88320u & (0u - (c & 1u)));
    }
    return ~c;
}
}  // namespace zsbgz

extern "C" int zsb_gzip_header(const void *in, size_t len, zsb_gz_header *h) {
    if (!h || (len && !in)) return ZSB_E_ARG;
    memset(h, 0, sizeof *h);
    const uint8_t *z = (const uint8_t *)in;
    auto bad = [&](int msg) { h->msg_id = msg; return -3; };
    if (len < 2) return -5;
    if (z[0] != 0x1f || z[1] != 0x8b) return bad(zsb::M_HDRCHECK);
    if (len < 4) return -5;
    if (z[2] != 8) return bad(zsb::M_METHOD);
    const uint32_t flg = z[3];
    if (flg & 0xe0) return bad(zsb::M_GZFLAGS);
    if (len < 10) return -5;
    h->text = flg & 1;
    h->time = zsbzip::rd32(z + 4);
    h->xflags = z[8];
    h->os = z[9];
    uint64_t p = 10;
    if (flg & 4) {                                                  // FEXTRA: 2-byte length, then the field
        if (len - p < 2) return -5;
        const uint32_t xl = zsbzip::rd16(z + p);
        p += 2;
        if (len - p < xl) return -5;
        h->extra_len = xl;
        h->extra_off = xl ? (uint32_t)p : 0;
        p += xl;
    }
    for (int f = 0; f < 2; f++) {                                   // FNAME, FCOMMENT: NUL-terminated
        if (!(flg & (f ? 16 : 8))) continue;
        const void *e = memchr(z + p, 0, (size_t)(len - p));
        if (!e) return -5;
        const uint64_t n = (uint64_t)((const uint8_t *)e - (z + p));
        if (n > 0xffffffffull || p > 0xffffffffull) return ZSB_E_ARG;
        (f ? h->comment_len : h->name_len) = (uint32_t)n;
        (f ? h->comment_off : h->name_off) = n ? (uint32_t)p : 0;
        p += n + 1;
    }
    if (flg & 2) {                                                  // FHCRC: low 16 bits of the CRC-32 of all before it
        if (len - p < 2) return -5;
        if (zsbzip::rd16(z + p) != (zsbgz::crc32_bytes(z, (size_t)p) & 0xffffu)) return bad(zsb::M_HCRC);
        h->hcrc = 1;
        p += 2;
    }
    if (p > 0xffffffffull) return ZSB_E_ARG;
    h->header_len = (uint32_t)p;
    return 0;
}
