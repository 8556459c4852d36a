"""Shared parity cases: (name, data, wbits, out_cap) lists used by the oracle, emulation and GPU tests."""
import hashlib
import json
import os
import random
import struct
import zlib

import gen

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def golden_fixtures():
    meta = json.load(open(os.path.join(GOLD, "ref_data_sha256.json")))
    out = []
    for name, m in meta.items():
        if name.startswith("_"):
            continue
        data = open(os.path.join(GOLD, "ref-data", name + ".deflate64"), "rb").read()
        assert hashlib.sha256(data).hexdigest() == m["in_sha256"]
        out.append((name, data, m))
    return out


def probe_vectors():
    """Hand-made streams from SURVEY.md §8a with the results the compiled reference gave there."""
    v = []
    # fixed block: literal 'a', length code 285, distance code 0 (dist 1)
    def fixed285(extra16=None):
        bw = gen.BitWriter()
        bw.bits(1, 1); bw.bits(1, 2)
        gen.fixed_lit(bw, ord("a"))
        gen.fixed_lit(bw, 285)
        if extra16 is not None:
            bw.bits(extra16, 16)
        bw.code(0, 5)
        gen.fixed_lit(bw, 256)
        return bw.done()
    v.append(("fixed285_deflate", fixed285(), -15, 70000, (1, 0, 4, 259)))
    v.append(("fixed285_d64_x0000", fixed285(0x0000), -16, 70000, (1, 0, 6, 4)))
    v.append(("fixed285_d64_xffff", fixed285(0xffff), -16, 70000, (1, 0, 6, 65539)))
    v.append(("fixed285_d64_cap1000", fixed285(0xffff), -16, 1000, (-5, 0, 5, 1000)))
    # distance code 30 after one literal
    bw = gen.BitWriter()
    bw.bits(1, 1); bw.bits(1, 2)
    gen.fixed_lit(bw, ord("a")); gen.fixed_lit(bw, 257); bw.code(30, 5); bw.bits(0, 14)
    gen.fixed_lit(bw, 256)
    d30 = bw.done()
    v.append(("dist30_deflate", d30, -15, 1000, (-3, 15, 3, 1)))
    v.append(("dist30_d64", d30, -16, 1000, (-3, 16, 5, 1)))
    v.append(("btype3", b"\x07", -15, 10, (-3, 6, 1, 0)))
    v.append(("bad_nlen", b"\x01\x05\x00\x00\x00abcde", -15, 10, (-3, 7, 5, 0)))
    gz = gen.deflate(b"hello hello hello", 6, 31)
    v.append(("gzip_as_zlib", gz, 15, 100, (-3, 1, 2, 0)))
    bad = bytearray(gz); bad[-5] ^= 1
    v.append(("gzip_bad_crc", bytes(bad), 31, 100, (-3, 17, len(gz) - 4, 17)))
    bad = bytearray(gz); bad[-1] ^= 1
    v.append(("gzip_bad_isize", bytes(bad), 31, 100, (-3, 18, len(gz), 17)))
    st = gen.deflate(random.Random(5).randbytes(70000), 0, -15)
    v.append(("stored_70000_deflate", st, -15, 80000, (1, 0, 70010, 70000)))
    v.append(("stored_70000_d64", st, -16, 80000, (1, 0, 70010, 70000)))
    # code sets at the reference's table-space limits (SURVEY.md §8 a6'(ii); tests/tablespace.py): a literal/length
    # set needing exactly 852 entries and a 32-symbol distance set needing 594 are rejected by Deflate64 mode only
    # (src/inftrees.c:211-212,226-228,295-297); results as the compiled reference gave them
    import tablespace
    want = {("286sym_root9_852", -15): (1, 0, 37, 0), ("286sym_root9_852", -16): (-3, 12, 35, 0),
            ("286sym_root9_850", -15): (1, 0, 39, 0), ("286sym_root9_850", -16): (1, 0, 39, 0),
            ("30sym_root6_592", -15): (1, 0, 23, 0), ("30sym_root6_592", -16): (1, 0, 23, 0),
            ("32sym_root6_592", -15): (-3, 8, 3, 0), ("32sym_root6_592", -16): (1, 0, 23, 0),
            ("32sym_root6_594", -15): (-3, 8, 3, 0), ("32sym_root6_594", -16): (-3, 13, 23, 0)}
    for name, data, wb, cap in tablespace.probe_cases():
        key = name[len("tabspace_"):name.rindex("_wb")]
        v.append((name, data, wb, cap, want[(key, wb)]))
    return v


def gzip_with_fields(payload, extra=None, name=None, comment=None, hcrc=False, seed=0):
    flg = (4 if extra is not None else 0) | (8 if name is not None else 0) | (16 if comment is not None else 0) | (2 if hcrc else 0)
    h = bytearray(b"\x1f\x8b\x08" + bytes([flg]) + struct.pack("<I", 12345) + b"\x00\x03")
    if extra is not None:
        h += struct.pack("<H", len(extra)) + extra
    if name is not None:
        h += name + b"\0"
    if comment is not None:
        h += comment + b"\0"
    if hcrc:
        h += struct.pack("<H", zlib.crc32(bytes(h)) & 0xffff)
    body = gen.deflate(payload, 6, -15)
    return bytes(h) + body + struct.pack("<II", zlib.crc32(payload) & 0xffffffff, len(payload) & 0xffffffff)


def valid_streams(scale=1):
    """(name, data, wbits, plain) across wrappers, block types and sizes."""
    out = []
    rnd = random.Random(99)
    sizes = [0, 1, 2, 17, 300, 4096, 65536, 200000]
    for i, sz in enumerate(sizes):
        plain = gen.mixed_plain(sz, 100 + i)
        for wb in (-15, 15, 31):
            for level, strat in ((6, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_DEFAULT_STRATEGY), (9, zlib.Z_FIXED),
                                 (0, zlib.Z_DEFAULT_STRATEGY), (6, zlib.Z_HUFFMAN_ONLY), (6, zlib.Z_RLE)):
                out.append(("mixed%d_wb%d_l%d_s%d" % (sz, wb, level, strat), gen.deflate(plain, level, wb, strat), wb, plain))
    for i in range(6 * scale):
        plain = gen.word_soup(65536, 1000 + i)
        out.append(("soup%d" % i, gen.deflate(plain, 6, -15), -15, plain))
        out.append(("soup%d_d64mode" % i, gen.deflate(plain, 6, -15), -16, plain))
    for wb in (-9, -12, 9, 12, 25, 28, 47, 44):
        plain = gen.mixed_plain(5000, wb + 500)
        eff = wb if wb < 0 else (wb & 15)
        enc_wb = -abs(eff) if wb < 0 else (eff + (16 if wb in (25, 28) else 0))
        out.append(("smallwin_wb%d" % wb, gen.deflate(plain, 6, enc_wb), wb, plain))
    p = gen.mixed_plain(3000, 77)
    out.append(("gz_extra", gzip_with_fields(p, extra=b"xy" * 10), 31, p))
    out.append(("gz_name_comment_hcrc", gzip_with_fields(p, name=b"file.txt", comment=b"hi there", hcrc=True), 31, p))
    out.append(("gz_all_auto", gzip_with_fields(p, extra=b"", name=b"", comment=b"", hcrc=True), 47, p))
    out.append(("zlib_auto", gen.deflate(p, 6, 15), 47, p))
    out.append(("zlib_wb0", gen.deflate(p, 6, 12), 0, p))
    return out


def damaged_streams(count=300):
    base = valid_streams()
    rnd = random.Random(4242)
    out = []
    for k in range(count):
        name, data, wb, plain = base[rnd.randrange(len(base))]
        out.append(("dmg%d_%s" % (k, name), gen.corrupt(data, 9000 + k), wb, len(plain) + 100))
    # systematic truncation and capacity limits on a few streams
    for name, data, wb, plain in base[::17]:
        for cut in sorted(set([0, 1, 2, 3, 5, 9, len(data) // 2, max(len(data) - 9, 0), max(len(data) - 5, 0),
                               max(len(data) - 4, 0), max(len(data) - 1, 0)])):
            out.append(("trunc%d_%s" % (cut, name), data[:cut], wb, len(plain) + 10))
        for cap in sorted(set([0, 1, 2, len(plain) // 2, max(len(plain) - 1, 0), len(plain)])):
            out.append(("cap%d_%s" % (cap, name), data, wb, cap))
    # damage concentrated on dynamic block headers (code-length code, repeats, code sets), both modes
    soup = gen.deflate(gen.word_soup(3000, 31337), 6, -15)
    fx = open(os.path.join(GOLD, "ref-data", "10k_lines.deflate64"), "rb").read()
    for k in range(count):
        for src, wbs in ((soup, (-15, -16)), (fx, (-16,))):
            b = bytearray(src)
            for _ in range(rnd.randrange(1, 3)):
                b[rnd.randrange(min(70, len(b)))] ^= 1 << rnd.randrange(8)
            for wb in wbs:
                out.append(("hdrfuzz%d_wb%d" % (k, wb), bytes(b), wb, 200000))
    gzh = gzip_with_fields(gen.mixed_plain(500, 3), extra=b"abcdef", name=b"n", comment=b"c", hcrc=True)
    for k in range(min(count, 200)):
        b = bytearray(gzh)
        b[rnd.randrange(30)] ^= 1 << rnd.randrange(8)
        out.append(("gzhdrfuzz%d" % k, bytes(b), 31 if k % 2 else 47, 1000))
        out.append(("gzhdrtrunc%d" % k, gzh[:k % 40], 31, 1000))
    # headers that ask for a dictionary / bad wrappers / invalid windowBits
    z = bytearray(gen.deflate(b"abc", 6, 15)); z[1] |= 0x20
    z[1] = (z[1] & 0xe0) | (31 - ((z[0] << 8 | (z[1] & 0xe0)) % 31)) % 31 if ((z[0] << 8 | (z[1] & 0xe0)) % 31) else z[1] & 0xe0
    out.append(("fdict", bytes(z) + b"\x01\x02\x03\x04rest", 15, 100))
    out.append(("fdict_short", bytes(z) + b"\x01\x02", 15, 100))
    for wb in (-17, -7, 7, 16 + 7, 48, 100, -16, -15, 0, 16):
        out.append(("wbits%d" % wb, gen.deflate(b"hello world", 6, -15), wb, 100))
    return out


def deflate64_streams(big=False):
    """(name, data, wbits, plain): Deflate64 streams from the test-only encoder (tests/d64enc.py) that use
    what the reference's fixtures never do: length code 285 with its 16 extra bits, distance codes 30/31
    (in dynamic and fixed blocks), matches longer than 258, distances beyond 32 KiB, all block kinds.
    Callers first check with the reference build that each one decodes back to `plain`."""
    import d64enc
    rnd = random.Random(64)
    blk = bytes(rnd.getrandbits(8) for _ in range(50000))
    plains = {
        "soup": gen.word_soup(90000, 3),
        "far": blk + gen.word_soup(12000, 4) + blk + blk[:40000],
        "runs": b"ab" * 40000 + bytes(70000) + b"xyz" * 30000 + bytes(range(256)) * 300,
        "mixed": gen.mixed_plain(150000, 5),
        "tiny": b"q",
        "empty": b"",
    }
    if big:                       # a C3-like member: 4 MiB, long matches at far distances
        seg = gen.word_soup(60000, 9)
        plains["c3_4MiB"] = (seg + blk[:5000]) * 64 + bytes(4 * 1024 * 1024 - 65000 * 64)
    out = []
    for name, plain in plains.items():
        for tag, kinds, bt in (("dyn", ("dynamic",), 4000), ("fix", ("fixed",), 100000),
                               ("mix", ("dynamic", "stored", "fixed"), 1500)):
            if name == "c3_4MiB" and tag != "dyn":
                continue
            out.append(("d64_%s_%s" % (name, tag), d64enc.compress(plain, True, bt, kinds), -16, plain))
        out.append(("d64_%s_f285" % name, d64enc.compress(plain, True, 5000, ("dynamic",), force285=True), -16, plain))
        # the same tokens restricted to deflate's limits, decoded in both modes
        comp = d64enc.compress(plain, False, 5000, ("dynamic", "fixed"))
        out.append(("d32_%s" % name, comp, -15, plain))
    return out


def deflate64_damaged(count=200):
    base = [c for c in deflate64_streams() if len(c[1]) > 8]
    rnd = random.Random(6464)
    out = []
    for k in range(count):
        name, data, wb, plain = base[rnd.randrange(len(base))]
        out.append(("dmg%d_%s" % (k, name), gen.corrupt(data, 700 + k), wb, len(plain) + 100))
    for name, data, wb, plain in base[::3]:
        for cut in (1, len(data) // 3, len(data) - 3, len(data) - 1):
            out.append(("trunc%d_%s" % (cut, name), data[:cut], wb, len(plain) + 10))
        for cap in (0, 1, 258, 259, len(plain) // 2, len(plain) - 1):
            out.append(("cap%d_%s" % (cap, name), data, wb, max(cap, 0)))
        out.append(("asdeflate_%s" % name, data, -15 if wb == -16 else -16, len(plain) + 10))
    return out
