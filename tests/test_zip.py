"""ZIP container walker (include/zsb.h zsb_zip_*): the host walker against Python's
zipfile on CPU, and whole archives (methods 0 / 8 / 9, ZIP64, damaged entries) decoded and CRC-checked on
the GPU against the original files."""
import ctypes
import io
import random
import zipfile
import zlib

import pytest

import gen
import zipgen
import zsblib

LIB = zsblib.load_product()


def _files(seed=7, count=40):
    rnd = random.Random(seed)
    out = []
    for i in range(count):
        n = rnd.choice((0, 1, 50, 3000, 70000, 300000)) + rnd.randrange(100)
        kind = i % 3
        plain = gen.word_soup(n, 900 + i) if kind == 0 else gen.mixed_plain(n, 900 + i) if kind == 1 else rnd.randbytes(n)
        out.append(("dir%d/file_%03d.bin" % (i % 4, i), plain))
    return out


def test_scan_matches_zipfile():
    files = _files()
    bio = io.BytesIO()
    with zipfile.ZipFile(bio, "w") as zf:
        for i, (name, plain) in enumerate(files):
            zf.writestr(zipfile.ZipInfo(name), plain, zipfile.ZIP_STORED if i % 5 == 0 else zipfile.ZIP_DEFLATED,
                        compresslevel=(1, 6, 9)[i % 3])
        zf.comment = b"archive comment"
    data = bio.getvalue()
    rc, ents = zsblib.zip_scan(LIB, data)
    assert rc == 0 and len(ents) == len(files)
    with zipfile.ZipFile(io.BytesIO(data)) as zf:
        for e, zi in zip(ents, zf.infolist()):
            assert (e.header_off, e.comp_size, e.uncomp_size, e.crc32, e.method) == (
                zi.header_offset, zi.compress_size, zi.file_size, zi.CRC, zi.compress_type)
            assert data[e.name_off:e.name_off + e.name_len].decode() == zi.filename
            raw = data[e.data_off:e.data_off + e.comp_size]
            plain = raw if e.method == 0 else zlib.decompress(raw, -15)
            assert plain == zf.read(zi)


def test_scan_zip64_and_local_extra():
    ents_in = [dict(name="a.txt", method=8, data=gen.deflate(b"hello " * 100, 6, -15), plain=b"hello " * 100),
               dict(name="b.bin", method=0, data=b"\x00\x01\x02", plain=b"\x00\x01\x02")]
    for z64 in (False, True):
        data = zipgen.build(ents_in, zip64=z64, comment=b"c" * 33, local_extra=b"\x99\x99\x04\x00abcd")
        rc, ents = zsblib.zip_scan(LIB, data)
        assert rc == 0 and len(ents) == 2
        for e, src in zip(ents, ents_in):
            assert data[e.data_off:e.data_off + e.comp_size] == src["data"]
            assert (e.uncomp_size, e.crc32, e.method) == (len(src["plain"]), zlib.crc32(src["plain"]), src["method"])
        if not z64:                      # zipfile agrees on the hand-written archive
            with zipfile.ZipFile(io.BytesIO(data)) as zf:
                assert zf.read("a.txt") == b"hello " * 100


def test_scan_rejects_malformed_archives():
    ents_in = [dict(name="a", method=8, data=gen.deflate(b"x" * 1000, 6, -15), plain=b"x" * 1000)]
    good = zipgen.build(ents_in)
    assert zsblib.zip_scan(LIB, good)[0] == 0
    ZIPERR = -103
    assert zsblib.zip_scan(LIB, b"")[0] == ZIPERR
    assert zsblib.zip_scan(LIB, good[:-1])[0] == ZIPERR                      # comment length no longer fits
    assert zsblib.zip_scan(LIB, b"PK" + good[2:-22] + bytes(22))[0] == ZIPERR   # no end record
    rnd = random.Random(3)
    for _ in range(300):                 # random damage: an error code or a directory that stays inside the archive
        b = bytearray(good)
        for _ in range(rnd.randrange(1, 4)):
            b[rnd.randrange(len(b))] = rnd.randrange(256)
        rc, ents = zsblib.zip_scan(LIB, bytes(b))
        assert rc in (0, ZIPERR)
        for e in ents:
            assert e.data_off + e.comp_size <= len(b)


def _archive_with_all_methods():
    import d64enc
    files = _files(seed=11, count=30)
    ents = []
    for i, (name, plain) in enumerate(files):
        k = i % 4
        if k == 0:
            ents.append(dict(name=name, method=0, data=plain, plain=plain))
        elif k == 3:
            ents.append(dict(name=name, method=9, data=d64enc.compress(plain, True, 5000, ("dynamic", "fixed")), plain=plain))
        else:
            ents.append(dict(name=name, method=8, data=gen.deflate(plain, (1, 6, 9)[i % 3], -15), plain=plain,
                             flags=8 if i % 7 == 1 else 0))
    return ents


@pytest.mark.gpu
@pytest.mark.parametrize("z64", [False, True], ids=["zip", "zip64"])
def test_gpu_zip_inflate_all_methods(z64):
    ents_in = _archive_with_all_methods()
    # damage: a flipped data byte in a stored entry (CRC), a wrong directory size, an encrypted flag, an unknown method
    ents_in[4] = dict(ents_in[4], data=bytes([ents_in[4]["data"][0] ^ 1]) + ents_in[4]["data"][1:])
    ents_in[5] = dict(ents_in[5], crc=(zlib.crc32(ents_in[5]["plain"]) ^ 0x10) & 0xffffffff)
    ents_in[6] = dict(ents_in[6], flags=1)
    ents_in[9] = dict(ents_in[9], method=12)
    ents_in[10] = dict(ents_in[10], plain=ents_in[10]["plain"] + b"!")          # directory says one byte more
    data = zipgen.build(ents_in, zip64=z64)
    rc, ents = zsblib.zip_scan(LIB, data)
    assert rc == 0 and len(ents) == len(ents_in)
    n = len(ents)
    offs, total = [], 0
    for e in ents:
        offs.append(total)
        total += e.uncomp_size + 3                    # ragged, unaligned outputs
    zbuf = ctypes.create_string_buffer(data, len(data))
    obuf = ctypes.create_string_buffer(max(total, 1))
    earr = (zsblib.ZipEntry * n)(*ents)
    oarr = (ctypes.c_uint64 * n)(*offs)
    res = (zsblib.Result * n)()
    eng = zsblib.Zsb(devices=[0])
    try:
        for path in (0, 2):
            eng.set_option("path", path)
            ctypes.memset(obuf, 0, max(total, 1))
            rc = LIB.zsb_zip_inflate(eng.ctx, zbuf, earr, n, obuf, oarr, res)
            assert rc == 0, eng.error()
            for i, (e, src) in enumerate(zip(ents, ents_in)):
                got = obuf.raw[offs[i]:offs[i] + res[i].total_out]
                if i == 4 or i == 5:
                    assert (res[i].ret, LIB.zsb_message(res[i].msg_id)) == (-3, b"incorrect data check"), (i, res[i])
                elif i in (6, 9):
                    assert res[i].ret == -2, (i, res[i])
                elif i == 10:
                    assert res[i].ret in (-3, -5) and res[i].ret != 1, (i, res[i])
                else:
                    assert res[i].ret == 1 and got == src["plain"], (path, i, src["name"], res[i])
                    assert res[i].check == zlib.crc32(src["plain"]) & 0xffffffff
    finally:
        eng.close()


@pytest.mark.gpu
def test_gpu_gunzip_files_of_several_members():
    """concatenated gzip members; the checker decodes member by member with the oracle
    (one inflateInit2/inflate/inflateEnd per member, advancing by total_in, as the reference would be driven)."""
    import orclib
    chk = orclib.ref() or orclib.orc()
    rnd = random.Random(21)
    files = []
    for f in range(24):
        parts = []
        for k in range(rnd.randrange(1, 6)):
            plain = gen.word_soup(rnd.randrange(0, 90000), 300 + 10 * f + k)
            parts.append(gen.deflate(plain, rnd.choice((1, 6, 9)), 31))
        data = b"".join(parts)
        if f % 6 == 4:
            data = data[:-5]                         # the last member's trailer is cut
        if f % 6 == 5:
            data += b"\x00\x00\x00\x00"              # something that is not a member follows
        files.append(data)
    files.append(b"")
    n = len(files)
    caps = []
    want = []
    for data in files:                               # member by member with the oracle
        pos, out, last, members = 0, b"", None, 0
        cap = 600000
        while True:
            r, o = chk.oneshot(data[pos:], cap - len(out), 31)
            last = r
            out += o
            pos += r.total_in
            if r.ret != 1:
                break
            members += 1
            if pos >= len(data):
                break
        want.append((last.ret, last.msg_id, pos, out, members))
        caps.append(cap)
    ibufs = [ctypes.create_string_buffer(d, max(len(d), 1)) for d in files]
    obufs = [ctypes.create_string_buffer(c) for c in caps]
    iarr = (ctypes.c_void_p * n)(*[ctypes.addressof(b) for b in ibufs])
    oarr = (ctypes.c_void_p * n)(*[ctypes.addressof(b) for b in obufs])
    ilen = (ctypes.c_size_t * n)(*[len(d) for d in files])
    ocap = (ctypes.c_size_t * n)(*caps)
    res = (zsblib.Result * n)()
    cnt = (ctypes.c_uint32 * n)()
    LIB.zsb_gunzip_files.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    eng = zsblib.Zsb(devices=[0])
    try:
        rc = LIB.zsb_gunzip_files(eng.ctx, n, iarr, ilen, oarr, ocap, 31, res, cnt)
        assert rc == 0, eng.error()
    finally:
        eng.close()
    for i, (ret, msg, tin, out, members) in enumerate(want):
        got = obufs[i].raw[:res[i].total_out]
        assert (res[i].ret, res[i].msg_id, res[i].total_in, cnt[i]) == (ret, msg, tin, members), (i, res[i], want[i][:3])
        assert got == out, i
    import gzip
    assert obufs[0].raw[:res[0].total_out] == gzip.decompress(files[0])
