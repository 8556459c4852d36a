"""zsb_gzip_header() (gzip header field extraction) against the system zlib's
inflateGetHeader() — the same header walk as the reference's src/inflate.c:678-816 — over every combination of
FTEXT / FHCRC / FEXTRA / FNAME / FCOMMENT, damaged headers and every truncation point.  Host-only: no GPU."""
import ctypes
import ctypes.util
import itertools
import struct
import zlib

import pytest

import zsblib


class GzHeaderOut(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint32) for n in ("text", "time", "xflags", "os", "extra_off", "extra_len", "name_off", "name_len",
                                                "comment_off", "comment_len", "hcrc", "header_len")] + \
               [("msg_id", ctypes.c_int32), ("reserved", ctypes.c_uint32)]


class ZlibGzHeader(ctypes.Structure):               # zlib.h gz_header
    _fields_ = [("text", ctypes.c_int), ("time", ctypes.c_ulong), ("xflags", ctypes.c_int), ("os", ctypes.c_int),
                ("extra", ctypes.c_void_p), ("extra_len", ctypes.c_uint), ("extra_max", ctypes.c_uint),
                ("name", ctypes.c_void_p), ("name_max", ctypes.c_uint), ("comment", ctypes.c_void_p), ("comm_max", ctypes.c_uint),
                ("hcrc", ctypes.c_int), ("done", ctypes.c_int)]


def build(text=0, hcrc=0, extra=None, name=None, comment=None, mtime=0x12345678, xfl=2, os_=3, payload=b"gzip header fields"):
    flg = text | (hcrc << 1) | ((extra is not None) << 2) | ((name is not None) << 3) | ((comment is not None) << 4)
    h = bytes([0x1f, 0x8b, 8, flg]) + struct.pack("<IBB", mtime, xfl, os_)
    if extra is not None:
        h += struct.pack("<H", len(extra)) + extra
    if name is not None:
        h += name + b"\0"
    if comment is not None:
        h += comment + b"\0"
    if hcrc:
        h += struct.pack("<H", zlib.crc32(h) & 0xffff)
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    body = c.compress(payload) + c.flush()
    return h, h + body + struct.pack("<II", zlib.crc32(payload), len(payload))


def ours(lib, data):
    h = GzHeaderOut()
    buf = ctypes.create_string_buffer(data, max(len(data), 1))
    rc = lib.zsb_gzip_header(buf, ctypes.c_size_t(len(data)), ctypes.byref(h))
    return rc, h


def system_zlib(data):
    """(ret, msg, fields) of inflate() with inflateGetHeader() registered, from the system libz."""
    z = ctypes.CDLL(ctypes.util.find_library("z") or "libz.so.1")
    strm = zsblib.ZStreamStruct()
    ver = z.zlibVersion
    ver.restype = ctypes.c_char_p
    assert z.inflateInit2_(ctypes.byref(strm), 31, ver(), ctypes.sizeof(strm)) == 0
    head = ZlibGzHeader()
    extra, name, comment = (ctypes.create_string_buffer(70000) for _ in range(3))
    head.extra, head.extra_max = ctypes.cast(extra, ctypes.c_void_p), 70000
    head.name, head.name_max = ctypes.cast(name, ctypes.c_void_p), 70000
    head.comment, head.comm_max = ctypes.cast(comment, ctypes.c_void_p), 70000
    assert z.inflateGetHeader(ctypes.byref(strm), ctypes.byref(head)) == 0
    src = ctypes.create_string_buffer(data, max(len(data), 1))
    out = ctypes.create_string_buffer(1 << 16)
    strm.next_in, strm.avail_in = ctypes.cast(src, ctypes.c_void_p), len(data)
    strm.next_out, strm.avail_out = ctypes.cast(out, ctypes.c_void_p), 1 << 16
    ret = z.inflate(ctypes.byref(strm), 0)
    msg = ctypes.cast(strm.msg, ctypes.c_char_p).value if strm.msg else None
    f = dict(done=head.done, text=head.text, time=head.time, xflags=head.xflags, os=head.os, hcrc=head.hcrc,
             extra=extra.raw[:head.extra_len] if head.done == 1 and head.extra else b"",
             name=name.value if head.done == 1 else b"", comment=comment.value if head.done == 1 else b"")
    z.inflateEnd(ctypes.byref(strm))
    return ret, msg, f


@pytest.fixture(scope="module")
def lib():
    lib = ctypes.CDLL(zsblib.LIB)
    lib.zsb_gzip_header.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(GzHeaderOut)]
    lib.zsb_message.restype = ctypes.c_char_p
    return lib


def variants():
    for text, hcrc, ex, nm, cm in itertools.product((0, 1), (0, 1), (None, b"", b"\x41\x70\x04\x00abcd"), (None, b"", b"file name.txt"),
                                                    (None, b"", b"a comment \xff\xfe")):
        yield dict(text=text, hcrc=hcrc, extra=ex, name=nm, comment=cm)
    yield dict(extra=bytes(range(256)) * 255, name=b"n" * 3000, comment=b"c" * 5000, hcrc=1)


def test_fields_match_inflateGetHeader(lib):
    for v in variants():
        hdr, member = build(**v)
        rc, h = ours(lib, member)
        ret, msg, f = system_zlib(member)
        assert (rc, ret, f["done"]) == (0, 1, 1), (v, rc, ret, msg)
        assert h.header_len == len(hdr)
        assert (h.text, h.time, h.xflags, h.os, h.hcrc) == (f["text"], f["time"], f["xflags"], f["os"], f["hcrc"]), v
        assert member[h.extra_off:h.extra_off + h.extra_len] == f["extra"] == (v.get("extra") or b"")
        assert member[h.name_off:h.name_off + h.name_len] == f["name"] == (v.get("name") or b"")
        assert member[h.comment_off:h.comment_off + h.comment_len] == f["comment"] == (v.get("comment") or b"")
        # the data the decoder is handed starts where the header ends
        assert zlib.decompress(member[h.header_len:], -15) == b"gzip header fields"


def test_every_truncation_point(lib):
    hdr, member = build(text=1, hcrc=1, extra=b"\x01\x02\x03", name=b"name", comment=b"comment")
    for n in range(len(hdr)):
        rc, h = ours(lib, member[:n])
        ret, msg, f = system_zlib(member[:n])
        assert rc == -5 and ret in (0, -5) and f["done"] == 0, (n, rc, ret)
    assert ours(lib, member[:len(hdr)])[0] == 0


def test_damaged_headers_report_inflates_message(lib):
    hdr, member = build(hcrc=1, name=b"x")
    cases = []
    for pos, val in ((0, 0x1e), (1, 0x8a), (2, 7), (3, 0x2a), (3, 0x4a), (3, 0x8a), (len(hdr) - 1, hdr[-1] ^ 1), (5, hdr[5] ^ 0x10)):
        b = bytearray(member)
        b[pos] = val
        cases.append(bytes(b))
    for data in cases:
        rc, h = ours(lib, data)
        ret, msg, f = system_zlib(data)
        assert (rc, ret) == (-3, -3), (data[:12], rc, ret, msg)
        assert lib.zsb_message(h.msg_id) == msg, (data[:12], h.msg_id, msg)


def test_argument_errors(lib):
    assert lib.zsb_gzip_header(None, 10, ctypes.byref(GzHeaderOut())) == -100          # ZSB_E_ARG
    buf = ctypes.create_string_buffer(b"\x1f\x8b", 2)
    assert lib.zsb_gzip_header(buf, 2, None) == -100
