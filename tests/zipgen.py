"""TEST TOOL: writes ZIP archives byte by byte (APPNOTE layout), so that tests control what Python's zipfile
cannot produce: method 9 (Deflate64) entries, data descriptors, ZIP64 records, extra fields that differ
between the local header and the directory, damaged entries."""
import struct
import zlib


def build(entries, zip64=False, comment=b"", local_extra=b"", central_extra=b""):
    """entries: list of dict(name, method, data=compressed bytes, plain=original bytes, [crc], [flags]).
    Returns archive bytes.  zip64=True writes sizes / offsets through ZIP64 extra fields and records."""
    out = bytearray()
    cd = bytearray()
    for e in entries:
        name = e["name"].encode()
        crc = e.get("crc", zlib.crc32(e["plain"]) & 0xffffffff)
        flags = e.get("flags", 0)
        csize, usize, off = len(e["data"]), len(e["plain"]), len(out)
        lx = local_extra
        cx = central_extra
        if zip64:
            lx = struct.pack("<HHQQ", 1, 16, usize, csize) + lx
            cx = struct.pack("<HHQQQ", 1, 24, usize, csize, off) + cx
        f32 = (lambda v: 0xffffffff) if zip64 else (lambda v: v)
        out += struct.pack("<IHHHHHIIIHH", 0x04034b50, 45 if zip64 else 20, flags, e["method"], 0, 0x21,
                           0 if flags & 8 else crc, 0 if flags & 8 else f32(csize), 0 if flags & 8 else f32(usize),
                           len(name), len(lx))
        out += name + lx + e["data"]
        if flags & 8:                       # data descriptor after the data
            out += struct.pack("<IIII", 0x08074b50, crc, csize & 0xffffffff, usize & 0xffffffff)
        cd += struct.pack("<IHHHHHHIIIHHHHHII", 0x02014b50, 45, 45 if zip64 else 20, flags, e["method"], 0, 0x21, crc,
                          f32(csize), f32(usize), len(name), len(cx), 0, 0, 0, 0, f32(off))
        cd += name + cx
    cd_off = len(out)
    out += cd
    n = len(entries)
    if zip64:
        e64 = len(out)
        out += struct.pack("<IQHHIIQQQQ", 0x06064b50, 44, 45, 45, 0, 0, n, n, len(cd), cd_off)
        out += struct.pack("<IIQI", 0x07064b50, 0, e64, 1)
        out += struct.pack("<IHHHHIIH", 0x06054b50, 0, 0, 0xffff, 0xffff, 0xffffffff, 0xffffffff, len(comment))
    else:
        out += struct.pack("<IHHHHIIH", 0x06054b50, 0, 0, n, n, len(cd), cd_off, len(comment))
    out += comment
    return bytes(out)
