"""TEST TOOL: zlib's z_stream API (inflate side, auxiliary entry points included) driven through ctypes on ANY
library that exports it — libzsb.so (the product) or oracle/_ref/libzsb_ref.so (the unmodified reference) — so
that one script runs against both and the observations can be compared field by field."""
import ctypes
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class ZStream(ctypes.Structure):
    _fields_ = [("next_in", ctypes.c_void_p), ("avail_in", ctypes.c_uint), ("total_in", ctypes.c_ulong),
                ("next_out", ctypes.c_void_p), ("avail_out", ctypes.c_uint), ("total_out", ctypes.c_ulong),
                ("msg", ctypes.c_char_p), ("state", ctypes.c_void_p), ("zalloc", ctypes.c_void_p),
                ("zfree", ctypes.c_void_p), ("opaque", ctypes.c_void_p), ("data_type", ctypes.c_int),
                ("adler", ctypes.c_ulong), ("reserved", ctypes.c_ulong)]


class GzHeader(ctypes.Structure):
    _fields_ = [("text", ctypes.c_int), ("time", ctypes.c_ulong), ("xflags", ctypes.c_int), ("os", ctypes.c_int),
                ("extra", ctypes.c_void_p), ("extra_len", ctypes.c_uint), ("extra_max", ctypes.c_uint),
                ("name", ctypes.c_void_p), ("name_max", ctypes.c_uint), ("comment", ctypes.c_void_p),
                ("comm_max", ctypes.c_uint), ("hcrc", ctypes.c_int), ("done", ctypes.c_int)]


def load(which):
    path = {"zsb": os.path.join(ROOT, "zlib-streams_b200", "libzsb.so"),
            "ref": os.path.join(ROOT, "oracle", "_ref", "libzsb_ref.so")}[which]
    if not os.path.exists(path):
        return None
    lib = ctypes.CDLL(path)
    P = ctypes.POINTER(ZStream)
    lib.inflateInit2_.argtypes = [P, ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
    lib.inflate.argtypes = [P, ctypes.c_int]
    for n in ("inflateEnd", "inflateReset", "inflateResetKeep", "inflateSync", "inflateSyncPoint"):
        getattr(lib, n).argtypes = [P]
    lib.inflateReset2.argtypes = [P, ctypes.c_int]
    lib.inflatePrime.argtypes = [P, ctypes.c_int, ctypes.c_int]
    lib.inflateSetDictionary.argtypes = [P, ctypes.c_char_p, ctypes.c_uint]
    lib.inflateGetDictionary.argtypes = [P, ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint)]
    lib.inflateGetHeader.argtypes = [P, ctypes.POINTER(GzHeader)]
    lib.inflateCopy.argtypes = [P, P]
    lib.inflateUndermine.argtypes = [P, ctypes.c_int]
    lib.inflateValidate.argtypes = [P, ctypes.c_int]
    lib.inflateMark.argtypes = [P]
    lib.inflateMark.restype = ctypes.c_long
    lib.inflateCodesUsed.argtypes = [P]
    lib.inflateCodesUsed.restype = ctypes.c_ulong
    return lib


class Z:
    """One z_stream of `lib`; every method returns plain Python values."""

    def __init__(self, lib, wbits):
        self.lib = lib
        self.s = ZStream()
        self.init = lib.inflateInit2_(ctypes.byref(self.s), wbits, b"1.3", ctypes.sizeof(ZStream))
        self.keep = []

    def call(self, data, out_cap, flush=0):
        """one inflate(): -> (ret, consumed, output bytes)"""
        ibuf = ctypes.create_string_buffer(data, max(len(data), 1))
        obuf = ctypes.create_string_buffer(max(out_cap, 1))
        s = self.s
        s.next_in, s.avail_in = ctypes.addressof(ibuf), len(data)
        s.next_out, s.avail_out = ctypes.addressof(obuf), out_cap
        ret = self.lib.inflate(ctypes.byref(s), flush)
        return ret, len(data) - s.avail_in, obuf.raw[:out_cap - s.avail_out]

    def observe(self):
        s = self.s
        p = ctypes.byref(s)
        return {"total_in": s.total_in, "total_out": s.total_out, "adler": s.adler & 0xffffffff, "data_type": s.data_type,
                "msg": s.msg, "mark": self.lib.inflateMark(p), "codes_used": self.lib.inflateCodesUsed(p),
                "sync_point": self.lib.inflateSyncPoint(p)}

    def drive(self, data, chunks, out_cap, flush=0, finish=True, max_calls=100000, each=None):
        """Feed data in chunk sizes (cycled); returns (trace of (ret, consumed, produced, observation), output)."""
        pos, k, pending, out, trace = 0, 0, b"", bytearray(), []
        for _ in range(max_calls):
            if not pending and pos < len(data):
                n = chunks[k % len(chunks)]
                k += 1
                pending = data[pos:pos + n]
                pos += len(pending)
            fl = 4 if (finish and pos >= len(data)) else flush
            ret, c, o = self.call(pending, out_cap, fl)
            out += o
            pending = pending[c:]
            trace.append((ret, c, len(o), tuple(sorted(self.observe().items()))))
            if each:
                each(self, trace)
            if ret not in (0, -5):
                break
            if ret == -5 and c == 0 and not o and pos >= len(data):
                break
        return trace, bytes(out)

    def end(self):
        return self.lib.inflateEnd(ctypes.byref(self.s))
