"""ctypes bindings for the two CPU checkers (oracle/oracle_api.h).  TEST INFRASTRUCTURE ONLY.

`orc` = this repo's plain-C restatement (oracle/liborc.so, always buildable with gcc);
`ref` = the unmodified reference compiled in place (oracle/_ref/libzsb_ref.so; built here where
/root/reference exists, travels prebuilt to the GPU box).
"""
import ctypes
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

Z_OK, Z_STREAM_END, Z_NEED_DICT = 0, 1, 2
Z_STREAM_ERROR, Z_DATA_ERROR, Z_MEM_ERROR, Z_BUF_ERROR, Z_VERSION_ERROR = -2, -3, -4, -5, -6
Z_NO_FLUSH, Z_SYNC_FLUSH, Z_FINISH, Z_BLOCK, Z_TREES = 0, 2, 4, 5, 6


class Result(ctypes.Structure):
    _fields_ = [("ret", ctypes.c_int32), ("msg_id", ctypes.c_int32),
                ("total_in", ctypes.c_uint64), ("total_out", ctypes.c_uint64),
                ("adler", ctypes.c_uint32), ("data_type", ctypes.c_int32)]

    def astuple(self):
        return (self.ret, self.msg_id, self.total_in, self.total_out, self.adler, self.data_type)

    def __repr__(self):
        return "Result(ret=%d msg=%d in=%d out=%d adler=%08x dt=%d)" % self.astuple()


class Checker:
    """One of the two CPU checkers behind the common oracle_api.h surface."""

    def __init__(self, path, prefix):
        self.lib = ctypes.CDLL(path)
        self.prefix = prefix
        g = lambda n: getattr(self.lib, prefix + "_" + n)
        self._oneshot = g("oneshot")
        self._oneshot.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                  ctypes.c_int, ctypes.POINTER(Result)]
        self._oneshot.restype = ctypes.c_int
        self._open = g("open")
        self._open.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        self._open.restype = ctypes.c_void_p
        self._step = g("step")
        self._step.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p,
                               ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(ctypes.c_size_t),
                               ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(Result)]
        self._step.restype = ctypes.c_int
        self._close = g("close")
        self._close.argtypes = [ctypes.c_void_p]
        self._bench = g("bench")
        self._bench.argtypes = [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        self._bench.restype = ctypes.c_double

    def codes_used(self, data, wbits):
        """(reference build only) inflateCodesUsed() after a one-shot decode -> (entries, Result)."""
        f = self.lib.ref_codes_used
        f.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(Result)]
        f.restype = ctypes.c_long
        r = Result()
        return f(data, len(data), wbits, ctypes.byref(r)), r

    def oneshot(self, data, out_cap, wbits):
        """inflateInit2(wbits) + inflate(Z_FINISH) + inflateEnd -> (Result, output bytes)."""
        buf = ctypes.create_string_buffer(max(out_cap, 1))
        r = Result()
        self._oneshot(data, len(data), buf, out_cap, wbits, ctypes.byref(r))
        return r, buf.raw[:min(r.total_out, out_cap)]

    def init_ret(self, wbits):
        rc = ctypes.c_int()
        h = self._open(wbits, ctypes.byref(rc))
        if h:
            self._close(h)
        return rc.value

    def stream(self, data, wbits, in_chunks, out_cap, finish=True, max_calls=1000000, flush=Z_NO_FLUSH):
        """Feed `data` in the given input chunk sizes (cycled) with an `out_cap`-byte rolling output
        buffer, `flush` (Z_NO_FLUSH, or Z_BLOCK / Z_TREES) while input remains, then Z_FINISH; returns the per-call trace
        [(ret,msg,total_in,total_out,adler,data_type,consumed,produced)...] and the output."""
        rc = ctypes.c_int()
        h = self._open(wbits, ctypes.byref(rc))
        if not h:
            return [("init", rc.value)], b""
        trace, out = [], bytearray()
        buf = ctypes.create_string_buffer(max(out_cap, 1))
        pos, k = 0, 0
        pending = b""
        c, p, r = ctypes.c_size_t(), ctypes.c_size_t(), Result()
        try:
            for _ in range(max_calls):
                if not pending and pos < len(data):
                    n = in_chunks[k % len(in_chunks)]
                    k += 1
                    pending = data[pos:pos + n]
                    pos += len(pending)
                fl = Z_FINISH if (finish and pos >= len(data)) else flush
                ret = self._step(h, pending, len(pending), buf, out_cap, fl,
                                 ctypes.byref(c), ctypes.byref(p), ctypes.byref(r))
                out += buf.raw[:p.value]
                trace.append(r.astuple() + (c.value, p.value))
                pending = pending[c.value:]
                if ret not in (Z_OK, Z_BUF_ERROR):
                    break
                if ret == Z_BUF_ERROR and c.value == 0 and p.value == 0 and pos >= len(data):
                    break
        finally:
            self._close(h)
        return trace, bytes(out)

    def bench(self, members, wbits, out_caps, threads, reps=1):
        """Time one pass over `members` (list of bytes) on `threads` host threads."""
        n = len(members)
        keep = [ctypes.create_string_buffer(m, len(m)) for m in members]
        ptrs = (ctypes.c_void_p * n)(*[ctypes.addressof(b) for b in keep])
        lens = (ctypes.c_size_t * n)(*[len(m) for m in members])
        caps = (ctypes.c_size_t * n)(*out_caps)
        wb = (ctypes.c_int * n)(*wbits)
        so, bad = ctypes.c_uint64(), ctypes.c_uint64()
        secs = self._bench(n, ptrs, lens, caps, wb, threads, reps, ctypes.byref(so), ctypes.byref(bad))
        return secs, so.value, bad.value


def build_oracles():
    """Compile the restatement (always) and the reference build (only where /root/reference is)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "orc"], check=True)
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True)


_cache = {}


def orc():
    if "orc" not in _cache:
        path = os.path.join(ORACLE_DIR, "liborc.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-s", "-C", ORACLE_DIR, "orc"], check=True)
        _cache["orc"] = Checker(path, "orc")
    return _cache["orc"]


def ref():
    """The compiled reference, or None when oracle/_ref was neither built nor shipped."""
    if "ref" not in _cache:
        path = os.path.join(ORACLE_DIR, "_ref", "libzsb_ref.so")
        if not os.path.exists(path) and os.path.isdir("/root/reference/src"):
            subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True)
        _cache["ref"] = Checker(path, "ref") if os.path.exists(path) else None
    return _cache["ref"]
