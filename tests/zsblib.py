"""ctypes bindings used by the tests and bench.py.

`Zsb`  = the product: zlib-streams_b200/libzsb.so (CUDA, C ABI of include/zsb.h).
`emul` = TEST TOOL: the decoder's device source compiled for the host with a group width of one lane
         (tests/emul/), so the CPU-only test tier can exercise the decoder's control logic.
"""
import ctypes
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "zlib-streams_b200")
LIB = os.environ.get("ZSB_LIB") or os.path.join(PKG, "libzsb.so")      # ZSB_LIB: a variant build (experiments)
EMUL_DIR = os.path.join(ROOT, "tests", "emul")

API_SYMBOLS = ["zsb_profile_read_seg", "zsb_deflate_batch", "zsb_deflate_bound", "zsb_deflate_batch_static", "zsb_deflate_static_limits", "zsb_profile_read_check", "zsb_create", "zsb_destroy", "zsb_create_error", "zsb_last_error", "zsb_num_devices",
               "zsb_inflate_batch", "zsb_inflate_batch_device", "zsb_set_option", "zsb_get_option",
               "zsb_last_launches", "zsb_plan_shards", "zsb_host_alloc", "zsb_host_free", "zsb_message",
               "zsb_profile_read", "zsb_zip_scan", "zsb_zip_members", "zsb_zip_inflate", "zsb_gunzip_files", "zsb_plan_big", "zsb_inflate_streams", "zsb_gzip_header"]
ZLIB_SYMBOLS = ["inflateInit_", "inflateInit2_", "inflate", "inflateEnd", "inflateReset", "inflateReset2",
                "inflateResetKeep", "inflatePrime", "inflateGetDictionary", "inflateSetDictionary", "inflateGetHeader",
                "inflateSync", "inflateSyncPoint", "inflateCopy", "inflateUndermine", "inflateValidate", "inflateMark",
                "inflateCodesUsed"]

ZSB_RAW_ONLY = 4
ZSB_HAS_WRAPPED = 1


class Member(ctypes.Structure):
    _fields_ = [("inp", ctypes.c_void_p), ("in_len", ctypes.c_size_t), ("out", ctypes.c_void_p),
                ("out_cap", ctypes.c_size_t), ("window_bits", ctypes.c_int32), ("flags", ctypes.c_int32)]


class Result(ctypes.Structure):
    _fields_ = [("ret", ctypes.c_int32), ("msg_id", ctypes.c_int32), ("total_in", ctypes.c_uint64),
                ("total_out", ctypes.c_uint64), ("check", ctypes.c_uint32), ("data_type", ctypes.c_int32)]

    def astuple(self):
        return (self.ret, self.msg_id, self.total_in, self.total_out, self.check, self.data_type)

    def __repr__(self):
        return "Result(ret=%d msg=%d in=%d out=%d check=%08x dt=%d)" % self.astuple()


class DeflateMember(ctypes.Structure):
    _fields_ = [("inp", ctypes.c_void_p), ("in_len", ctypes.c_size_t), ("out", ctypes.c_void_p),
                ("out_cap", ctypes.c_size_t), ("window_bits", ctypes.c_int32), ("level", ctypes.c_int32)]


class DeflateResult(ctypes.Structure):
    _fields_ = [("ret", ctypes.c_int32), ("blocks", ctypes.c_int32), ("total_in", ctypes.c_uint64),
                ("total_out", ctypes.c_uint64), ("check", ctypes.c_uint32), ("reserved", ctypes.c_uint32)]


class ZipEntry(ctypes.Structure):
    _fields_ = [("header_off", ctypes.c_uint64), ("data_off", ctypes.c_uint64), ("comp_size", ctypes.c_uint64),
                ("uncomp_size", ctypes.c_uint64), ("crc32", ctypes.c_uint32), ("method", ctypes.c_uint16),
                ("gp_flags", ctypes.c_uint16), ("name_off", ctypes.c_uint32), ("name_len", ctypes.c_uint16),
                ("pad", ctypes.c_uint16)]


def zip_scan(lib, data):
    """-> (rc, [ZipEntry]) for archive bytes `data` (host walker, no GPU)."""
    buf = ctypes.create_string_buffer(data, max(len(data), 1))
    cnt = ctypes.c_size_t(0)
    rc = lib.zsb_zip_scan(buf, len(data), None, 0, ctypes.byref(cnt))
    if rc:
        return rc, []
    arr = (ZipEntry * max(cnt.value, 1))()
    rc = lib.zsb_zip_scan(buf, len(data), arr, cnt.value, ctypes.byref(cnt))
    return rc, [arr[i] for i in range(cnt.value)]


def build_product():
    subprocess.run(["make", "-s", "-C", os.path.join(PKG, "csrc")], check=True)


def load_product():
    if not os.path.exists(LIB):
        build_product()
    lib = ctypes.CDLL(LIB)
    lib.zsb_create.restype = ctypes.c_void_p
    lib.zsb_create.argtypes = [ctypes.POINTER(ctypes.c_int), ctypes.c_int]
    lib.zsb_destroy.argtypes = [ctypes.c_void_p]
    lib.zsb_create_error.restype = ctypes.c_char_p
    lib.zsb_last_error.restype = ctypes.c_char_p
    lib.zsb_last_error.argtypes = [ctypes.c_void_p]
    lib.zsb_num_devices.argtypes = [ctypes.c_void_p]
    lib.zsb_inflate_batch.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
    lib.zsb_inflate_batch_device.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint,
                                             ctypes.c_void_p]
    lib.zsb_set_option.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_long]
    lib.zsb_get_option.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
    lib.zsb_get_option.restype = ctypes.c_long
    lib.zsb_last_launches.argtypes = [ctypes.c_void_p]
    lib.zsb_last_launches.restype = ctypes.c_long
    lib.zsb_plan_shards.argtypes = [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    lib.zsb_host_alloc.restype = ctypes.c_void_p
    lib.zsb_host_alloc.argtypes = [ctypes.c_size_t]
    lib.zsb_host_free.argtypes = [ctypes.c_void_p]
    lib.zsb_message.restype = ctypes.c_char_p
    lib.zsb_message.argtypes = [ctypes.c_int]
    lib.zsb_zip_scan.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                 ctypes.POINTER(ctypes.c_size_t)]
    lib.zsb_zip_members.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p,
                                    ctypes.c_void_p]
    lib.zsb_zip_inflate.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                    ctypes.c_void_p, ctypes.c_void_p]
    lib.zsb_plan_big.restype = ctypes.c_size_t
    lib.zsb_plan_big.argtypes = [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    lib.zsb_deflate_batch.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
    lib.zsb_deflate_batch_static.argtypes = [ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                             ctypes.c_char_p, ctypes.c_size_t]
    lib.zsb_deflate_static_limits.restype = ctypes.c_size_t
    lib.zsb_deflate_static_limits.argtypes = [ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_size_t)]
    lib.zsb_deflate_bound.restype = ctypes.c_size_t
    lib.zsb_deflate_bound.argtypes = [ctypes.c_size_t, ctypes.c_int]
    lib.zsb_profile_read_check.restype = ctypes.c_long
    lib.zsb_profile_read_check.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    lib.zsb_profile_read_seg.restype = ctypes.c_long
    lib.zsb_profile_read_seg.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    lib.zsb_profile_read.restype = ctypes.c_long
    lib.zsb_profile_read.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_double),
                                     ctypes.POINTER(ctypes.c_double)]
    return lib


def plan_shards(lib, costs, nshards):
    n = len(costs)
    c = (ctypes.c_uint64 * n)(*costs)
    so = (ctypes.c_int32 * n)()
    load = (ctypes.c_uint64 * nshards)()
    rc = lib.zsb_plan_shards(n, c, nshards, so, load)
    if rc:
        raise RuntimeError("zsb_plan_shards rc=%d" % rc)
    return list(so), list(load)


class Zsb:
    """The engine through its C ABI.  Raises when no CUDA device is usable: there is no CPU path."""

    def __init__(self, devices=None, lanes=None):
        self.lib = load_product()
        if devices:
            arr = (ctypes.c_int * len(devices))(*devices)
            self.ctx = self.lib.zsb_create(arr, len(devices))
        else:
            self.ctx = self.lib.zsb_create(None, 0)
        if not self.ctx:
            raise RuntimeError("zsb_create failed: " + self.lib.zsb_create_error().decode())
        if lanes:
            self.set_option("lanes_per_member", lanes)

    def close(self):
        if self.ctx:
            self.lib.zsb_destroy(self.ctx)
            self.ctx = None

    def set_option(self, name, value):
        rc = self.lib.zsb_set_option(self.ctx, name.encode(), value)
        if rc:
            raise ValueError("zsb_set_option(%s,%r) rc=%d" % (name, value, rc))

    def get_option(self, name):
        return self.lib.zsb_get_option(self.ctx, name.encode())

    def error(self):
        return self.lib.zsb_last_error(self.ctx).decode()

    def inflate_batch(self, members, wbits, out_caps):
        """members: list of bytes; host buffers.  Returns [(Result, output bytes)]."""
        n = len(members)
        if n == 0:
            return []
        ins = [ctypes.create_string_buffer(m, max(len(m), 1)) for m in members]
        outs = [ctypes.create_string_buffer(max(c, 1)) for c in out_caps]
        arr = (Member * n)()
        for i in range(n):
            arr[i].inp = ctypes.addressof(ins[i]) if len(members[i]) else (ctypes.addressof(ins[i]))
            arr[i].in_len = len(members[i])
            arr[i].out = ctypes.addressof(outs[i])
            arr[i].out_cap = out_caps[i]
            arr[i].window_bits = wbits[i]
        res = (Result * n)()
        rc = self.lib.zsb_inflate_batch(self.ctx, n, arr, res)
        if rc:
            raise RuntimeError("zsb_inflate_batch rc=%d: %s" % (rc, self.error()))
        return [(res[i], outs[i].raw[:min(res[i].total_out, out_caps[i])]) for i in range(n)]


def deflate_batch(eng, plains, wbits, levels, caps=None):
    """zsb_deflate_batch over host buffers -> [(DeflateResult, compressed bytes)]."""
    n = len(plains)
    ins = [ctypes.create_string_buffer(p, max(len(p), 1)) for p in plains]
    caps = caps or [eng.lib.zsb_deflate_bound(len(p), w) for p, w in zip(plains, wbits)]
    outs = [ctypes.create_string_buffer(max(c, 1)) for c in caps]
    arr = (DeflateMember * n)()
    for i in range(n):
        arr[i].inp, arr[i].in_len = ctypes.addressof(ins[i]), len(plains[i])
        arr[i].out, arr[i].out_cap = ctypes.addressof(outs[i]), caps[i]
        arr[i].window_bits, arr[i].level = wbits[i], levels[i]
    res = (DeflateResult * n)()
    rc = eng.lib.zsb_deflate_batch(eng.ctx, n, arr, res)
    if rc:
        raise RuntimeError("zsb_deflate_batch rc=%d: %s" % (rc, eng.error()))
    return [(res[i], outs[i].raw[:res[i].total_out]) for i in range(n)]


class Emul:
    def __init__(self):
        name = "libemul_asan.so" if os.environ.get("ZSB_EMUL_SANITIZE") else "libemul.so"   # ASan build: LD_PRELOAD libasan
        path = os.path.join(EMUL_DIR, name)
        subprocess.run(["make", "-s", "-C", EMUL_DIR, path], check=True)
        self.lib = ctypes.CDLL(path)
        self.lib.emul_decode.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                         ctypes.c_int, ctypes.POINTER(Result), ctypes.c_void_p]
        self.lib.emul_decode_lane.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                                              ctypes.c_int, ctypes.POINTER(Result), ctypes.c_void_p,
                                              ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint32)]
        self.lib.emul_finish.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.POINTER(Result),
                                         ctypes.c_void_p, ctypes.c_uint32]

    @staticmethod
    def seq_bound(in_len, out_cap):
        """zsb_host.cu seq_bound(): most LZ sequences a member can produce."""
        return min(out_cap // 3 + in_len // 2 + 6, 8 * in_len + 4, out_cap + 2)

    def oneshot(self, data, out_cap, wbits, lane=False):
        import zlib
        # copy the input to a word-aligned buffer (the decoder reads whole aligned words)
        buf_in = ctypes.create_string_buffer(data, max((len(data) + 3) & ~3, 4))      # whole aligned words, as on the device
        buf = ctypes.create_string_buffer(max(out_cap, 1))
        r = Result()
        aux = (ctypes.c_uint32 * 4)()
        if lane:                                    # True = k_lane's decoder, "big" = k_big's (lane 0 + direct tables)
            self.lib.emul_set_big(1 if lane == "big" else 0)
            nseq = ctypes.c_uint32(0)
            rc = self.lib.emul_decode_lane(buf_in, len(data), buf if out_cap else None, out_cap, wbits,
                                           ctypes.byref(r), aux, self.seq_bound(len(data), out_cap), ctypes.byref(nseq))
            assert rc != -1000, "sequence bound too small: %d > %d" % (nseq.value, self.seq_bound(len(data), out_cap))
        else:
            self.lib.emul_decode(buf_in, len(data), buf if out_cap else None, out_cap, wbits, ctypes.byref(r), aux)
        out = buf.raw[:min(r.total_out, out_cap)]
        if aux[0]:
            chk = (zlib.adler32(out) if aux[0] == 1 else zlib.crc32(out)) & 0xffffffff
            self.lib.emul_finish(buf_in, len(data), ctypes.byref(r), aux, chk)
        return r, out


    def oneshot_seg(self, data, out_cap, wbits, segb, seg_q=None):
        """The segmented path (zsb_seg.cuh) on the host: -> (Result, output, info) with info = (starts found, chain held,
        segments, live segments)."""
        import zlib
        buf_in = ctypes.create_string_buffer(data, max((len(data) + 3) & ~3, 4))
        buf = ctypes.create_string_buffer(max(out_cap, 1))
        r = Result()
        aux = (ctypes.c_uint32 * 4)()
        info = (ctypes.c_uint32 * 4)()
        self.lib.emul_decode_seg.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                             ctypes.c_uint32, ctypes.c_uint32, ctypes.POINTER(Result), ctypes.c_void_p, ctypes.c_void_p]
        rc = self.lib.emul_decode_seg(buf_in, len(data), buf if out_cap else None, out_cap, wbits, segb,
                                      seg_q if seg_q is not None else max(segb * 3 // 4, 64), ctypes.byref(r), aux, info)
        assert rc != -1000
        out = buf.raw[:min(r.total_out, out_cap)]
        if aux[0]:
            chk = (zlib.adler32(out) if aux[0] == 1 else zlib.crc32(out)) & 0xffffffff
            self.lib.emul_finish(buf_in, len(data), ctypes.byref(r), aux, chk)
        return r, out, tuple(info)


_emul = None


def emul():
    global _emul
    if _emul is None:
        _emul = Emul()
    return _emul


# ---- the z_stream surface of libzsb.so (inflateInit2_ / inflate / inflateEnd), through ctypes ----
class ZStreamStruct(ctypes.Structure):
    _fields_ = [("next_in", ctypes.c_void_p), ("avail_in", ctypes.c_uint), ("total_in", ctypes.c_ulong),
                ("next_out", ctypes.c_void_p), ("avail_out", ctypes.c_uint), ("total_out", ctypes.c_ulong),
                ("msg", ctypes.c_char_p), ("state", ctypes.c_void_p), ("zalloc", ctypes.c_void_p),
                ("zfree", ctypes.c_void_p), ("opaque", ctypes.c_void_p), ("data_type", ctypes.c_int),
                ("adler", ctypes.c_ulong), ("reserved", ctypes.c_ulong)]


class ZStreamApi:
    """Drives libzsb.so's zlib entry points the way the reference's callers do
    (src/wasm/wasm_stream_common.c:37-63, test/payload_decompress.c:124-192)."""

    def __init__(self):
        self.lib = load_product()
        self.lib.inflateInit2_.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
        self.lib.inflate.argtypes = [ctypes.c_void_p, ctypes.c_int]
        self.lib.inflateEnd.argtypes = [ctypes.c_void_p]
        self.lib.inflateReset.argtypes = [ctypes.c_void_p]
        self.msgs = {self.lib.zsb_message(i): i for i in range(1, 19)}

    def init_ret(self, wbits):
        s = ZStreamStruct()
        rc = self.lib.inflateInit2_(ctypes.byref(s), wbits, b"1.3", ctypes.sizeof(s))
        if rc == 0:
            self.lib.inflateEnd(ctypes.byref(s))
        return rc

    def stream(self, data, wbits, in_chunks, out_cap, finish=True, max_calls=1000000, flush=0):
        """Same driver and trace format as orclib.Checker.stream()."""
        mid_flush = flush
        s = ZStreamStruct()
        rc = self.lib.inflateInit2_(ctypes.byref(s), wbits, b"1.3", ctypes.sizeof(s))
        if rc != 0:
            return [("init", rc)], b""
        trace, out = [], bytearray()
        buf = ctypes.create_string_buffer(max(out_cap, 1))
        pos, k = 0, 0
        pending = b""
        try:
            for _ in range(max_calls):
                if not pending and pos < len(data):
                    n = in_chunks[k % len(in_chunks)]
                    k += 1
                    pending = data[pos:pos + n]
                    pos += len(pending)
                flush = 4 if (finish and pos >= len(data)) else mid_flush
                ibuf = ctypes.create_string_buffer(pending, max(len(pending), 1))
                s.next_in = ctypes.addressof(ibuf)
                s.avail_in = len(pending)
                s.next_out = ctypes.addressof(buf)
                s.avail_out = out_cap
                ret = self.lib.inflate(ctypes.byref(s), flush)
                c, p = len(pending) - s.avail_in, out_cap - s.avail_out
                out += buf.raw[:p]
                trace.append((ret, self.msgs.get(s.msg, 0) if s.msg else 0, s.total_in, s.total_out,
                              s.adler & 0xffffffff, s.data_type, c, p))
                pending = pending[c:]
                if ret not in (0, -5):
                    break
                if ret == -5 and c == 0 and p == 0 and pos >= len(data):
                    break
        finally:
            self.lib.inflateEnd(ctypes.byref(s))
        return trace, bytes(out)


    def stream_multi(self, jobs, finish=True, max_ticks=1000000):
        """jobs: [(data, wbits, in_chunks, out_cap)].  All streams advance together: every tick is ONE
        zsb_inflate_streams() call over the streams still running.  Per stream the driver and the trace format
        are those of stream() / orclib.Checker.stream()."""
        n = len(jobs)
        self.lib.zsb_inflate_streams.argtypes = [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        st = []
        for data, wbits, in_chunks, out_cap in jobs:
            s = ZStreamStruct()
            rc = self.lib.inflateInit2_(ctypes.byref(s), wbits, b"1.3", ctypes.sizeof(s))
            st.append(dict(s=s, init=rc, trace=[("init", rc)] if rc else [], out=bytearray(), pos=0, k=0, pending=b"",
                           buf=ctypes.create_string_buffer(max(out_cap, 1)), done=rc != 0, data=data, chunks=in_chunks,
                           cap=out_cap, ibuf=None))
        try:
            for _ in range(max_ticks):
                act = [x for x in st if not x["done"]]
                if not act:
                    break
                # streams are grouped by the flush value their driver would pass
                for flush in (0, 4):
                    grp = []
                    for x in act:
                        if not x["pending"] and x["pos"] < len(x["data"]):
                            c = x["chunks"][x["k"] % len(x["chunks"])]
                            x["k"] += 1
                            x["pending"] = x["data"][x["pos"]:x["pos"] + c]
                            x["pos"] += len(x["pending"])
                        f = 4 if (finish and x["pos"] >= len(x["data"])) else 0
                        if f == flush and not x.get("ticked"):
                            grp.append(x)
                    if not grp:
                        continue
                    ptrs = (ctypes.c_void_p * len(grp))()
                    rets = (ctypes.c_int * len(grp))()
                    for i, x in enumerate(grp):
                        s = x["s"]
                        x["ibuf"] = ctypes.create_string_buffer(x["pending"], max(len(x["pending"]), 1))
                        s.next_in = ctypes.addressof(x["ibuf"])
                        s.avail_in = len(x["pending"])
                        s.next_out = ctypes.addressof(x["buf"])
                        s.avail_out = x["cap"]
                        ptrs[i] = ctypes.addressof(s)
                        x["ticked"] = True
                    rc = self.lib.zsb_inflate_streams(len(grp), ptrs, flush, rets)
                    assert rc == 0
                    for i, x in enumerate(grp):
                        s, ret = x["s"], rets[i]
                        c, p = len(x["pending"]) - s.avail_in, x["cap"] - s.avail_out
                        x["out"] += x["buf"].raw[:p]
                        x["trace"].append((ret, self.msgs.get(s.msg, 0) if s.msg else 0, s.total_in, s.total_out,
                                           s.adler & 0xffffffff, s.data_type, c, p))
                        x["pending"] = x["pending"][c:]
                        if ret not in (0, -5) or (ret == -5 and c == 0 and p == 0 and x["pos"] >= len(x["data"])):
                            x["done"] = True
                for x in act:
                    x["ticked"] = False
        finally:
            for x in st:
                if x["init"] == 0:
                    self.lib.inflateEnd(ctypes.byref(x["s"]))
        return [(x["trace"], bytes(x["out"])) for x in st]
