"""TEST TOOL: the reference's own callers of the inflate path, compiled unmodified by `make -C oracle callers`
against libzsb.so ("zsb") and against the reference itself ("ref"):
  * test/payload_decompress.c and test/payload_decompress_nowindow.c as binaries (a 64 KiB rolling output window),
  * the WASM glue (src/wasm/wasm_stream_common.c + inflate_stream_wasm.c + inflate9_stream_wasm.c + allocator.c)
    as a native shared object, driven the way src/wasm/api/zlib-streams.js:73-138 drives it: 32 KiB input slices
    with flush 0, then flush 4 until the packed result says end or nothing was produced."""
import ctypes
import hashlib
import os
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFDIR = os.path.join(ROOT, "oracle", "_ref")


def built(variant):
    return all(os.path.exists(os.path.join(REFDIR, n % variant)) for n in
               ("payload_decompress_%s", "payload_decompress_nowindow_%s", "libglue_%s.so"))


def run_binary(name, variant, data):
    """-> (exit code, sha256 of the output file, output length)"""
    with tempfile.TemporaryDirectory() as d:
        src, dst = os.path.join(d, "in"), os.path.join(d, "out")
        open(src, "wb").write(data)
        p = subprocess.run([os.path.join(REFDIR, "%s_%s" % (name, variant)), src, dst], capture_output=True, timeout=600)
        out = open(dst, "rb").read() if os.path.exists(dst) else b""
        return p.returncode, hashlib.sha256(out).hexdigest(), len(out)


class Glue:
    def __init__(self, variant):
        # RTLD_DEEPBIND: the glue's inflateInit2_/inflate/inflateEnd must bind to the library it was linked against,
        # not to the system libz.so.1 that CPython has already mapped into the global scope
        self.lib = ctypes.CDLL(os.path.join(REFDIR, "libglue_%s.so" % variant), mode=os.RTLD_NOW | os.RTLD_DEEPBIND)
        up = ctypes.c_size_t
        for fam in ("inflate", "inflate9"):
            getattr(self.lib, fam + "_new").restype = up
            for fn in ("_init", "_init_raw", "_end", "_last_consumed") + (("_init_gzip",) if fam == "inflate" else ()):
                getattr(self.lib, fam + fn).argtypes = [up]
            getattr(self.lib, fam + "_last_consumed").restype = ctypes.c_uint
            getattr(self.lib, fam + "_process").argtypes = [up, up, ctypes.c_uint, up, ctypes.c_uint, ctypes.c_int]

    def replay(self, family, init, data, out_size=65536, in_slice=32768):
        """The JS TransformStream's transform() + flush() loops; returns (init code, [packed result words],
        [last_consumed...], output bytes, end code)."""
        L = self.lib
        h = getattr(L, family + "_new")()
        rc = getattr(L, family + init)(h)
        words, consumed, out = [], [], bytearray()
        obuf = ctypes.create_string_buffer(out_size)
        process, last = getattr(L, family + "_process"), getattr(L, family + "_last_consumed")
        failed = rc != 0
        off = 0
        while not failed and off < len(data):
            chunk = data[off:off + in_slice]
            ibuf = ctypes.create_string_buffer(chunk, len(chunk))
            w = process(h, ctypes.addressof(ibuf), len(chunk), ctypes.addressof(obuf), out_size, 0) & 0xffffffff
            words.append(w)
            out += obuf.raw[:w & 0xffffff]
            code = (w >> 24) & 0xff
            if code & 0x80:
                failed = True
                break
            c = last(h)
            consumed.append(c)
            if c == 0:
                break
            off += c
        while not failed:
            w = process(h, 0, 0, ctypes.addressof(obuf), out_size, 4) & 0xffffffff
            words.append(w)
            out += obuf.raw[:w & 0xffffff]
            code = (w >> 24) & 0xff
            if code & 0x80 or code == 1 or (w & 0xffffff) == 0:
                break
        end = getattr(L, family + "_end")(h)
        return rc, words, consumed, bytes(out), end
