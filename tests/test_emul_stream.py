"""CPU tier: the z_stream surface's DEVICE code (zlib-streams_b200/csrc/zsb_stream.cuh: resumable decode,
carry, trailer, window accounting), compiled for the host by tests/emul with a warp of one lane,
against the oracle call by call: ret, msg, total_in, total_out, adler, data_type, consumed and
produced of every inflate() call under random chunkings, plus the output bytes.  The same traces are
checked on the GPU through libzsb.so's inflateInit2_/inflate/inflateEnd in test_gpu_parity.py."""
import ctypes
import os
import random

import cases
import orclib
import zsblib

CHK = orclib.ref() or orclib.orc()


class EmulStream(orclib.Checker):
    """open/step/close of tests/emul (prefix emulzs) behind orclib.Checker.stream()."""

    def __init__(self):
        zsblib.emul()                      # builds libemul.so
        self.lib = zsblib.emul().lib
        self._open = self.lib.emulzs_open
        self._open.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        self._open.restype = ctypes.c_void_p
        self._step = self.lib.emulzs_step
        self._step.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t,
                               ctypes.c_int, ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_size_t),
                               ctypes.POINTER(orclib.Result)]
        self._step.restype = ctypes.c_int
        self._close = self.lib.emulzs_close
        self._close.argtypes = [ctypes.c_void_p]


EM = EmulStream()


def pool():
    p = cases.valid_streams() + [(n, d, w, b"x" * c) for n, d, w, c in cases.damaged_streams(150)[:400]]
    p += [(n, d, -16, b"x" * m["out_len"]) for n, d, m in cases.golden_fixtures()]
    p += [(n, d, w, b"x" * c) for n, d, w, c, _ in cases.probe_vectors()]
    return p


def compare(name, data, wb, chunks, ocap, finish=True, flush=0):
    ta, oa = CHK.stream(data, wb, chunks, ocap, finish, max_calls=400000, flush=flush)
    tb, ob = EM.stream(data, wb, chunks, ocap, finish, max_calls=400000, flush=flush)
    if ta != tb:
        k = next((i for i, (x, y) in enumerate(zip(ta, tb)) if x != y), min(len(ta), len(tb)))
        raise AssertionError((name, wb, chunks, ocap, "call", k, ta[k - 1:k + 1], tb[k - 1:k + 1]))
    assert oa == ob, (name, "output differs")


def test_stream_traces_random_chunkings():
    # (ZSB_TEST_SCALE=full: every second stream, two chunkings each — 2.5 minutes; the default keeps the CPU tier short)
    full = os.environ.get("ZSB_TEST_SCALE") == "full"
    rnd = random.Random(11)
    for name, data, wb, plain in pool()[::2 if full else 5]:
        for trial in range(2 if full else 1):
            chunks = [rnd.choice([1, 2, 3, 7, 64, 1000, 32768]) for _ in range(5)]
            ocap = rnd.choice([1, 5, 300, 65536])
            compare(name, data, wb, chunks, ocap)


def test_stream_wasm_pattern():
    """The reference's JS caller: 32 KiB input slices, 64 KiB output, Z_NO_FLUSH then Z_FINISH
    (src/wasm/api/zlib-streams.js:73-138)."""
    for name, data, wb, plain in pool()[1::4]:
        compare(name, data, wb, [32768], 65536)


def test_stream_byte_at_a_time():
    for name, data, wb, plain in pool()[::9]:
        if len(data) <= 20000:
            compare(name, data, wb, [1], 1)
            compare(name, data, wb, [1], 70000)


def test_stream_block_and_trees_flush():
    """inflate(Z_BLOCK) / inflate(Z_TREES): the early returns at block boundaries and header ends and the
    128 / 256 bits of data_type (reference src/inflate.c:832,854-857,883,1032,1269-1271), call by call."""
    rnd = random.Random(56)
    stops = {5: 0, 6: 0}
    for name, data, wb, plain in pool()[::3]:
        for flush in (5, 6):
            chunks = [rnd.choice([1, 3, 64, 1000, 32768, 1 << 20]) for _ in range(4)]
            ocap = rnd.choice([5, 300, 65536, 1 << 20])
            compare(name, data, wb, chunks, ocap, finish=rnd.random() < 0.5, flush=flush)
    # the fixtures whole: every block boundary is a stop
    for name, data, m in cases.golden_fixtures():
        for flush in (5, 6):
            compare(name, data, -16, [len(data)], m["out_len"] + 10, finish=False, flush=flush)
            t, _ = EM.stream(data, -16, [len(data)], m["out_len"] + 10, finish=False, flush=flush)
            stops[flush] += sum(1 for c in t if c[0] == 0 and (c[5] & 0x180))
    assert stops[5] > 10 and stops[6] > stops[5]
