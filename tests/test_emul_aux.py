"""CPU tier: what inflateMark(), inflateCodesUsed() and inflateSyncPoint() report after every inflate() call —
the device code of the z_stream surface (zsb_stream.cuh, compiled for the host by tests/emul) against the
UNMODIFIED reference, over chunked traces with Z_NO_FLUSH and Z_BLOCK."""
import ctypes
import os
import random

import pytest

import cases
import orclib
import zapi
import zsblib

REF = zapi.load("ref")
pytestmark = pytest.mark.skipif(REF is None, reason="oracle/_ref not built (no /root/reference here)")


def emul_trace(data, wb, chunks, ocap, flush):
    lib = zsblib.emul().lib
    lib.emulzs_open.restype = ctypes.c_void_p
    lib.emulzs_open.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    lib.emulzs_step.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(orclib.Result)]
    lib.emulzs_aux.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_long)]
    lib.emulzs_close.argtypes = [ctypes.c_void_p]
    rc = ctypes.c_int()
    h = lib.emulzs_open(wb, ctypes.byref(rc))
    if not h:
        return rc.value
    pos, k, pending, trace = 0, 0, b"", []
    buf = ctypes.create_string_buffer(max(ocap, 1))
    c, p, r, aux = ctypes.c_size_t(), ctypes.c_size_t(), orclib.Result(), (ctypes.c_long * 3)()
    for _ in range(40000):
        if not pending and pos < len(data):
            n = chunks[k % len(chunks)]
            k += 1
            pending = data[pos:pos + n]
            pos += len(pending)
        fl = 4 if pos >= len(data) else flush
        ret = lib.emulzs_step(h, pending, len(pending), buf, ocap, fl, ctypes.byref(c), ctypes.byref(p), ctypes.byref(r))
        lib.emulzs_aux(h, aux)
        trace.append((ret, c.value, p.value, r.data_type, aux[0], aux[1], aux[2]))
        pending = pending[c.value:]
        if ret not in (0, -5) or (ret == -5 and c.value == 0 and p.value == 0 and pos >= len(data)):
            break
    lib.emulzs_close(h)
    return trace


def ref_trace(data, wb, chunks, ocap, flush):
    z = zapi.Z(REF, wb)
    if z.init:
        return z.init
    t, _ = z.drive(data, chunks, ocap, flush=flush, max_calls=40000)
    z.end()
    out = []
    for ret, c, p, obs in t:
        o = dict(obs)
        out.append((ret, c, p, o["data_type"], o["mark"], o["codes_used"], o["sync_point"]))
    return out


def test_mark_codes_used_sync_point_traces():
    # (ZSB_TEST_SCALE=full: 300 damaged streams and every fifth valid one — 2 minutes; the default keeps the CPU tier short)
    full = os.environ.get("ZSB_TEST_SCALE") == "full"
    rnd = random.Random(13)
    pool = [(n, d, w) for n, d, w, p in cases.valid_streams()[1::5 if full else 14]] + [(n, d, -16) for n, d, m in cases.golden_fixtures()[:5 if full else 2]]
    pool += [(n, d, w) for n, d, w, c in cases.damaged_streams(300)[:300 if full else 90]]
    pool += [(n, d, w) for n, d, w, c, _ in cases.probe_vectors()]
    for name, data, wb in pool:
        for flush in (0, 5, 6):
            chunks = [rnd.choice([1, 5, 64, 1000, 32768]) for _ in range(4)] if len(data) < 30000 else [32768]
            ocap = rnd.choice([7, 300, 65536])
            a, b = ref_trace(data, wb, chunks, ocap, flush), emul_trace(data, wb, chunks, ocap, flush)
            if a != b:
                k = next((i for i, (x, y) in enumerate(zip(a, b)) if x != y), min(len(a), len(b))) if isinstance(a, list) else 0
                raise AssertionError((name, wb, flush, chunks, ocap, "call", k, "ref", a[max(k - 1, 0):k + 1] if isinstance(a, list) else a,
                                      "emul", b[max(k - 1, 0):k + 1] if isinstance(b, list) else b))
