"""GPU tier, SURVEY.md §8 a13: the auxiliary zlib entry points of libzsb.so against the UNMODIFIED reference
(oracle/_ref/libzsb_ref.so), both driven by the same script through tests/zapi.py: inflateSetDictionary /
GetDictionary, inflateGetHeader, inflateCopy, inflateSync / SyncPoint, inflatePrime, inflateValidate,
inflateUndermine, and inflateMark / inflateCodesUsed observed after EVERY inflate() call of chunked traces."""
import ctypes
import random
import struct
import zlib

import pytest

import cases
import gen
import zapi

pytestmark = pytest.mark.gpu
REF = zapi.load("ref")
ZSB = zapi.load("zsb")
if __import__("os").environ.get("ZSB_AUX_SELFTEST"):     # harness check without a GPU: the reference against itself
    ZSB = REF
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built")


def first_diff(a, b, path=""):
    """Path and values of the first leaf where two nested observations differ."""
    if type(a) != type(b):
        return "%s: %r != %r" % (path, a, b)
    if isinstance(a, (list, tuple)):
        if len(a) != len(b):
            return "%s: lengths %d != %d (%r ... | %r ...)" % (path, len(a), len(b), a[-1:], b[-1:])
        for i, (x, y) in enumerate(zip(a, b)):
            if x != y:
                return first_diff(x, y, "%s[%d]" % (path, i))
    if isinstance(a, dict):
        for k in a:
            if a[k] != b.get(k):
                return first_diff(a[k], b.get(k), "%s[%r]" % (path, k))
    if isinstance(a, bytes) and a != b:
        i = next((i for i, (x, y) in enumerate(zip(a, b)) if x != y), min(len(a), len(b)))
        return "%s: bytes differ at %d of %d/%d: %r != %r" % (path, i, len(a), len(b), a[i:i + 12], b[i:i + 12])
    return "%s: %r != %r" % (path, a, b)


def both(fn, what=""):
    """Run fn(lib) on the reference and on the product; the two observations must be equal."""
    a, b = fn(REF), fn(ZSB)
    assert a == b, "%s reference != product at %s" % (what, first_diff(a, b))
    return b


@needs_ref
def test_mark_codes_used_sync_point_after_every_call():
    rnd = random.Random(13)
    pool = [(n, d, w) for n, d, w, p in cases.valid_streams()[1::7]] + [(n, d, -16) for n, d, m in cases.golden_fixtures()[:4]]
    pool += [(n, d, w) for n, d, w, c in cases.damaged_streams(40)[:40:3]]
    for name, data, wb in pool:
        for flush in (0, 5):
            chunks = [rnd.choice([1, 5, 64, 1000, 32768]) for _ in range(4)] if len(data) < 30000 else [32768]
            ocap = rnd.choice([7, 300, 65536])

            def run(lib):
                z = zapi.Z(lib, wb)
                if z.init:
                    return z.init
                t, out = z.drive(data, chunks, ocap, flush=flush, max_calls=40000)
                z.end()
                return t, out
            both(run, "%s wb=%d flush=%d chunks=%r ocap=%d" % (name, wb, flush, chunks, ocap))


@needs_ref
def test_set_and_get_dictionary():
    plain = gen.mixed_plain(40000, 21)
    zdict = gen.word_soup(20000, 22)
    for wb in (15, -15):
        co = zlib.compressobj(6, zlib.DEFLATED, wb, 8, zlib.Z_DEFAULT_STRATEGY, zdict)
        comp = co.compress(plain) + co.flush()

        def run(lib):
            z = zapi.Z(lib, wb)
            obs = []
            if wb < 0:
                obs.append(lib.inflateSetDictionary(ctypes.byref(z.s), zdict, len(zdict)))
            ret, c, o = z.call(comp, 1000)
            obs.append((ret, c, o, z.observe()))
            if ret == 2:
                obs.append(lib.inflateSetDictionary(ctypes.byref(z.s), b"wrong" * 10, 50))      # Z_DATA_ERROR
                obs.append(z.call(comp[c:], 1000)[0])                                            # still Z_NEED_DICT
                obs.append(lib.inflateSetDictionary(ctypes.byref(z.s), zdict, len(zdict)))
                obs.append(z.observe())
            else:
                obs.append(lib.inflateSetDictionary(ctypes.byref(z.s), zdict, len(zdict)) if wb > 0 else None)
            t, out = z.drive(comp[c:], [700, 5000], 3000)
            obs.append((t, out))
            buf = ctypes.create_string_buffer(70000)
            n = ctypes.c_uint(0)
            obs.append(lib.inflateGetDictionary(ctypes.byref(z.s), buf, ctypes.byref(n)))
            obs.append((n.value, buf.raw[:n.value]))
            z.end()
            return obs
        obs = both(run)
        assert (o for o in obs)
    # the decoded text is the plaintext
    z = zapi.Z(ZSB, -15)
    ZSB.inflateSetDictionary(ctypes.byref(z.s), zdict, len(zdict))
    co = zlib.compressobj(6, zlib.DEFLATED, -15, 8, zlib.Z_DEFAULT_STRATEGY, zdict)
    comp = co.compress(plain) + co.flush()
    t, out = z.drive(comp, [len(comp)], len(plain) + 10)
    assert out == plain and t[-1][0] == 1


@needs_ref
def test_get_dictionary_tracks_the_window():
    plain = gen.mixed_plain(150000, 5)
    for wb, enc in ((-15, -15), (31, 31), (-16, -15), (-9, -9)):
        comp = gen.deflate(plain, 6, enc)

        def run(lib):
            z = zapi.Z(lib, wb)
            obs = []
            pos = 0
            for step in (100, 3000, 40000, 1 << 20):
                ret, c, o = z.call(comp[pos:pos + step], 1 << 20)
                pos += c
                buf = ctypes.create_string_buffer(70000)
                n = ctypes.c_uint(0)
                rc = lib.inflateGetDictionary(ctypes.byref(z.s), buf, ctypes.byref(n))
                obs.append((ret, c, len(o), rc, n.value, buf.raw[:n.value]))
            obs.append(lib.inflateResetKeep(ctypes.byref(z.s)))
            n = ctypes.c_uint(0)
            lib.inflateGetDictionary(ctypes.byref(z.s), None, ctypes.byref(n))
            obs.append(n.value)                          # inflateResetKeep keeps the window
            obs.append(lib.inflateReset(ctypes.byref(z.s)))
            lib.inflateGetDictionary(ctypes.byref(z.s), None, ctypes.byref(n))
            obs.append(n.value)                          # inflateReset empties it
            z.end()
            return obs
        both(run)


@needs_ref
def test_get_header():
    p = gen.mixed_plain(3000, 77)
    variants = [cases.gzip_with_fields(p), cases.gzip_with_fields(p, extra=b"xy" * 40, name=b"a-file-name.txt", comment=b"some comment", hcrc=True),
                cases.gzip_with_fields(p, name=b"", comment=b"c"), cases.gzip_with_fields(p, extra=b""), gen.deflate(p, 6, 15)]
    for data in variants:
        for wb in (31, 47):
            for chunks in ([len(data)], [1], [7, 13]):
                for maxes in ((64, 64, 64), (5, 4, 3), (0, 0, 0)):
                    def run(lib):
                        z = zapi.Z(lib, wb)
                        h = zapi.GzHeader()
                        eb, nb, cb = (ctypes.create_string_buffer(b"\xaa" * 80, 80) for _ in range(3))
                        h.extra, h.extra_max = ctypes.addressof(eb), maxes[0]
                        h.name, h.name_max = ctypes.addressof(nb), maxes[1]
                        h.comment, h.comm_max = ctypes.addressof(cb), maxes[2]
                        if maxes == (0, 0, 0):
                            h.extra = h.name = h.comment = None
                        rc = lib.inflateGetHeader(ctypes.byref(z.s), ctypes.byref(h))
                        t, out = z.drive(data, chunks, 4096)
                        z.end()
                        return (rc, t[-1][0], out, h.text, h.time, h.xflags, h.os, bool(h.extra), h.extra_len, bool(h.name),
                                bool(h.comment), h.hcrc, h.done, eb.raw, nb.raw, cb.raw)
                    both(run)
    # not a gzip-capable stream: Z_STREAM_ERROR
    both(lambda lib: lib.inflateGetHeader(ctypes.byref(zapi.Z(lib, 15).s), ctypes.byref(zapi.GzHeader())))


@needs_ref
def test_copy_mid_stream():
    plain = gen.mixed_plain(120000, 9)
    for wb in (-15, 15, 31):
        comp = gen.deflate(plain, 6, wb)

        def run(lib):
            z = zapi.Z(lib, wb)
            ret, c, o = z.call(comp[:9000], 50000)
            d = zapi.Z.__new__(zapi.Z)
            d.lib, d.s, d.keep = lib, zapi.ZStream(), []
            rc = lib.inflateCopy(ctypes.byref(d.s), ctypes.byref(z.s))
            t1, o1 = z.drive(comp[c:], [4000], 30000)
            t2, o2 = d.drive(comp[c:], [1 << 20], 1 << 20)
            e = (z.end(), d.end())
            return rc, ret, c, o, t1, o1, t2[-1], o2, e
        r = both(run)
        assert r[3] + r[5] == plain and r[3] + r[7] == plain


@needs_ref
def test_sync_after_damage():
    co = zlib.compressobj(6, zlib.DEFLATED, 15)
    parts = [gen.mixed_plain(20000, 30 + i) for i in range(4)]
    comp = b""
    marks = []
    for p in parts:
        comp += co.compress(p) + co.flush(zlib.Z_FULL_FLUSH)
        marks.append(len(comp))
    comp += co.flush()
    bad = bytearray(comp)
    bad[marks[0] + 40] ^= 0x55                     # damage inside the second part
    bad = bytes(bad)
    for wb in (15, 47):
        def run(lib):
            z = zapi.Z(lib, wb)
            obs = []
            ret, c, o = z.call(bad, 1 << 20)
            obs.append((ret, c, len(o), z.observe()))
            pos = c
            for _ in range(6):
                ibuf = ctypes.create_string_buffer(bad[pos:pos + 7000], max(len(bad[pos:pos + 7000]), 1))
                z.s.next_in, z.s.avail_in = ctypes.addressof(ibuf), len(bad[pos:pos + 7000])
                rc = lib.inflateSync(ctypes.byref(z.s))
                used = len(bad[pos:pos + 7000]) - z.s.avail_in
                pos += used
                obs.append((rc, used, z.observe()))
                if rc == 0:
                    break
            t, out = z.drive(bad[pos:], [5000], 1 << 20)
            obs.append((t, out))
            z.end()
            return obs
        r = both(run)
        assert r[-1][1].endswith(parts[3])
    # no input and nothing held: Z_BUF_ERROR
    both(lambda lib: lib.inflateSync(ctypes.byref(zapi.Z(lib, -15).s)))


@needs_ref
def test_prime_a_raw_stream_at_a_bit_offset():
    plain = gen.mixed_plain(30000, 41)
    raw = gen.deflate(plain, 6, -15)
    for k in (1, 3, 7, 8, 13, 16):
        # the stream's first k bits go in through inflatePrime, the rest is the stream shifted down by k bits
        v = int.from_bytes(raw, "little")
        first = v & ((1 << k) - 1)
        rest = (v >> k).to_bytes(len(raw), "little")

        def run(lib):
            z = zapi.Z(lib, -15)
            a = lib.inflatePrime(ctypes.byref(z.s), k if k <= 16 else 16, first)
            t, out = z.drive(rest, [len(rest)], len(plain) + 10, finish=False)
            z.end()
            return a, [x[:3] for x in t], out
        r = both(run)
        assert r[2] == plain
    def limits(lib):
        z = zapi.Z(lib, -15)
        p = ctypes.byref(z.s)
        return (lib.inflatePrime(p, 17, 0), lib.inflatePrime(p, 16, 1), lib.inflatePrime(p, 16, 2), lib.inflatePrime(p, 1, 0),
                lib.inflatePrime(p, -1, 0), lib.inflatePrime(p, 0, 0), z.end())
    both(limits)


@needs_ref
def test_validate_and_undermine():
    p = gen.mixed_plain(5000, 3)
    for wb in (15, 31):
        good = gen.deflate(p, 6, wb)
        bad = bytearray(good)
        bad[-2 if wb == 15 else -6] ^= 1             # the check value
        for data in (good, bytes(bad)):
            for check in (0, 1):
                def run(lib):
                    z = zapi.Z(lib, wb)
                    a = lib.inflateValidate(ctypes.byref(z.s), check)
                    u = lib.inflateUndermine(ctypes.byref(z.s), 1)
                    t, out = z.drive(data, [3000], 8000)
                    z.end()
                    return a, u, t, out
                both(run)
