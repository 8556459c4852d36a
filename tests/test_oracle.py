"""CPU tier: pins the oracle.  (1) the reference's own golden fixtures, (2) the plain-C restatement
(oracle/orc_inflate.c) against the unmodified reference compiled in place (oracle/_ref), on valid,
damaged, truncated and capacity-limited streams, one-shot and in random chunkings."""
import hashlib
import random

import pytest

import cases
import orclib

ORC = orclib.orc()
REF = orclib.ref()
needs_ref = pytest.mark.skipif(REF is None, reason="oracle/_ref not built (no /root/reference here)")
CHECKERS = [("orc", ORC)] + ([("ref", REF)] if REF is not None else [])


@pytest.mark.parametrize("cname,chk", CHECKERS)
def test_golden_fixtures(cname, chk):
    for name, data, m in cases.golden_fixtures():
        r, out = chk.oneshot(data, m["out_len"] + 64, -16)
        assert (r.ret, r.msg_id, r.total_in, r.total_out) == (1, 0, m["in_len"], m["out_len"]), (cname, name, r)
        assert hashlib.sha256(out).hexdigest() == m["out_sha256"], (cname, name)


def test_regenerable_plaintexts():
    fx = {n: (d, m) for n, d, m in cases.golden_fixtures()}
    want = {"10k_lines": b"".join(b"line-%06d\n" % i for i in range(10000)),
            "100k_lines": b"".join(b"Sample content %d.\n" % i for i in range(100000)),
            "zeros_100k": bytes(100000)}
    for name, plain in want.items():
        data, m = fx[name]
        r, out = ORC.oneshot(data, len(plain) + 1, -16)
        assert out == plain


@pytest.mark.parametrize("cname,chk", CHECKERS)
def test_probe_vectors(cname, chk):
    for name, data, wb, cap, want in cases.probe_vectors():
        r, out = chk.oneshot(data, cap, wb)
        assert (r.ret, r.msg_id, r.total_in, r.total_out) == want, (cname, name, r)


@pytest.mark.parametrize("cname,chk", CHECKERS)
def test_valid_streams_roundtrip(cname, chk):
    for name, data, wb, plain in cases.valid_streams():
        r, out = chk.oneshot(data, len(plain) + 16, wb)
        assert r.ret == 1 and out == plain and r.total_in == len(data), (cname, name, r)


@needs_ref
def test_restatement_equals_reference_oneshot():
    n = 0
    for name, data, wb, plain in cases.valid_streams():
        for cap in (len(plain) + 16, len(plain), max(len(plain) - 1, 0)):
            a, ao = REF.oneshot(data, cap, wb)
            b, bo = ORC.oneshot(data, cap, wb)
            assert a.astuple() == b.astuple() and ao == bo, (name, cap, a, b)
            n += 1
    for name, data, wb, cap in cases.damaged_streams(600):
        a, ao = REF.oneshot(data, cap, wb)
        b, bo = ORC.oneshot(data, cap, wb)
        assert a.astuple() == b.astuple() and ao == bo, (name, a, b)
        n += 1
    assert n > 1000


@needs_ref
def test_restatement_equals_reference_streaming():
    rnd = random.Random(7)
    pool = cases.valid_streams() + [(n, d, w, b"x" * c) for n, d, w, c in cases.damaged_streams(80)[:80]]
    fx = [(n, d, -16, b"x" * m["out_len"]) for n, d, m in cases.golden_fixtures()]
    for name, data, wb, plain in pool[::3] + fx:
        for trial in range(2):
            chunks = [rnd.choice([1, 2, 3, 7, 64, 1000, 32768]) for _ in range(5)]
            ocap = rnd.choice([1, 5, 300, 65536])
            ta, oa = REF.stream(data, wb, chunks, ocap, max_calls=400000)
            tb, ob = ORC.stream(data, wb, chunks, ocap, max_calls=400000)
            assert ta == tb and oa == ob, (name, chunks, ocap)


def test_init_codes():
    for wb in (-17, -16, -15, -8, -7, 0, 7, 8, 15, 16, 23, 24, 31, 47, 48, 63):
        a = ORC.init_ret(wb)
        if REF is not None:
            assert a == REF.init_ret(wb), wb
