"""Deflate64-specific parity (SURVEY.md §7 hard part 5, §8 C3): streams from the test-only encoder that
use length code 285 + 16 extra bits, distance codes 30/31, matches up to 65538 and distances up to 64 KiB.
CPU tier: the encoder is admitted by the REFERENCE build (it must decode every stream back to the
plaintext), then the restatement and the decoders' device source (host emulation) must agree with it on
the valid, damaged, truncated and capacity-limited streams.  GPU tier: the same through libzsb.so."""
import pytest

import cases
import orclib
import zsblib

REF = orclib.ref()
CHK = REF or orclib.orc()


def test_encoder_admitted_by_reference():
    for name, data, wb, plain in cases.deflate64_streams():
        r, out = CHK.oneshot(data, len(plain) + 8, wb)
        assert (r.ret, r.total_in, r.total_out) == (1, len(data), len(plain)), (name, r)
        assert out == plain, name


@pytest.mark.skipif(REF is None, reason="reference build absent (GPU box): the restatement is the checker there")
def test_restatement_matches_reference():
    orc = orclib.orc()
    items = [(n, d, w, len(p) + 8) for n, d, w, p in cases.deflate64_streams()] + cases.deflate64_damaged(120)
    for name, data, wb, cap in items:
        a, ao = REF.oneshot(data, cap, wb)
        b, bo = orc.oneshot(data, cap, wb)
        assert a.astuple() == b.astuple() and ao == bo, (name, a, b)


@pytest.mark.parametrize("lane", [True, False, "big"], ids=["lane", "group", "big"])
def test_emulated_device_source(lane):
    em = zsblib.emul()
    items = [(n, d, w, len(p) + 8) for n, d, w, p in cases.deflate64_streams()] + cases.deflate64_damaged(120)
    for name, data, wb, cap in items:
        a, ao = CHK.oneshot(data, cap, wb)
        b, bo = em.oneshot(data, cap, wb, lane)
        assert a.astuple() == b.astuple(), (name, wb, cap, a, b)
        assert ao == bo, (name, "output differs")


@pytest.mark.gpu
@pytest.mark.parametrize("lanes", [0, "big", 16])
def test_gpu_deflate64(lanes):
    eng = zsblib.Zsb()
    try:
        if lanes == "big":                       # every member through k_big
            eng.set_option("path", 0)
            eng.set_option("split_big", 2)
        else:
            eng.set_option("path", 1 if lanes == 0 else 2)
            if lanes:
                eng.set_option("lanes_per_member", lanes)
        items = [(n, d, w, len(p) + 8) for n, d, w, p in cases.deflate64_streams(big=(lanes != 16))]
        items += cases.deflate64_damaged(200)
        got = eng.inflate_batch([d for _, d, _, _ in items], [w for _, _, w, _ in items], [c for _, _, _, c in items])
        for (name, data, wb, cap), (r, out) in zip(items, got):
            a, ao = CHK.oneshot(data, cap, wb)
            assert a.astuple() == r.astuple(), (lanes, name, wb, cap, a, r)
            assert ao == out, (lanes, name, "output differs")
    finally:
        eng.close()
