"""The reference's own callers (tests/callers.py), compiled unmodified.  CPU tier: the builds linked against the
reference reproduce the golden hashes (the harness is sound).  GPU tier: the SAME sources linked against
libzsb.so give the same files / the same packed result words — the drop-in boundary exercised by the reference's
code rather than by ctypes."""
import json
import os

import pytest

import callers
import cases

GOLD = json.load(open(os.path.join(cases.GOLD, "ref_data_sha256.json")))
needs_ref_builds = pytest.mark.skipif(not callers.built("ref"), reason="oracle/_ref callers not built")
needs_zsb_builds = pytest.mark.skipif(not callers.built("zsb"), reason="oracle/_ref callers not built")


def glue_jobs():
    import gen
    jobs = [("inflate9", "_init_raw", d, n) for n, d, m in cases.golden_fixtures()]
    p = gen.mixed_plain(300000, 8)
    jobs += [("inflate", "_init", gen.deflate(p, 6, 15), "zlib300k"), ("inflate", "_init_raw", gen.deflate(p, 6, -15), "raw300k"),
             ("inflate", "_init_gzip", gen.deflate(p, 6, 31), "gzip300k"), ("inflate9", "_init", gen.deflate(p, 6, -15), "raw_as_d64"),
             ("inflate", "_init", gen.deflate(p, 6, 31), "gzip_as_zlib"), ("inflate", "_init_gzip", gen.corrupt(gen.deflate(p, 6, 31), 5), "gzip_damaged")]
    return jobs


@needs_ref_builds
def test_reference_callers_on_the_reference_reproduce_the_golden_hashes():
    for name, data, m in cases.golden_fixtures():
        for binary in ("payload_decompress", "payload_decompress_nowindow"):
            rc, sha, n = callers.run_binary(binary, "ref", data)
            assert (rc, sha, n) == (0, m["out_sha256"], m["out_len"]), (binary, name)
    g = callers.Glue("ref")
    for fam, init, data, name in glue_jobs()[:len(cases.golden_fixtures())]:
        rc, words, consumed, out, end = g.replay(fam, init, data)
        assert rc == 0 and end == 0 and len(out) == GOLD[name]["out_len"], name


@pytest.mark.gpu
@needs_zsb_builds
def test_reference_binaries_linked_against_libzsb():
    """test/payload_decompress.c and payload_decompress_nowindow.c of the reference, unmodified, -lzsb."""
    for name, data, m in cases.golden_fixtures():
        for binary in ("payload_decompress", "payload_decompress_nowindow"):
            rc, sha, n = callers.run_binary(binary, "zsb", data)
            assert (rc, sha, n) == (0, m["out_sha256"], m["out_len"]), (binary, name, rc, n)
    # a damaged payload fails the same way in both builds
    import gen
    bad = gen.corrupt(cases.golden_fixtures()[0][1], 3)
    assert callers.run_binary("payload_decompress", "zsb", bad)[0] == callers.run_binary("payload_decompress", "ref", bad)[0]


@pytest.mark.gpu
@needs_zsb_builds
def test_wasm_glue_linked_against_libzsb_returns_the_reference_words():
    """The WASM glue built natively: the JS driver loop's packed (code << 24 | produced) words, last_consumed values
    and output bytes are the same with libzsb.so underneath as with the reference underneath."""
    a, b = callers.Glue("ref"), callers.Glue("zsb")
    for fam, init, data, name in glue_jobs():
        ra, rb = a.replay(fam, init, data), b.replay(fam, init, data)
        assert ra == rb, (name, ra[:3], rb[:3])
