"""CPU tier: the segmented path of large members (zlib-streams_b200/csrc/zsb_seg.cuh) through the host emulation —
block starts found near multiples of `segb` compressed bytes, the segments decoded on their own, chained, resolved —
must give the oracle's record and bytes for valid, damaged and truncated streams, whether the chain holds or the
member falls back to the serial decoder."""
import random
import zlib

import pytest

import cases
import gen
import orclib
import zsblib

CHK = orclib.ref() or orclib.orc()


def check(name, data, wb, cap, segb, want_chain=None, seg_q=None):
    a, ao = CHK.oneshot(data, cap, wb)
    r, out, info = zsblib.emul().oneshot_seg(data, cap, wb, segb, seg_q)
    assert a.astuple() == r.astuple(), (name, wb, cap, segb, info, a, r)
    assert ao == out, (name, "output differs", info)
    if want_chain is not None:
        assert bool(info[1]) == want_chain, (name, info)
    return info


def big_streams():
    rnd = random.Random(11)
    soup = b"".join(gen.word_soup(65536, 40 + i) for i in range(6))
    out = []
    for wb in (-15, 15, 31):
        out.append(("soup6_wb%d" % wb, gen.deflate(soup, 6, wb), wb, soup))
    out.append(("soup1", gen.deflate(soup, 1, -15), -15, soup))
    out.append(("soup9", gen.deflate(soup, 9, 31), 31, soup))
    mixed = gen.mixed_plain(300000, 5)
    out.append(("mixed6", gen.deflate(mixed, 6, 15), 15, mixed))
    fixed = gen.deflate(soup[:200000], 6, -15, zlib.Z_FIXED)
    out.append(("fixed", fixed, -15, soup[:200000]))
    rb = rnd.randbytes(200000)
    out.append(("stored", gen.deflate(rb, 0, 15), 15, rb))
    return out


def test_valid_members_chain_and_match():
    chained = 0
    for name, data, wb, plain in big_streams():
        for segb in (4096, 16384):
            info = check(name, data, wb, len(plain) + 16, segb)
            if segb == 16384 and (name.startswith("soup") or name.startswith("mixed") or name == "fixed"):
                assert info[1] == 1 and info[0] >= 1, (name, segb, info)     # starts found, the chain holds
            if name == "stored":
                assert info[0] == 0, info                                      # nothing to find in stored blocks
            chained += info[1]
    assert chained >= 7


def test_exact_capacity_and_short_capacity():
    name, data, wb, plain = big_streams()[0]
    check(name, data, wb, len(plain), 8192, want_chain=True)
    check(name, data, wb, len(plain) - 1, 8192, want_chain=False)        # output beyond out_cap: the serial decoder's verdict
    check(name, data, wb, 100000, 8192, want_chain=False)
    check(name, data, wb, 0, 8192)


def test_damaged_and_truncated_members():
    rnd = random.Random(5)
    for name, data, wb, plain in big_streams()[:6]:
        for k in range(12):
            bad = bytearray(data)
            at = rnd.randrange(len(bad))
            bad[at] ^= 1 << rnd.randrange(8)
            check("%s_flip%d" % (name, at), bytes(bad), wb, len(plain) + 64, 8192)
        for cut in (len(data) // 3, len(data) // 2, len(data) - 7, len(data) - 1):
            check("%s_cut%d" % (name, cut), data[:cut], wb, len(plain) + 64, 8192)
        check(name + "_trailing", data + b"\0" * 5000 + data[:3000], wb, len(plain) + 64, 8192)


def test_deflate64_members():
    import d64enc
    rnd = random.Random(3)
    pool = [gen.word_soup(65536, 70 + i) for i in range(3)]
    plain = b"".join(pool) + pool[0] + bytes(90000) + pool[1][:50000] * 3
    comp = d64enc.compress(plain, True, 3000, ("dynamic", "dynamic", "fixed", "stored"))
    info = check("d64_mixed", comp, -16, len(plain) + 16, 8192)
    assert info[1] == 1 and info[0] >= 2, info
    check("d64_mixed_small_segments", comp, -16, len(plain) + 16, 1024)


def test_small_history_and_full_arena_fall_back():
    # a second block start before 64 KiB of output exist: the chain must not trust the assumed history
    plain = gen.word_soup(50000, 9)
    comp = gen.deflate(plain, 6, -15)
    check("short_history", comp, -15, len(plain) + 16, 2048)
    # an arena range far too small for a segment's sequences: EV_SEGFULL, serial fallback
    name, data, wb, p = big_streams()[0]
    check(name, data, wb, len(p) + 16, 8192, want_chain=False, seg_q=64)


def test_every_fixture_and_probe_vector_survives_tiny_segments():
    for name, data, meta in cases.golden_fixtures():
        check(name, data, -16, meta["out_len"] + 16, 512)
    for name, data, wb, cap, _ in cases.probe_vectors():
        check(name, data, wb, cap, 256)


def stream_with_a_false_start():
    """A valid raw deflate stream whose stored block carries the bytes of ANOTHER deflate stream: the search finds that
    stream's dynamic header inside the stored data — a false start, like the header-shaped bit patterns that turn up by
    chance about once per 8 MB searched."""
    p1 = b"".join(gen.word_soup(65536, 300 + i) for i in range(5))
    inner = gen.deflate(gen.word_soup(100000, 77), 6, -15)
    assert len(inner) < 65536
    p3 = b"".join(gen.word_soup(65536, 400 + i) for i in range(6))
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    piece1 = c.compress(p1) + c.flush(zlib.Z_SYNC_FLUSH)                  # ends byte-aligned, BFINAL = 0 everywhere
    n = len(inner)
    piece2 = bytes([0, n & 255, n >> 8, (n & 255) ^ 255, (n >> 8) ^ 255]) + inner
    piece3 = gen.deflate(p3, 6, -15)
    return piece1 + piece2 + piece3, p1 + inner + p3


def test_false_start_is_skipped_by_the_chain():
    data, plain = stream_with_a_false_start()
    assert CHK.oneshot(data, len(plain), -15)[1] == plain
    for segb in (8192, 16384, 32768):
        info = check("false_start", data, -15, len(plain), segb)
        assert info[0] >= 3, info
    assert check("false_start", data, -15, len(plain), 8192)[1] == 1       # small segments: the chain holds around it
