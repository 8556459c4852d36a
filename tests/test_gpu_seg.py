"""GPU tier: large members decoded in segments (zlib-streams_b200/csrc/zsb_seg.cuh) — block starts found inside a member,
the stretches between them decoded by a lane each, chained, resolved — against the oracle, record for record and byte
for byte: valid, damaged and truncated members, all three wrappers and Deflate64, members whose chain cannot hold
(fixed / stored blocks only, short capacity) and therefore take the serial fallback, with small segments (many
boundaries per member) and with the default geometry."""
import random
import zlib

import pytest

import gen
import orclib
import zsblib

pytestmark = pytest.mark.gpu
CHK = orclib.ref() or orclib.orc()


def members():
    import d64enc
    rnd = random.Random(21)
    soup = b"".join(gen.word_soup(65536, 140 + i) for i in range(12))
    out = []
    for wb in (-15, 15, 31):
        out.append((gen.deflate(soup, 6, wb), wb, len(soup) + 16))
    out.append((gen.deflate(soup, 1, -15), -15, len(soup) + 16))
    out.append((gen.deflate(soup, 9, 31), 31, len(soup)))                      # exact capacity
    out.append((gen.deflate(soup, 6, 15), 15, len(soup) - 1))                  # one byte short: serial verdict
    mixed = gen.mixed_plain(900000, 5)
    out.append((gen.deflate(mixed, 6, 15), 15, len(mixed) + 16))
    out.append((gen.deflate(soup[:400000], 6, -15, zlib.Z_FIXED), -15, 400016))
    rb = rnd.randbytes(300000)
    out.append((gen.deflate(rb, 0, 15), 15, 300016))
    plain = soup[:200000] + soup[:65536] + bytes(90000) + soup[300000:350000] * 3
    out.append((d64enc.compress(plain, True, 3000, ("dynamic", "dynamic", "fixed", "stored")), -16, len(plain) + 16))
    import test_emul_seg
    fs, fs_plain = test_emul_seg.stream_with_a_false_start()                  # a false start inside a stored block
    out.append((fs, -15, len(fs_plain)))
    base = list(out[:4]) + [out[6], out[9]]
    for comp, wb, cap in base:                                                # damage and truncation
        for k in range(6):
            bad = bytearray(comp)
            at = rnd.randrange(len(bad))
            bad[at] ^= 1 << rnd.randrange(8)
            out.append((bytes(bad), wb, cap + 64))
        for cut in (len(comp) // 3, len(comp) - 6, len(comp) - 1):
            out.append((comp[:cut], wb, cap + 64))
        out.append((comp + b"\0" * 3000 + comp[:2000], wb, cap + 64))
    return out


@pytest.mark.parametrize("seg_bytes,seg_min,arena", [(8192, 32768, 0), (32768, 131072, 0), (2048, 16384, 0), (0, 65536, 0), (8192, 32768, 6 << 20)])
def test_segmented_members_match_the_oracle(seg_bytes, seg_min, arena):
    import torch
    import batchlib
    eng = zsblib.Zsb(devices=[0])
    try:
        eng.set_option("seg_bytes", seg_bytes)
        eng.set_option("seg_min", seg_min)
        if arena:                                     # a small sequence arena: the segmented members go in several waves
            eng.set_option("seq_arena_bytes", arena)
        small = [(gen.deflate(gen.word_soup(3000 + 100 * i, i), 6, -15), -15, 3000 + 100 * i + 8) for i in range(40)]
        distinct = [(c, wb, bytes(cap)) for c, wb, cap in members() + small]
        w = {"distinct": distinct, "tiles": 3, "desc": "large members in segments"}
        dev = torch.device("cuda", 0)
        batch = batchlib.DeviceBatch(w, dev).to_device()
        stream = torch.cuda.Stream(dev)
        for segment in (1, 0, 1):
            eng.set_option("segment", segment)
            batch.clear_outputs()
            torch.cuda.synchronize()                  # (the clears run on torch's current stream, the step on `stream`)
            batch.step(eng, stream)
            stream.synchronize()
            bad = batch.verify()
            assert not bad, (segment, bad[:3])
    finally:
        eng.close()


def test_device_batch_of_every_test_stream():
    """All valid and damaged test streams (tests/cases.py) as one device-resident batch of several thousand members,
    with the default geometry, with tiny segments, and with segments off."""
    import torch
    import batchlib
    import cases
    distinct = [(d, wb, bytes(len(p) + 16)) for _, d, wb, p in cases.valid_streams()]
    distinct += [(d, wb, bytes(c)) for _, d, wb, c in cases.damaged_streams(150)]
    w = {"distinct": distinct, "tiles": -(-9000 // len(distinct)), "desc": "valid + damaged streams"}
    dev = torch.device("cuda", 0)
    batch = batchlib.DeviceBatch(w, dev).to_device()
    stream = torch.cuda.Stream(dev)
    eng = zsblib.Zsb(devices=[0])
    try:
        # (the last: a small sequence arena — the lane path decodes in slices)
        for opts in ((), (("seg_bytes", 4096), ("seg_min", 16384)), (("segment", 0),), (("seq_arena_bytes", 3 << 20),)):
            for k, v in opts:
                eng.set_option(k, v)
            batch.clear_outputs()
            torch.cuda.synchronize()                  # (the clears run on torch's current stream, the step on `stream`)
            batch.step(eng, stream)
            stream.synchronize()
            bad = batch.verify()
            assert not bad, (opts, bad[:3])
    finally:
        eng.close()


def test_host_batch_with_large_members():
    """Host buffers: the large members are staged whole and decoded in segments, the rest goes through the chunk
    pipeline; every record and byte as the oracle's, in the caller's order."""
    eng = zsblib.Zsb()
    try:
        small = [(gen.deflate(gen.word_soup(3000 + 100 * i, i), 6, -15), -15, 3000 + 100 * i + 8) for i in range(300)]
        big = members()
        items = []
        for i in range(max(len(small), len(big))):            # interleaved
            if i < len(big):
                items.append(big[i])
            if i < len(small):
                items.append(small[i])
        for opts in ((), (("seg_bytes", 8192), ("seg_min", 32768)), (("segment", 0),)):
            for k, v in opts:
                eng.set_option(k, v)
            got = eng.inflate_batch([d for d, _, _ in items], [w for _, w, _ in items], [c for _, _, c in items])
            for j, ((data, wb, cap), (r, out)) in enumerate(zip(items, got)):
                a, ao = CHK.oneshot(data, cap, wb)
                assert a.astuple() == r.astuple(), (opts, j, wb, cap, a, r)
                assert ao == out, (opts, j, "output differs")
    finally:
        eng.close()


def test_checksums_of_large_members_in_parts():
    """Wrapped members of 2 MiB of output or more have their Adler-32 / CRC-32 computed in parts of 1 MiB by a CTA each
    and combined (k_check_parts, k_check): sizes around the part boundaries, both wrappers, compressible and stored
    data, good and damaged trailers — records (check value, ret, msg, total_in) as the oracle's."""
    import torch
    import batchlib
    rnd = random.Random(8)
    distinct = []
    for size in (2 << 20, (2 << 20) + 1, (3 << 20) - 1, 3 << 20, (5 << 20) + 12345, (1 << 20) + 77):
        stored = rnd.randbytes(size)
        soup = (gen.word_soup(65536, size & 1023) * (size // 65536 + 1))[:size]
        for wb in (15, 31):
            for plain, level in ((stored, 0), (soup, 6)):
                comp = gen.deflate(plain, level, wb)
                distinct.append((comp, wb, plain))
                bad = bytearray(comp)
                bad[-1] ^= 0x40                                   # the trailer: "incorrect data check" / "incorrect length check"
                distinct.append((bytes(bad), wb, plain))
                distinct.append((comp[:-3], wb, plain))           # the trailer cut short
    w = {"distinct": distinct, "tiles": 1, "desc": "large wrapped members"}
    dev = torch.device("cuda", 0)
    batch = batchlib.DeviceBatch(w, dev).to_device()
    eng = zsblib.Zsb(devices=[0])
    try:
        st = torch.cuda.current_stream(dev)
        for segment in (1, 0):
            eng.set_option("segment", segment)
            batch.clear_outputs()
            batch.step(eng, st)
            torch.cuda.synchronize()
            bad = batch.verify()
            assert not bad, (segment, bad[:3])
    finally:
        eng.close()
