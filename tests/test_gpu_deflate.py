"""GPU tier, SURVEY.md §8(f) rank 4: zsb_deflate_batch() — every stream the GPU writes decodes to its input on the
oracle (the reference's own inflate) in all three wrappers, with the check value the format asks for; bad
parameters and short output buffers answer as documented; the compression ratio on the C2 plaintexts is compared
with zlib level 6; the streams also go back through the GPU inflate (a GPU-resident round trip)."""
import ctypes
import random
import zlib

import pytest

import gen
import orclib
import zsblib

pytestmark = pytest.mark.gpu
CHK = orclib.ref() or orclib.orc()


@pytest.fixture(scope="module")
def eng():
    e = zsblib.Zsb()
    yield e
    e.close()


def plaintexts():
    rnd = random.Random(3)
    return [b"", b"a", b"ab", b"abc", b"aaaa", b"abcd" * 3, bytes(1000), bytes(range(256)) * 5, rnd.randbytes(70000),
            gen.word_soup(65536, 1), gen.word_soup(200000, 2), gen.mixed_plain(150000, 3), b"ab" * 40000 + bytes(70000) + b"xyz" * 30000,
            gen.word_soup(65535, 4), gen.word_soup(65537, 5), bytes(65536 * 2), rnd.randbytes(65535 * 2 + 7), gen.word_soup(3 << 20, 6)]


def test_every_stream_decodes_on_the_reference(eng):
    items = [(p, wb, lv) for p in plaintexts() for wb in (-15, 15, 31) for lv in (6, 0, 1, 9, -1)]
    got = zsblib.deflate_batch(eng, [p for p, _, _ in items], [w for _, w, _ in items], [l for _, _, l in items])
    for (plain, wb, lv), (r, comp) in zip(items, got):
        assert r.ret == 1 and r.total_in == len(plain) and r.total_out == len(comp), (len(plain), wb, lv, r.ret)
        a, out = CHK.oneshot(comp, len(plain) + 8, wb)
        assert a.ret == 1 and out == plain and a.total_in == len(comp), (len(plain), wb, lv, a)
        if wb != -15:
            want = (zlib.adler32(plain) if wb == 15 else zlib.crc32(plain)) & 0xffffffff
            assert r.check == want and a.adler == want
        assert len(comp) <= eng.lib.zsb_deflate_bound(len(plain), wb)
    # a GPU-resident round trip: what the GPU compressed, the GPU inflates
    back = eng.inflate_batch([c for _, c in got], [w for _, w, _ in items], [len(p) + 8 for p, _, _ in items])
    for (plain, wb, lv), (r, out) in zip(items, back):
        assert r.ret == 1 and out == plain


def test_parameters_and_capacity(eng):
    p = gen.word_soup(5000, 9)
    got = zsblib.deflate_batch(eng, [p] * 6, [-15, 14, -16, 15, 31, -15], [6, 6, 6, 10, -2, 6],
                               caps=[6000, 6000, 6000, 6000, 6000, 100])
    assert [r.ret for r, _ in got] == [1, -2, -2, -2, -2, -5]
    assert got[5][0].total_out == 0


def test_ratio_on_c2_plaintexts(eng):
    plains = [gen.word_soup(65536, 1000 + i) for i in range(256)]
    got = zsblib.deflate_batch(eng, plains, [-15] * len(plains), [6] * len(plains))
    ours = sum(len(c) for _, c in got)
    ref = sum(len(zlib.compress(p, 6, wbits=-15)) for p in plains)
    print("compressed size vs zlib level 6: %.4f" % (ours / ref))
    assert ours <= 1.10 * ref, (ours, ref)
    for p, (r, c) in zip(plains[::17], got[::17]):
        assert zlib.decompress(c, -15) == p


def test_device_resident_batch(eng):
    import torch
    plains = [gen.word_soup(65536, 50 + i) for i in range(64)] + [gen.mixed_plain(300000, 7)]
    dev = torch.device("cuda", 0)
    keep, n = [], len(plains)
    arr = (zsblib.DeflateMember * n)()
    for i, p in enumerate(plains):
        ti = torch.frombuffer(bytearray(p), dtype=torch.uint8).to(dev)
        cap = eng.lib.zsb_deflate_bound(len(p), 31)
        to = torch.zeros(cap, dtype=torch.uint8, device=dev)
        keep.append((ti, to))
        arr[i].inp, arr[i].in_len, arr[i].out, arr[i].out_cap = ti.data_ptr(), len(p), to.data_ptr(), cap
        arr[i].window_bits, arr[i].level = 31, 6
    res = (zsblib.DeflateResult * n)()
    torch.cuda.synchronize()
    assert eng.lib.zsb_deflate_batch(eng.ctx, n, arr, res) == 0, eng.error()
    for i, p in enumerate(plains):
        comp = keep[i][1][:res[i].total_out].cpu().numpy().tobytes()
        assert res[i].ret == 1 and zlib.decompress(comp, 31) == p
