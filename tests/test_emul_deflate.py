"""CPU tier: the GPU deflate's device code (zsb_deflate.cuh: LZ77 tiles, Huffman construction, header, emission,
block alignment), compiled for the host by tests/emul with a warp of one lane.  Every stream it writes must decode
to the input on the oracle (the reference's own inflate where built) and on the system zlib, in all three wrappers;
the compression ratio is compared with zlib level 6 on the C2 plaintext."""
import ctypes
import random
import zlib

import cases
import gen
import orclib
import zsblib

CHK = orclib.ref() or orclib.orc()


def emul_deflate(data, wbits, level=6):
    lib = zsblib.emul().lib
    lib.emul_deflate.restype = ctypes.c_long
    lib.emul_deflate.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_int]
    cap = len(data) + len(data) // 4096 * 16 + 256
    buf = ctypes.create_string_buffer(cap)
    n = lib.emul_deflate(data, len(data), buf, cap, wbits, level)
    assert n >= 0
    return buf.raw[:n]


def plaintexts():
    rnd = random.Random(3)
    p = [b"", b"a", b"ab", b"abc", b"aaaa", b"abcd" * 3, bytes(1000), bytes(range(256)) * 5, rnd.randbytes(70000),
         gen.word_soup(65536, 1), gen.word_soup(200000, 2), gen.mixed_plain(150000, 3), b"ab" * 40000 + bytes(70000) + b"xyz" * 30000,
         gen.word_soup(65535, 4), gen.word_soup(65537, 5), bytes(65536 * 2), rnd.randbytes(65535 * 2 + 7)]
    return p


def test_streams_decode_to_their_input():
    for plain in plaintexts():
        for wb in (-15, 15, 31):
            for level in (6, 0, 1, 9):
                comp = emul_deflate(plain, wb, level)
                r, out = CHK.oneshot(comp, len(plain) + 8, wb)
                assert r.ret == 1 and out == plain and r.total_in == len(comp), (len(plain), wb, level, r)
                assert zlib.decompress(comp, wb) == plain
                if wb != -15:
                    want = zlib.adler32(plain) if wb == 15 else zlib.crc32(plain)
                    assert r.adler == (want & 0xffffffff)


def test_ratio_against_zlib_level_6():
    total, ref = 0, 0
    for seed in range(8):
        plain = gen.word_soup(65536, 700 + seed)
        total += len(emul_deflate(plain, -15))
        ref += len(zlib.compress(plain, 6, wbits=-15))
    # the matcher keeps 4 candidates per hash bucket where zlib level 6 walks chains of up to 128
    assert total <= 1.15 * ref, (total, ref)
    print("ratio vs zlib level 6: %.3f" % (total / ref))


def test_length_limits_are_repaired_to_complete_codes():
    """Skewed symbol statistics push the unlimited Huffman depths past 15 (literal/length, distance) and 7 (the
    code-length code); every code set written must still be complete."""
    rnd = random.Random(8)
    fib = [1, 1]
    while len(fib) < 40:
        fib.append(fib[-1] + fib[-2])
    plains = []
    for k in range(12):
        # byte values with Fibonacci-like counts: the deepest possible Huffman trees
        parts = [bytes([(7 * i + k) % 256]) * min(fib[i + (k % 3)], 30000) for i in range(20 + k)]
        rnd.shuffle(parts)
        plains.append(b"".join(parts)[:65536])
        plains.append(bytes(rnd.choice(b"abcdefgh" * (k + 1) + bytes(range(k, k + 40))) for _ in range(3000 + 997 * k)))
    plains.append(open("gpurun_out/deflate_bad_plain.bin", "rb").read() if __import__("os").path.exists("gpurun_out/deflate_bad_plain.bin") else b"x")
    for plain in plains:
        comp = emul_deflate(plain, -15)
        assert zlib.decompress(comp, -15) == plain
        r, out = CHK.oneshot(comp, len(plain) + 8, -15)
        assert r.ret == 1 and out == plain


def test_code_lengths_are_complete_and_within_the_limit():
    lib = zsblib.emul().lib
    lib.emul_build_lengths.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_uint, ctypes.c_void_p]
    rnd = random.Random(77)
    repaired = 0
    for trial in range(3000):
        n, maxbits = rnd.choice([(19, 7), (30, 15), (286, 15), (19, 7)])
        k = rnd.randrange(1, n + 1)
        shape = rnd.randrange(4)
        freq = [0] * n
        for s in rnd.sample(range(n), k):
            freq[s] = [rnd.randrange(1, 10), 1 << rnd.randrange(0, 24), int(1.7 ** rnd.randrange(0, 40)) + 1, rnd.randrange(1, 100000)][shape]
        f = (ctypes.c_uint32 * n)(*freq)
        lens = (ctypes.c_uint8 * n)()
        lib.emul_build_lengths(f, n, maxbits, lens)
        ls = list(lens)
        coded = [l for l in ls if l]
        assert all(l <= maxbits for l in coded) and len(coded) >= 2
        assert all(ls[s] for s in range(n) if freq[s]), "a used symbol without a code"
        assert sum(1 << (maxbits - l) for l in coded) == 1 << maxbits, (trial, n, maxbits, freq, ls)
        repaired += max(coded) == maxbits
    assert repaired > 100
