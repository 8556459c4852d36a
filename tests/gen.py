"""Seeded synthetic inputs for tests and bench.py.  No file or network access."""
import random
import zlib

import numpy as np


_VOCAB = None


def _vocab():
    global _VOCAB
    if _VOCAB is None:
        vr = np.random.default_rng(12345)
        nv = 4096
        lens = vr.integers(2, 11, nv)
        offs = np.concatenate(([0], np.cumsum(lens)))
        flat = vr.integers(97, 123, int(offs[-1]), dtype=np.uint8)
        w = 1.0 / (np.arange(nv) + 1.0)
        _VOCAB = (lens.astype(np.int64), offs[:-1].astype(np.int64), flat, w / w.sum())
    return _VOCAB


def word_soup(nbytes, seed):
    """C2-style plaintext: 4096-word vocabulary of random lowercase words (2-10
    letters), Zipf weights 1/(k+1), separators 90% space / 10% newline, 5% of the tokens replaced by
    4-32 raw random bytes.  Deterministic in (nbytes, seed)."""
    lens, offs, flat, w = _vocab()
    rng = np.random.default_rng(seed)
    ntok = nbytes // 3 + 16
    ids = rng.choice(len(lens), size=ntok, p=w)
    sep = np.where(rng.random(ntok) < 0.9, 32, 10).astype(np.uint8)
    raw = rng.random(ntok) < 0.05
    rawlen = rng.integers(4, 33, ntok)
    tl = np.where(raw, rawlen, lens[ids]) + 1            # token length incl. separator
    ends = np.cumsum(tl)
    total = int(ends[-1])
    starts = ends - tl
    # position within token for every output byte
    tok = np.repeat(np.arange(ntok), tl)
    within = np.arange(total) - starts[tok]
    src = offs[ids][tok] + within
    body = flat[np.minimum(src, len(flat) - 1)]
    rnd_bytes = rng.integers(0, 256, total, dtype=np.uint8)
    out = np.where(raw[tok], rnd_bytes, body)
    is_sep = within == (tl[tok] - 1)
    out = np.where(is_sep, sep[tok], out).astype(np.uint8)
    assert total >= nbytes
    return out[:nbytes].tobytes()


def deflate(data, level=6, wbits=-15, strategy=zlib.Z_DEFAULT_STRATEGY, memlevel=8):
    c = zlib.compressobj(level, zlib.DEFLATED, wbits, memlevel, strategy)
    return c.compress(data) + c.flush()


def mixed_plain(nbytes, seed):
    """Plaintext with runs, repeats at long distances, random and text parts."""
    rnd = random.Random(seed)
    out = bytearray()
    while len(out) < nbytes:
        k = rnd.randrange(5)
        n = rnd.randrange(1, 3000)
        if k == 0:
            out += bytes([rnd.randrange(256)]) * n
        elif k == 1:
            out += rnd.randbytes(n)
        elif k == 2 and len(out) > 10:
            d = rnd.randrange(1, min(len(out), 40000) + 1)
            for _ in range(n):
                out.append(out[-d])
        elif k == 3:
            out += word_soup(n, rnd.randrange(1 << 30))
        else:
            pat = rnd.randbytes(rnd.randrange(1, 9))
            out += pat * (n // len(pat) + 1)
    return bytes(out[:nbytes])


def corrupt(data, seed):
    """Return a damaged copy: bit flip, truncation, byte overwrite or trailing garbage."""
    rnd = random.Random(seed)
    b = bytearray(data)
    k = rnd.randrange(4)
    if k == 0 and b:
        i = rnd.randrange(len(b))
        b[i] ^= 1 << rnd.randrange(8)
    elif k == 1 and b:
        del b[rnd.randrange(len(b)):]
    elif k == 2 and b:
        i = rnd.randrange(len(b))
        b[i] = rnd.randrange(256)
    else:
        b += rnd.randbytes(rnd.randrange(1, 9))
    return bytes(b)


class BitWriter:
    """LSB-first bit packer for hand-made deflate streams."""

    def __init__(self):
        self.acc = 0
        self.n = 0
        self.out = bytearray()

    def bits(self, value, nbits):
        self.acc |= (value & ((1 << nbits) - 1)) << self.n
        self.n += nbits
        while self.n >= 8:
            self.out.append(self.acc & 0xff)
            self.acc >>= 8
            self.n -= 8

    def code(self, code, nbits):
        """Huffman code given MSB-first."""
        for i in range(nbits - 1, -1, -1):
            self.bits((code >> i) & 1, 1)

    def align(self):
        if self.n:
            self.bits(0, 8 - self.n)

    def done(self):
        self.align()
        return bytes(self.out)


def fixed_lit(bw, sym):
    if sym < 144:
        bw.code(0x30 + sym, 8)
    elif sym < 256:
        bw.code(0x190 + sym - 144, 9)
    elif sym < 280:
        bw.code(sym - 256, 7)
    else:
        bw.code(0xc0 + sym - 280, 8)
