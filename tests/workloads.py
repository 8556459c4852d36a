"""TEST TOOL: synthetic batches shaped like BASELINE.json's configs C3, C4 and C5.

A workload is K DISTINCT seeded members (compressed bytes, windowBits, plaintext) and a member count
n = tiles * K: member i is distinct member i % K, so a batch is the distinct blob tiled to distinct HBM
addresses and the expected output of every tile is the same byte string (one comparison per tile checks
all n members bit for bit).  Expected per-member results come from the oracle on the K distinct members.
Nothing here is used by the product.
"""
import os
import random
import zlib
from concurrent.futures import ProcessPoolExecutor

import numpy as np

import gen

_POOL = None


def _pool_block(seed):
    return gen.word_soup(65536, seed)


def soup_pool(blocks=128):
    """`blocks` word-soup blocks of 64 KiB (seeds 5000+i); large plaintexts are assembled from them."""
    global _POOL
    if _POOL is None or len(_POOL) < blocks:
        with ProcessPoolExecutor(min(os.cpu_count() or 1, 32)) as ex:
            _POOL = list(ex.map(_pool_block, [5000 + i for i in range(blocks)], chunksize=4))
    return _POOL


def soup_bytes(nbytes, rnd, pool):
    """nbytes of word soup: whole pool blocks in a seeded order, the last one cut."""
    parts, left = [], nbytes
    while left > 0:
        b = pool[rnd.randrange(len(pool))]
        parts.append(b[:left])
        left -= len(parts[-1])
    return b"".join(parts)


def log_uniform(rnd, lo, hi):
    return int(round(lo * (hi / lo) ** rnd.random()))


# ---------------------------------------------------------------------------------------------------
# C2: 64 KiB word-soup members, raw deflate, zlib level 6 (the headline configuration)
# ---------------------------------------------------------------------------------------------------
def _c2_member(seed):
    plain = gen.word_soup(65536, seed)
    return gen.deflate(plain, 6, -15), -15, plain


def c2(distinct=1024, tiles=64, seed_base=1000):
    with ProcessPoolExecutor(min(os.cpu_count() or 1, 32)) as ex:
        mem = list(ex.map(_c2_member, [seed_base + i for i in range(distinct)], chunksize=16))
    return {"name": "C2", "desc": "%d x 64 KiB synthetic raw-deflate (zlib level 6, word-soup) members" % (distinct * tiles),
            "distinct": mem, "tiles": tiles}


# ---------------------------------------------------------------------------------------------------
# C3: Deflate64 members with a 64 KiB window and long matches (tests/d64enc.py, the test-only encoder)
# ---------------------------------------------------------------------------------------------------
def _c3_member(args):
    seed, nbytes, pool_blocks = args
    import d64enc
    rnd = random.Random(seed)
    pool = [gen.word_soup(65536, 5000 + rnd.randrange(64)) for _ in range(pool_blocks)]
    parts, total = [], 0
    while total < nbytes:
        kind = rnd.randrange(4)
        if kind == 0:                                   # word soup (as C2)
            p = soup_bytes(rnd.randrange(32768, 196608), rnd, pool)
        elif kind == 1:                                 # a 48-64 KiB block repeated: distances 32769..65536
            blk = soup_bytes(rnd.randrange(49152, 65537), rnd, pool)
            p = blk * rnd.randrange(3, 9)
        elif kind == 2:                                 # long runs / short periods: matches up to 65538
            unit = bytes(rnd.getrandbits(8) for _ in range(rnd.choice((1, 2, 3, 7, 64, 300))))
            p = unit * (rnd.randrange(70000, 400000) // len(unit))
        else:                                           # a far copy of something older than 32 KiB
            p = soup_bytes(rnd.randrange(40000, 60000), rnd, pool)
            p = p + bytes(rnd.getrandbits(8) for _ in range(200)) + p
        parts.append(p)
        total += len(p)
    plain = b"".join(parts)[:nbytes]
    comp = d64enc.compress(plain, True, 8000, ("dynamic", "dynamic", "fixed", "stored"))
    return comp, -16, plain


def c3(distinct=8, tiles=512, member_bytes=4 << 20):
    with ProcessPoolExecutor(min(os.cpu_count() or 1, distinct)) as ex:
        mem = list(ex.map(_c3_member, [(3000 + i, member_bytes, 4) for i in range(distinct)]))
    return {"name": "C3", "desc": "%d x %d MiB synthetic Deflate64 members (64 KiB window, long matches)" % (
        distinct * tiles, member_bytes >> 20), "distinct": mem, "tiles": tiles}


# ---------------------------------------------------------------------------------------------------
# C4: gzip / zlib wrapped members, 1 KiB - 16 MiB log-uniform, dynamic + fixed + stored blocks, checksums
# verified; about 1 % of the distinct members are damaged (trailer flips, truncation, bit flips)
# ---------------------------------------------------------------------------------------------------
def _c4_member(args):
    i, seed, nbytes = args
    rnd = random.Random(seed)
    pool = [gen.word_soup(65536, 5000 + rnd.randrange(64)) for _ in range(3)]
    wb = 31 if i % 2 == 0 else 15
    third = i % 3
    if third == 0:                                      # word soup, level 6: dynamic blocks
        plain = soup_bytes(nbytes, rnd, pool)
        comp = gen.deflate(plain, 6, wb)
    elif third == 1:                                    # Z_FIXED: fixed blocks
        plain = soup_bytes(nbytes, rnd, pool)
        comp = gen.deflate(plain, 6, wb, zlib.Z_FIXED)
    else:                                               # random bytes: stored blocks (level 0 or what level 6 makes of them)
        plain = np.random.default_rng(seed).integers(0, 256, nbytes, dtype=np.uint8).tobytes()
        comp = gen.deflate(plain, 0 if i % 2 else 6, wb)
    if i % 97 == 5:                                     # ~1 %: damaged
        k = (i // 97) % 4
        b = bytearray(comp)
        if k == 0:
            b[-1] ^= 0x40                               # last trailer byte (ISIZE for gzip, Adler-32 for zlib)
        elif k == 1:
            b[-6 if wb == 31 else -3] ^= 0x01           # the data check
        elif k == 2:
            del b[len(b) * 2 // 3:]                     # truncated
        else:
            b[len(b) // 2] ^= 0x10                      # somewhere in the middle
        comp = bytes(b)
    return comp, wb, plain


def c4(distinct=512, tiles=32, lo=1024, hi=16 << 20):
    rnd = random.Random(44)
    sizes = [log_uniform(rnd, lo, hi) for _ in range(distinct)]
    with ProcessPoolExecutor(min(os.cpu_count() or 1, 32)) as ex:
        mem = list(ex.map(_c4_member, [(i, 4000 + i, sizes[i]) for i in range(distinct)], chunksize=2))
    return {"name": "C4", "desc": "%d gzip/zlib members, %d B - %d MiB log-uniform, dynamic + fixed + stored blocks, "
            "checksums verified, ~1%% damaged" % (distinct * tiles, lo, hi >> 20), "distinct": mem, "tiles": tiles}


# ---------------------------------------------------------------------------------------------------
# C5: ZIP-like mix, 80 % raw deflate / 20 % Deflate64, 1 KiB - 256 KiB log-uniform
# ---------------------------------------------------------------------------------------------------
def _c5_member(args):
    i, seed, nbytes = args
    rnd = random.Random(seed)
    pool = [gen.word_soup(65536, 5000 + rnd.randrange(64)) for _ in range(2)]
    plain = soup_bytes(nbytes, rnd, pool)
    if i % 5 == 4:
        import d64enc
        return d64enc.compress(plain, True, 6000, ("dynamic",)), -16, plain
    return gen.deflate(plain, 6, -15), -15, plain


def c5(distinct=2048, tiles=512, lo=1024, hi=256 << 10):
    rnd = random.Random(55)
    sizes = [log_uniform(rnd, lo, hi) for _ in range(distinct)]
    with ProcessPoolExecutor(min(os.cpu_count() or 1, 32)) as ex:
        mem = list(ex.map(_c5_member, [(i, 6000 + i, sizes[i]) for i in range(distinct)], chunksize=8))
    return {"name": "C5", "desc": "%d-member ZIP-like mix (80%% raw deflate, 20%% Deflate64), %d B - %d KiB log-uniform" % (
        distinct * tiles, lo, hi >> 10), "distinct": mem, "tiles": tiles}


BUILDERS = {"C2": c2, "C3": c3, "C4": c4, "C5": c5}
