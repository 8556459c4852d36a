"""Test-side tool for SURVEY.md §8 a6'(ii): code sets at the reference's table-space limits.

The reference's two-level table builder counts the entries it would allocate (`used`,
/root/reference/src/inftrees.c:211-228,275-297) and rejects a code set whose count exceeds ENOUGH_LENS = 852
/ ENOUGH_DISTS = 592 — and, in Deflate64 mode only, 851 / 593 (`enough_lens`, `enough_dists`, :211-212).  So a
literal/length set that needs exactly 852 entries is accepted by windowBits -15 and rejected by -16, and a
32-symbol distance set that needs 594 is rejected by -16 (a 30-symbol set can need at most 592).

`table_space()` restates that count from the per-length code counts alone (the canonical code is fixed by
them); `search()` climbs over complete count vectors (the method of zlib's examples/enough.c, randomised)
to reach a wanted count; `header_stream()` emits a one-block raw deflate stream — a dynamic header carrying
the two sets followed by the end-of-block code — that drives the set through any decoder.  The vectors the
search found are stored below (VECTORS) so the tests do not depend on the search's luck; run this file to
regenerate them.  The compiled reference confirms every one (tests/test_tablespace.py)."""
import random

import d64enc
import gen


def table_space(count, root):
    """Entries the reference allocates for a COMPLETE code with count[l] codes of length l (l = 1..15)."""
    count = list(count) + [0] * (17 - len(count))
    mx = max(l for l in range(1, 16) if count[l])
    mn = min(l for l in range(1, 16) if count[l])
    root = min(max(root, mn), mx)
    cnt = count[:]
    used = 1 << root
    mask = used - 1
    huff, length, drop, low = 0, mn, 0, -1
    while True:
        incr = 1 << (length - 1)               # backwards increment of the bit-reversed code
        while huff & incr:
            incr >>= 1
        huff = (huff & (incr - 1)) + incr if incr else 0
        cnt[length] -= 1
        if cnt[length] == 0:
            if length == mx:
                break
            length += 1
            while cnt[length] == 0:
                length += 1
        if length > root and (huff & mask) != low:
            if drop == 0:
                drop = root
            curr = length - drop
            left = 1 << curr
            while curr + drop < mx:
                left -= cnt[curr + drop]
                if left <= 0:
                    break
                curr += 1
                left <<= 1
            used += 1 << curr
            low = huff & mask
    return used


def kraft_complete(count):
    return sum(c << (15 - l) for l, c in enumerate(count) if l) == 1 << 15


def search(nsym, root, want, seed=1, tries=200000):
    """A count vector (index = length 0..15) of a complete code over at most nsym symbols whose table needs
    exactly `want` entries; None when the climb does not get there."""
    rnd = random.Random(seed)

    def random_start():
        c = [0] * 16
        c[1] = 2
        n = 2
        while n < nsym and rnd.random() < 0.995:
            ls = [l for l in range(1, 15) if c[l]]
            l = rnd.choice(ls)
            c[l] -= 1
            c[l + 1] += 2
            n += 1
        return c

    def moves(c):
        n = sum(c)
        out = []
        for l in range(1, 15):
            if c[l] and n < nsym:
                out.append((l, "split"))
            if c[l + 1] >= 2:
                out.append((l, "merge"))
        return out

    def apply(c, mv):
        l, k = mv
        d = c[:]
        if k == "split":
            d[l] -= 1
            d[l + 1] += 2
        else:
            d[l] += 1
            d[l + 1] -= 2
        return d

    best = None
    for _ in range(tries):
        c = random_start()
        u = table_space(c, root)
        for _ in range(400):
            if u == want:
                return c
            mv = moves(c)
            if not mv:
                break
            d = apply(c, rnd.choice(mv))
            if rnd.random() < 0.5:            # two-step moves keep the symbol count where it is
                mv2 = moves(d)
                if mv2:
                    d = apply(d, rnd.choice(mv2))
            v = table_space(d, root)
            if abs(v - want) <= abs(u - want) or rnd.random() < 0.02:
                c, u = d, v
            if best is None or abs(u - want) < best:
                best = abs(u - want)
    return None


def lengths_from_counts(count, nsym, must_have=()):
    """Code lengths for symbols 0..nsym-1 with count[l] codes of length l: the symbols in `must_have` get
    codes first (the longest ones), the remaining codes go to the lowest symbols."""
    pool = []
    for l in range(15, 0, -1):
        pool += [l] * count[l]
    assert len(pool) <= nsym
    lens = [0] * nsym
    order = list(must_have) + [s for s in range(nsym) if s not in must_have]
    for s, l in zip(order, pool):
        lens[s] = l
    return lens


def header_stream(ll, dl):
    """BFINAL=1 dynamic block with literal/length lengths ll (>= 257 of them, ll[256] != 0) and distance
    lengths dl, followed by the end-of-block code; returns (bytes, bit position just behind the header)."""
    bw = gen.BitWriter()
    bw.bits(1, 1)
    bw.bits(2, 2)
    nlen, ndist = len(ll), len(dl)
    seq = d64enc.rle_lengths(list(ll) + list(dl))
    cf = [0] * 19
    for s, _, _ in seq:
        cf[s] += 1
    cl = d64enc.huff_lengths(cf, 7)
    ncode = 19
    while ncode > 4 and cl[d64enc.ORDER[ncode - 1]] == 0:
        ncode -= 1
    bw.bits(nlen - 257, 5)
    bw.bits(ndist - 1, 5)
    bw.bits(ncode - 4, 4)
    for i in range(ncode):
        bw.bits(cl[d64enc.ORDER[i]], 3)
    cc = d64enc.canonical(cl)
    for s, x, xb in seq:
        bw.code(cc[s], cl[s])
        if xb:
            bw.bits(x, xb)
    lc = d64enc.canonical(list(ll))
    bw.code(lc[256], ll[256])
    return bw.done()


# count vectors found by search() (index = code length 0..15), keyed by (symbols, root, table entries)
VECTORS = {
    (286, 9, 852): [0, 1, 1, 1, 0, 0, 0, 0, 0, 0, 117, 1, 17, 17, 1, 130],     # the reference's ENOUGH_LENS
    (286, 9, 850): [0, 1, 1, 1, 1, 1, 0, 0, 0, 0, 19, 1, 1, 33, 33, 194],      # 851 cannot occur: counts are even
    (30, 6, 592): [0, 1, 1, 1, 1, 0, 0, 1, 9, 9, 1, 1, 1, 1, 1, 2],            # ENOUGH_DISTS
    (32, 6, 594): [0, 1, 1, 1, 1, 0, 0, 3, 1, 17, 1, 1, 1, 1, 1, 2],           # ENOUGH_DISTS_9 (593 cannot occur)
    (32, 6, 592): [0, 1, 1, 1, 1, 1, 0, 1, 1, 1, 17, 1, 1, 1, 1, 2],
}


def litlen_stream(count, dist_lens=(1, 1)):
    ll = lengths_from_counts(count, 286, must_have=(256,))
    nlen = max(257, max(s for s in range(286) if ll[s]) + 1)
    return header_stream(ll[:nlen], list(dist_lens))


def dist_stream(count, nsym):
    ll = [1] + [0] * 255 + [1]                 # a complete two-symbol literal/length code: 0 and end-of-block
    dl = lengths_from_counts(count, nsym)
    nd = max(s for s in range(nsym) if dl[s]) + 1
    return header_stream(ll, dl[:nd])


def probe_cases():
    """(name, data, wbits, out_cap) for every stored vector in both modes."""
    out = []
    for (nsym, root, used), count in sorted(VECTORS.items()):
        if root == 9:
            data = litlen_stream(count)
        else:
            data = dist_stream(count, nsym)
        for wb in (-15, -16):
            out.append(("tabspace_%dsym_root%d_%d_wb%d" % (nsym, root, used, wb), data, wb, 64))
    return out


if __name__ == "__main__":
    wanted = [(286, 9, 852), (286, 9, 851), (286, 9, 850), (30, 6, 592), (30, 6, 591), (32, 6, 592), (32, 6, 593), (32, 6, 594)]
    for nsym, root, used in wanted:
        c = None
        for seed in range(1, 40):
            c = search(nsym, root, used, seed=seed, tries=300)
            if c:
                break
        print("    (%d, %d, %d): %r," % (nsym, root, used, c))
