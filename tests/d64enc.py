"""TEST TOOL: a small Deflate64 (and plain deflate) encoder.

The reference ships no Deflate64 encoder that can run here (its generator needs 7-Zip, SURVEY.md §7 hard
part 5) and its fixtures never use length code 285's 16 extra bits or distance codes 30/31 in dynamic
blocks.  This encoder produces such streams; every stream it makes is admitted to a test only after the
reference build (oracle/_ref, windowBits -16) decodes it back to the plaintext (tests/test_deflate64.py).

Format facts restated from the reference's tables (src/inftrees.c:73-87): in Deflate64 length code 285
is base 3 + 16 extra bits (lengths 3..65538), distance codes 30/31 are 32769 + 14 bits and 49153 + 14
bits, the window is 64 KiB; everything else is RFC 1951.
"""
import heapq

from gen import BitWriter

LBASE = [3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227]
LEXT = [0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5]
DBASE = [1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073,
         4097, 6145, 8193, 12289, 16385, 24577, 32769, 49153]
DEXT = [0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13, 14, 14]
ORDER = [16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15]


def length_code(n, d64, force285=False):
    """-> (symbol, extra_value, extra_bits)"""
    if d64 and (n > 258 or force285 or n == 258):
        return 285, n - 3, 16
    if n == 258:
        return 285, 0, 0
    for c in range(27, -1, -1):
        if n >= LBASE[c]:
            return 257 + c, n - LBASE[c], LEXT[c]
    raise ValueError(n)


def dist_code(dist, d64):
    top = 32 if d64 else 30
    for c in range(top - 1, -1, -1):
        if dist >= DBASE[c]:
            return c, dist - DBASE[c], DEXT[c]
    raise ValueError(dist)


def lz77(data, d64=True, max_chain=16, min_len=3):
    """Greedy LZ77: list of int literals / (length, distance) tuples."""
    n = len(data)
    window = 65536 if d64 else 32768
    max_len = 65538 if d64 else 258
    heads = {}
    out = []
    i = 0
    while i < n:
        best_len, best_dist = 0, 0
        if i + min_len <= n:
            key = data[i:i + 3]
            cands = heads.get(key)
            if cands:
                for p in reversed(cands[-max_chain:]):
                    dist = i - p
                    if dist > window:
                        break
                    limit = min(max_len, n - i)
                    k = 0
                    # compare in slabs (fast for long matches), overlapping matches handled bytewise
                    while k < limit and data[p + k] == data[i + k]:
                        k += 1
                        if k >= 32 and dist >= limit:
                            # both ranges are plain slices of data: extend with slice comparison
                            while k < limit:
                                step = min(4096, limit - k)
                                if data[p + k:p + k + step] == data[i + k:i + k + step]:
                                    k += step
                                else:
                                    break
                            while k < limit and data[p + k] == data[i + k]:
                                k += 1
                            break
                    if k > best_len:
                        best_len, best_dist = k, dist
                        if k >= limit:
                            break
        if best_len >= min_len:
            out.append((best_len, best_dist))
            step_end = i + best_len
            # index a bounded number of positions inside long matches
            stride = 1 if best_len <= 64 else best_len // 64
            for j in range(i, step_end, stride):
                if j + 3 <= n:
                    heads.setdefault(data[j:j + 3], []).append(j)
            i = step_end
        else:
            out.append(data[i])
            if i + 3 <= n:
                heads.setdefault(data[i:i + 3], []).append(i)
            i += 1
    return out


def huff_lengths(freq, maxbits):
    """Code lengths (<= maxbits) for the symbols with freq > 0: Huffman, then a Kraft-sum repair."""
    syms = [s for s, f in enumerate(freq) if f]
    lens = [0] * len(freq)
    if not syms:
        return lens
    if len(syms) == 1:
        lens[syms[0]] = 1
        return lens
    heap = [(freq[s], s, None, None) for s in syms]
    heapq.heapify(heap)
    nxt = len(freq)
    while len(heap) > 1:
        a = heapq.heappop(heap)
        b = heapq.heappop(heap)
        heapq.heappush(heap, (a[0] + b[0], nxt, a, b))
        nxt += 1
    stack = [(heap[0], 0)]
    while stack:
        node, d = stack.pop()
        if node[2] is None:
            lens[node[1]] = max(d, 1)
        else:
            stack.append((node[2], d + 1))
            stack.append((node[3], d + 1))
    # limit: clamp, then lengthen the cheapest codes until the Kraft sum fits, then shorten where possible
    over = [s for s in syms if lens[s] > maxbits]
    for s in over:
        lens[s] = maxbits
    unit = 1 << maxbits
    kraft = sum(unit >> lens[s] for s in syms)
    order = sorted(syms, key=lambda s: (freq[s], s))
    while kraft > unit:
        for s in order:
            if lens[s] < maxbits:
                kraft -= unit >> (lens[s] + 1)
                lens[s] += 1
                if kraft <= unit:
                    break
    for s in reversed(order):            # use up the slack so that the code is complete
        while lens[s] > 1 and kraft + (unit >> lens[s]) <= unit:
            kraft += unit >> lens[s]
            lens[s] -= 1
    assert kraft == unit, (kraft, unit)
    return lens


def canonical(lens):
    maxl = max(lens) if lens else 0
    cnt = [0] * (maxl + 2)
    for l in lens:
        cnt[l] += 1
    cnt[0] = 0
    code, nxt = 0, [0] * (maxl + 2)
    for l in range(1, maxl + 1):
        code = (code + cnt[l - 1]) << 1
        nxt[l] = code
    codes = [0] * len(lens)
    for s, l in enumerate(lens):
        if l:
            codes[s] = nxt[l]
            nxt[l] += 1
    return codes


def rle_lengths(lens):
    """RFC 1951 3.2.7 run-length coding of a code-length sequence: list of (sym, extra, extra_bits)."""
    out = []
    i, n = 0, len(lens)
    while i < n:
        v = lens[i]
        j = i
        while j < n and lens[j] == v:
            j += 1
        run = j - i
        if v == 0:
            while run >= 11:
                r = min(run, 138)
                out.append((18, r - 11, 7))
                run -= r
            if run >= 3:
                out.append((17, run - 3, 3))
                run = 0
            out.extend([(0, 0, 0)] * run)
        else:
            out.append((v, 0, 0))
            run -= 1
            while run >= 3:
                r = min(run, 6)
                out.append((16, r - 3, 2))
                run -= r
            out.extend([(v, 0, 0)] * run)
        i = j
    return out


def fixed_tables():
    ll = [8] * 144 + [9] * 112 + [7] * 24 + [8] * 8
    dl = [5] * 32
    return ll, dl


def write_block(bw, tokens, final, d64, kind="dynamic", raw=None, force285=False):
    """One block.  kind: dynamic | fixed | stored (stored needs raw = the block's plaintext, <= 65535 B)."""
    bw.bits(1 if final else 0, 1)
    if kind == "stored":
        bw.bits(0, 2)
        bw.align()
        bw.bits(len(raw), 16)
        bw.bits(len(raw) ^ 0xffff, 16)
        for b in raw:
            bw.bits(b, 8)
        return
    enc = []
    lf, df = [0] * 286, [0] * 32
    for t in tokens:
        if isinstance(t, tuple):
            ls, lx, lxb = length_code(t[0], d64, force285)
            ds, dx, dxb = dist_code(t[1], d64)
            lf[ls] += 1
            df[ds] += 1
            enc.append((ls, lx, lxb, ds, dx, dxb))
        else:
            lf[t] += 1
            enc.append(t)
    lf[256] += 1
    if kind == "fixed":
        bw.bits(1, 2)
        ll, dl = fixed_tables()
    else:
        bw.bits(2, 2)
        ll = huff_lengths(lf, 15)
        dl = huff_lengths(df, 15)
        nd_max = 32 if d64 else 30
        nlen = max(257, max(s for s in range(286) if ll[s]) + 1)
        ndist = max(1, max([s for s in range(nd_max) if dl[s]] or [0]) + 1)
        if not any(dl):
            dl[0] = 1                          # a lone unused one-bit code keeps the set legal in both modes
        seq = rle_lengths(ll[:nlen] + dl[:ndist])
        cf = [0] * 19
        for s, _, _ in seq:
            cf[s] += 1
        cl = huff_lengths(cf, 7)
        ncode = 19
        while ncode > 4 and cl[ORDER[ncode - 1]] == 0:
            ncode -= 1
        bw.bits(nlen - 257, 5)
        bw.bits(ndist - 1, 5)
        bw.bits(ncode - 4, 4)
        for i in range(ncode):
            bw.bits(cl[ORDER[i]], 3)
        cc = canonical(cl)
        for s, x, xb in seq:
            bw.code(cc[s], cl[s])
            if xb:
                bw.bits(x, xb)
    lc, dc = canonical(ll), canonical(dl)
    for e in enc:
        if isinstance(e, tuple):
            ls, lx, lxb, ds, dx, dxb = e
            bw.code(lc[ls], ll[ls])
            if lxb:
                bw.bits(lx, lxb)
            bw.code(dc[ds], dl[ds])
            if dxb:
                bw.bits(dx, dxb)
        else:
            bw.code(lc[e], ll[e])
    bw.code(lc[256], ll[256])


def compress(data, d64=True, block_tokens=20000, kinds=("dynamic",), force285=False, max_chain=16):
    """Raw Deflate64 (d64=True, for windowBits -16) or raw deflate stream of `data`.
    Blocks of about block_tokens tokens cycle through `kinds`."""
    tokens = lz77(data, d64, max_chain)
    bw = BitWriter()
    if not tokens:
        write_block(bw, [], True, d64, "fixed")
        return bw.done()
    pos, i, b = 0, 0, 0
    n = len(tokens)
    while i < n:
        kind = kinds[b % len(kinds)]
        j = min(n, i + block_tokens)
        if kind == "stored":
            size, k = 0, i
            while k < n and k < j:
                step = tokens[k][0] if isinstance(tokens[k], tuple) else 1
                if size + step > 65535:
                    break
                size += step
                k += 1
            if k == i:                           # a single match longer than a stored block can hold
                kind = "dynamic"
            else:
                j = k
        chunk = tokens[i:j]
        size = sum(t[0] if isinstance(t, tuple) else 1 for t in chunk)
        write_block(bw, chunk, j == n, d64, kind, raw=data[pos:pos + size] if kind == "stored" else None,
                    force285=force285)
        pos += size
        i = j
        b += 1
    return bw.done()
