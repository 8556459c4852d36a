"""GPU tier: the CUDA path through the C ABI (libzsb.so) against the oracle, bit for bit: output bytes,
ret, msg, total_in, total_out, check (strm->adler) and data_type.  Every group width is exercised."""
import ctypes
import hashlib

import pytest

import cases
import orclib
import zsblib

pytestmark = pytest.mark.gpu
CHK = orclib.ref() or orclib.orc()
# decoder configurations: 0 = the lane path (k_lane + k_lz), "big" = every member through k_big (a warp each,
# direct tables) + k_lz, else the group path with that many lanes
LANES = [0, "big", 32, 1, 4, 8, 2, 16]


def select(eng, lanes):
    if lanes == 0:
        eng.set_option("path", 1)
    elif lanes == "big":
        eng.set_option("path", 0)
        eng.set_option("split_big", 2)
    else:
        eng.set_option("path", 2)
        eng.set_option("lanes_per_member", lanes)


@pytest.fixture(scope="module")
def eng():
    e = zsblib.Zsb()
    yield e
    e.close()


def run_cases(eng, lanes, items):
    """items: (name, data, wbits, cap).  One batch call for all of them."""
    select(eng, lanes)
    got = eng.inflate_batch([d for _, d, _, _ in items], [w for _, _, w, _ in items], [c for _, _, _, c in items])
    for (name, data, wb, cap), (r, out) in zip(items, got):
        a, ao = CHK.oneshot(data, cap, wb)
        assert a.astuple() == r.astuple(), (lanes, name, wb, cap, a, r)
        assert ao == out, (lanes, name, "output differs")


@pytest.mark.parametrize("lanes", LANES)
def test_golden_fixtures(eng, lanes):
    items = []
    for name, data, m in cases.golden_fixtures():
        items.append((name, data, -16, m["out_len"] + 64))
        items.append((name + "_exact", data, -16, m["out_len"]))
        items.append((name + "_as_deflate", data, -15, m["out_len"]))
    select(eng, lanes)
    got = eng.inflate_batch([d for _, d, _, _ in items], [w for _, _, w, _ in items], [c for _, _, _, c in items])
    for (name, data, m), (r, out) in zip(cases.golden_fixtures(), got[::3]):
        assert (r.ret, r.total_in, r.total_out) == (1, m["in_len"], m["out_len"]), (name, r)
        assert hashlib.sha256(out).hexdigest() == m["out_sha256"], name
    run_cases(eng, lanes, items)


@pytest.mark.parametrize("lanes", LANES)
def test_probe_vectors(eng, lanes):
    items = [(n, d, w, c) for n, d, w, c, _ in cases.probe_vectors()]
    select(eng, lanes)
    got = eng.inflate_batch([d for _, d, _, _ in items], [w for _, _, w, _ in items], [c for _, _, _, c in items])
    for (name, _, _, _, want), (r, _) in zip(cases.probe_vectors(), got):
        assert (r.ret, r.msg_id, r.total_in, r.total_out) == want, (name, r)
    run_cases(eng, lanes, items)


@pytest.mark.parametrize("lanes", LANES)
def test_valid_streams(eng, lanes):
    items = []
    for name, data, wb, plain in cases.valid_streams():
        for cap in (len(plain) + 16, len(plain), max(len(plain) - 1, 0)):
            items.append((name, data, wb, cap))
    run_cases(eng, lanes, items)


@pytest.mark.parametrize("lanes", LANES)
def test_damaged_streams(eng, lanes):
    run_cases(eng, lanes, cases.damaged_streams(400 if lanes in (0, 1, 32) else 120))


def test_device_pointer_batch(eng):
    """Members already resident in HBM (torch tensors as plain device memory)."""
    import ctypes
    import torch
    select(eng, 0)
    vs = cases.valid_streams()[:60]
    blob = b"".join(d + b"\0" * (-len(d) % 16) for _, d, _, _ in vs)
    t_in = torch.frombuffer(bytearray(blob), dtype=torch.uint8).cuda()
    caps = [len(p) + 16 for _, _, _, p in vs]
    t_out = torch.zeros(sum((c + 15) // 16 * 16 for c in caps), dtype=torch.uint8, device="cuda")
    arr = (zsblib.Member * len(vs))()
    io = oo = 0
    offs = []
    for i, (name, d, wb, p) in enumerate(vs):
        arr[i].inp = t_in.data_ptr() + io
        arr[i].in_len = len(d)
        arr[i].out = t_out.data_ptr() + oo
        arr[i].out_cap = caps[i]
        arr[i].window_bits = wb
        offs.append(oo)
        io += len(d) + (-len(d) % 16)
        oo += (caps[i] + 15) // 16 * 16
    res = (zsblib.Result * len(vs))()
    torch.cuda.synchronize()
    rc = eng.lib.zsb_inflate_batch(eng.ctx, len(vs), arr, res)
    assert rc == 0, eng.error()
    host = t_out.cpu().numpy().tobytes()
    for i, (name, d, wb, p) in enumerate(vs):
        a, ao = CHK.oneshot(d, caps[i], wb)
        assert a.astuple() == res[i].astuple(), (name, a, res[i])
        assert host[offs[i]:offs[i] + len(p)] == p, name


def test_many_small_and_unaligned(eng):
    """Ragged batch: thousands of tiny members at odd host addresses, empty members included."""
    import random
    rnd = random.Random(1)
    items = []
    for k in range(3000):
        n = rnd.choice([0, 1, 5, 40, 200, 1500])
        plain = rnd.randbytes(n) if k % 3 else bytes([k & 255]) * n
        wb = rnd.choice([-15, 15, 31, -16])
        enc = -15 if wb == -16 else wb
        items.append(("tiny%d" % k, b"" if k % 97 == 0 else zsblib_gen().deflate(plain, rnd.choice([0, 1, 6, 9]), enc), wb, n + rnd.choice([0, 3])))
    for lanes in (0, "big", 1, 8, 32):
        run_cases(eng, lanes, items)


def test_call_level_edge_cases(eng):
    """Empty batches, NULL buffers, sizes beyond zlib's 32-bit counters, bad arguments: what the reference's
    entry points would answer (src/inflate.c:186-224 Z_STREAM_ERROR on bad parameters), never a crash."""
    select(eng, 0)
    lib = eng.lib
    assert lib.zsb_inflate_batch(eng.ctx, 0, None, None) == 0
    assert lib.zsb_inflate_batch(None, 1, None, None) == -100            # ZSB_E_ARG
    data = zsblib_gen().deflate(b"abc" * 100, 6, -15)
    ibuf = ctypes.create_string_buffer(data, len(data))
    obuf = ctypes.create_string_buffer(4 * 300)
    m = (zsblib.Member * 4)()
    r = (zsblib.Result * 4)()
    for i in range(4):
        m[i].inp, m[i].in_len, m[i].out, m[i].out_cap, m[i].window_bits = (
            ctypes.addressof(ibuf), len(data), ctypes.addressof(obuf) + 300 * i, 300, -15)
    m[1].in_len = 1 << 33                      # more than a uInt avail_in can say
    m[2].window_bits = -17                     # src/inflate.c:159-162
    m[3].inp = None                            # NULL next_in with avail_in != 0
    for path in (0, 2):
        eng.set_option("path", path)
        assert lib.zsb_inflate_batch(eng.ctx, 4, m, r) == 0
        assert (r[0].ret, r[0].total_out) == (1, 300) and obuf.raw[:300] == b"abc" * 100
        assert [r[i].ret for i in (1, 2, 3)] == [-2, -2, -2]
        assert [r[i].total_out for i in (1, 2, 3)] == [0, 0, 0]
    eng.set_option("path", 0)


def test_lane_path_slices(eng):
    """A sequence arena smaller than the batch needs: the lane path decodes it in slices."""
    items = []
    for name, data, wb, plain in cases.valid_streams()[::5]:
        items.append((name, data, wb, len(plain) + 16))
    select(eng, 0)
    old = eng.get_option("seq_arena_bytes")
    eng.set_option("seq_arena_bytes", 1 << 20)
    try:
        run_cases(eng, 0, items)
    finally:
        eng.set_option("seq_arena_bytes", old)


def zsblib_gen():
    import gen
    return gen


# ---- the z_stream surface (inflateInit2_ / inflate / inflateEnd of libzsb.so) ----
def zs_compare(api, name, data, wb, chunks, ocap):
    ta, oa = CHK.stream(data, wb, chunks, ocap, True, max_calls=200000)
    tb, ob = api.stream(data, wb, chunks, ocap, True, max_calls=200000)
    if ta != tb:
        k = next((i for i, (x, y) in enumerate(zip(ta, tb)) if x != y), min(len(ta), len(tb)))
        raise AssertionError((name, wb, chunks, ocap, "call", k, ta[max(k - 1, 0):k + 1], tb[max(k - 1, 0):k + 1]))
    assert oa == ob, (name, "output differs")


def test_zstream_wasm_pattern():
    """32 KiB in / 64 KiB out, Z_NO_FLUSH then Z_FINISH: what src/wasm/api/zlib-streams.js:73-138 does."""
    api = zsblib.ZStreamApi()
    items = [(n, d, -16) for n, d, m in cases.golden_fixtures()]
    items += [(n, d, w) for n, d, w, p in cases.valid_streams()[::7]]
    items += [(n, d, w) for n, d, w, c in cases.damaged_streams(60)[:60]]
    for name, data, wb in items:
        zs_compare(api, name, data, wb, [32768], 65536)


def test_zstream_random_chunkings():
    import random
    api = zsblib.ZStreamApi()
    rnd = random.Random(5)
    pool = [(n, d, w) for n, d, w, p in cases.valid_streams()[3::11]] + \
           [(n, d, w) for n, d, w, c in cases.damaged_streams(40)[:40:2]]
    for name, data, wb in pool:
        if len(data) > 30000:
            continue
        chunks = [rnd.choice([1, 3, 64, 1000, 32768]) for _ in range(4)]
        zs_compare(api, name, data, wb, chunks, rnd.choice([5, 300, 65536]))


def test_zstream_many_streams_per_call():
    """zsb_inflate_streams: hundreds of z_streams advance together, one call per tick;
    every stream's call-by-call trace and output must equal what the reference gives that stream alone."""
    import random
    api = zsblib.ZStreamApi()
    rnd = random.Random(9)
    jobs, names = [], []
    for n, d, m in cases.golden_fixtures():
        jobs.append((d, -16, [32768], 65536))
        names.append(n)
    for n, d, w, p in cases.valid_streams():
        jobs.append((d, w, [32768], 65536))
        names.append(n)
    for n, d, w, c in cases.damaged_streams(120)[:120]:
        jobs.append((d, w, [32768], 65536))
        names.append(n)
    for n, d, w, p in cases.valid_streams()[1::5]:           # odd chunkings and small output windows
        if len(d) <= 30000:
            jobs.append((d, w, [rnd.choice([1, 7, 500, 4096]) for _ in range(3)], rnd.choice([9, 700, 65536])))
            names.append(n + "/odd")
    got = api.stream_multi(jobs)
    assert len(got) == len(jobs) > 150
    for name, (data, wb, chunks, ocap), (tb, ob) in zip(names, jobs, got):
        ta, oa = CHK.stream(data, wb, chunks, ocap, True, max_calls=200000)
        if ta != tb:
            k = next((i for i, (x, y) in enumerate(zip(ta, tb)) if x != y), min(len(ta), len(tb)))
            raise AssertionError((name, wb, chunks, ocap, "call", k, ta[max(k - 1, 0):k + 1], tb[max(k - 1, 0):k + 1]))
        assert oa == ob, (name, "output differs")


def test_zstream_init_codes():
    api = zsblib.ZStreamApi()
    for wb in (-17, -16, -15, -8, -7, 0, 7, 8, 15, 16, 23, 24, 31, 47, 48, 63):
        assert api.init_ret(wb) == CHK.init_ret(wb), wb
    s = zsblib.ZStreamStruct()
    assert api.lib.inflateInit2_(ctypes.byref(s), 15, b"2.0", ctypes.sizeof(s)) == -6      # Z_VERSION_ERROR
    assert api.lib.inflateInit2_(ctypes.byref(s), 15, b"1.3", 100) == -6
    assert api.lib.inflateEnd(ctypes.byref(s)) == -2                                     # never initialised


def test_host_batch_pipelined(eng):
    """Host-pointer batches cut into chunks whose copies and decodes overlap (option "chunks")."""
    items = [(n, d, w, len(p) + 8) for n, d, w, p in cases.valid_streams()]
    items += cases.damaged_streams(100)[:400]
    try:
        for k in (3, 8, 40):
            eng.set_option("chunks", k)
            run_cases(eng, 0, items)
        eng.set_option("chunks", 5)              # the chunks' members through k_big
        run_cases(eng, "big", items)
        eng.set_option("split_big", 1)
        # a small sequence arena: the chunks share fewer regions / streams and decode in slices
        arena = eng.get_option("seq_arena_bytes")
        eng.set_option("seq_arena_bytes", 1 << 20)
        eng.set_option("chunks", 8)
        run_cases(eng, 0, items)
        eng.set_option("seq_arena_bytes", arena)
    finally:
        eng.set_option("chunks", 0)


def test_host_batch_member_bound_above_the_arena_option(eng):
    """A member whose own sequence bound exceeds "seq_arena_bytes" (the option's minimum, 4096 bytes) through the
    pipelined host path: its stream's region must still hold it (round-1 advisor finding: sequences were written
    past the region).  Incompressible 64 KiB members write close to their bound."""
    import random
    rnd = random.Random(77)
    items = []
    for i in range(24):
        plain = rnd.randbytes(65536) if i % 2 else zsblib_gen().word_soup(65536, 900 + i)
        items.append(("arena%d" % i, zsblib_gen().deflate(plain, 6 if i % 4 else 1, -15), -15, 65536))
    arena = eng.get_option("seq_arena_bytes")
    try:
        eng.set_option("seq_arena_bytes", 4096)
        for k in (1, 4):
            eng.set_option("chunks", k)
            run_cases(eng, 0, items)
    finally:
        eng.set_option("seq_arena_bytes", arena)
        eng.set_option("chunks", 0)


def test_zstream_block_and_trees_flush():
    """inflate(strm, Z_BLOCK) / inflate(strm, Z_TREES): the early returns and the 128 / 256 bits of data_type
    (reference src/inflate.c:832,854-857,883,1032,1269-1271), call by call against the oracle."""
    import random
    api = zsblib.ZStreamApi()
    rnd = random.Random(56)
    stops = 0
    pool = [(n, d, w) for n, d, w, p in cases.valid_streams()[2::9]] + [(n, d, -16) for n, d, m in cases.golden_fixtures()]
    for name, data, wb in pool:
        for flush in (5, 6):
            chunks = [len(data)] if rnd.random() < 0.5 else [rnd.choice([3, 64, 1000, 32768]) for _ in range(3)]
            ocap = rnd.choice([300, 65536, 1 << 20])
            ta, oa = CHK.stream(data, wb, chunks, ocap, False, max_calls=100000, flush=flush)
            tb, ob = api.stream(data, wb, chunks, ocap, False, max_calls=100000, flush=flush)
            assert ta == tb and oa == ob, (name, wb, flush, chunks, ocap)
            stops += sum(1 for c in tb if c[0] == 0 and (c[5] & 0x180))
    assert stops > 20


def test_pipelined_uniform_batch_stays_on_the_lane_path(eng):
    """The chunks of a pipelined host batch overlap, so large members are chosen for the batch as a whole: a
    uniform batch must not send whole chunks to k_big (that serialised the pipeline once: 96 -> 186 ms on C2)."""
    import gen
    select(eng, 0)
    eng.set_option("path", 0)
    eng.set_option("split_big", 1)
    data = gen.deflate(gen.word_soup(20000, 5), 6, -15)
    n, K = 4096, 4
    try:
        eng.set_option("chunks", K)
        got = eng.inflate_batch([data] * n, [-15] * n, [20000] * n)
        assert all(r.ret == 1 and r.total_out == 20000 for r, _ in got)
        assert eng.lib.zsb_last_launches(eng.ctx) == 2 * K          # k_lane + k_lz per chunk, no k_big
    finally:
        eng.set_option("chunks", 0)


def test_one_context_over_all_gpus_of_the_box():
    """zsb_create(devices, N) + ONE zsb_inflate_batch(): members sharded over the box's GPUs by zsb_plan_shards
    (no collective), results and bytes back in the caller's order.  Skipped on a one-GPU box."""
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs at least two GPUs")
    e = zsblib.Zsb(devices=list(range(ndev)))
    try:
        assert e.lib.zsb_num_devices(e.ctx) == ndev
        items = [(n, d, w, len(p) + 8) for n, d, w, p in cases.valid_streams()]
        items += cases.damaged_streams(200)[:300]
        for chunks in (0, 3):
            e.set_option("chunks", chunks)
            run_cases(e, 0, items * 4)
            run_cases(e, "big", items)
        # members already resident in the HBM of different GPUs: each is decoded by the device that owns it
        select(e, 0)
        e.set_option("path", 0)
        e.set_option("split_big", 1)
        keep, mem = [], (zsblib.Member * len(items))()
        for i, (name, data, wb, cap) in enumerate(items):
            dv = torch.device("cuda", i % ndev)
            ti = torch.frombuffer(bytearray(data) or bytearray(1), dtype=torch.uint8).to(dv)
            to = torch.zeros(max(cap, 1), dtype=torch.uint8, device=dv)
            keep.append((ti, to))
            mem[i].inp, mem[i].in_len, mem[i].out, mem[i].out_cap = ti.data_ptr(), len(data), to.data_ptr(), cap
            mem[i].window_bits, mem[i].flags = wb, 0
        res = (zsblib.Result * len(items))()
        torch.cuda.synchronize()
        rc = e.lib.zsb_inflate_batch(e.ctx, len(items), mem, res)
        assert rc == 0, e.error()
        for i, (name, data, wb, cap) in enumerate(items):
            a, ao = CHK.oneshot(data, cap, wb)
            assert a.astuple() == res[i].astuple(), (name, a, res[i])
            assert keep[i][1][:a.total_out].cpu().numpy().tobytes() == ao, name
    finally:
        e.close()

