#!/usr/bin/env python
"""Extra (non-headline) bench lines: BASELINE.json configs C3, C4, C5 on one B200, members resident in HBM,
every member checked against the oracle.  usage: bench_configs.py [C3 C4 C5] [--path 0|1|2] [--small]
One JSON line per config on stdout."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def zip_bench(args):
    """One archive of C5's members as ZIP entries (method 8 / 9), archive and outputs resident in HBM:
    zsb_zip_scan on the host copy, zsb_zip_inflate (decode + CRC-32 of every entry on the GPU)."""
    import ctypes
    import time
    import zlib
    import numpy as np
    import torch
    import workloads
    import zipgen
    import zsblib
    w = workloads.c5(distinct=2048, tiles=4 if args.small else 32)
    ents = []
    for t in range(w["tiles"]):
        for i, (comp, wb, plain) in enumerate(w["distinct"]):
            ents.append(dict(name="t%03d/f%05d" % (t, i), method=9 if wb == -16 else 8, data=comp, plain=plain,
                             crc=zlib.crc32(plain) & 0xffffffff))
    data = zipgen.build(ents, zip64=len(ents) >= 65535)
    lib = zsblib.load_product()
    harr = np.frombuffer(data, dtype=np.uint8)
    cnt = ctypes.c_size_t(0)
    zarr = (zsblib.ZipEntry * len(ents))()
    t0 = time.perf_counter()
    rc = lib.zsb_zip_scan(ctypes.c_void_p(harr.ctypes.data), len(data), zarr, len(ents), ctypes.byref(cnt))
    scan_ms = (time.perf_counter() - t0) * 1e3
    assert rc == 0 and cnt.value == len(ents)
    zents = list(zarr)
    n = len(zents)
    sizes = np.array([e.uncomp_size for e in zents], dtype=np.uint64)
    offs = np.concatenate(([0], np.cumsum(sizes)[:-1])).astype(np.uint64)
    total = int(sizes.sum())
    t_zip = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    t_out = torch.zeros(total, dtype=torch.uint8, device="cuda")
    earr = zarr
    res = (zsblib.Result * n)()
    eng = zsblib.Zsb(devices=[0])
    eng.set_option("path", args.path)

    def step():
        rc = lib.zsb_zip_inflate(eng.ctx, ctypes.c_void_p(t_zip.data_ptr()), earr, n, ctypes.c_void_p(t_out.data_ptr()),
                                 offs.ctypes.data, res)
        assert rc == 0, eng.error()
    step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    ms = (time.perf_counter() - t0) / args.steps * 1e3
    bad = [i for i in range(n) if res[i].ret != 1]
    exp = torch.frombuffer(bytearray(b"".join(m[2] for m in w["distinct"])), dtype=torch.uint8).cuda()
    B = exp.numel()
    same = all(bool(torch.equal(t_out[t * B:(t + 1) * B], exp)) for t in range(w["tiles"]))
    launches = int(lib.zsb_last_launches(eng.ctx))
    eng.close()
    return {"config": "ZIP", "workload": "one ZIP archive, %d entries (80%% method 8, 20%% method 9), 1-256 KiB, archive and "
            "outputs in HBM; decode + CRC-32 of every entry on the GPU" % n, "members": n, "archive_bytes": len(data),
            "out_bytes": total, "scan_ms_host": scan_ms, "ms_per_step": ms, "decompressed_gbs": total / ms / 1e6,
            "gpu_launches_per_step": launches, "timer": "host clock around the synchronous zsb_zip_inflate() call",
            "parity": "ok" if (not bad and same) else "MISMATCH", "mismatches": bad[:5], "path": args.path}


def streams_bench(args):
    """N z_streams fed the way the reference's JS wrapper feeds one (32 KiB in, 64 KiB out per call): all of them
    per tick through zsb_inflate_streams(), against inflate() stream by stream."""
    import ctypes
    import time
    import zlib
    import gen
    import zsblib
    api = zsblib.ZStreamApi()
    lib = api.lib
    lib.zsb_inflate_streams.argtypes = [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    n = 256 if args.small else 1024
    plain = b"".join(gen.word_soup(65536, 70 + i) for i in range(16))          # 1 MiB
    comp = gen.deflate(plain, 6, 15)
    cbuf = ctypes.create_string_buffer(comp, len(comp))

    def run(multi, count):
        ss = [zsblib.ZStreamStruct() for _ in range(count)]
        outs = [ctypes.create_string_buffer(65536) for _ in range(count)]
        for s in ss:
            assert lib.inflateInit2_(ctypes.byref(s), 15, b"1.3", ctypes.sizeof(s)) == 0
        ptrs = (ctypes.c_void_p * count)(*[ctypes.addressof(s) for s in ss])
        rets = (ctypes.c_int * count)()
        pos = [0] * count
        total, calls, crc = 0, 0, [0] * count
        live = list(range(count))
        t0 = time.perf_counter()
        while live:
            for i in live:
                s = ss[i]
                if s.avail_in == 0 and pos[i] < len(comp):
                    k = min(32768, len(comp) - pos[i])
                    s.next_in = ctypes.addressof(cbuf) + pos[i]
                    s.avail_in = k
                    pos[i] += k
                s.next_out = ctypes.addressof(outs[i])
                s.avail_out = 65536
            if multi:
                sub = (ctypes.c_void_p * len(live))(*[ptrs[i] for i in live])
                assert lib.zsb_inflate_streams(len(live), sub, 0, rets) == 0
                rr = [rets[k] for k in range(len(live))]
            else:
                rr = [lib.inflate(ctypes.byref(ss[i]), 0) for i in live]
            calls += 1 if multi else len(live)
            nxt = []
            for i, r in zip(live, rr):
                p = 65536 - ss[i].avail_out
                total += p
                crc[i] = zlib.crc32(outs[i].raw[:p], crc[i])
                assert r in (0, 1), r
                if r == 0:
                    nxt.append(i)
            live = nxt
        dt = time.perf_counter() - t0
        ok = all(c == zlib.crc32(plain) for c in crc) and total == count * len(plain)
        for s in ss:
            lib.inflateEnd(ctypes.byref(s))
        return dt, total, calls, ok
    dm, tm, cm, okm = run(True, n)
    ds, ts, cs, oks = run(False, max(8, n // 16))
    return {"config": "STREAMS", "workload": "%d zlib streams of 1 MiB (word soup, level 6), 32 KiB in / 64 KiB out per call "
            "(src/wasm/api/zlib-streams.js:73-83), driver in Python" % n, "streams": n,
            "multiplexed": {"seconds": dm, "out_bytes": tm, "calls": cm, "decompressed_gbs": tm / dm / 1e9, "parity": okm},
            "one_stream_per_call": {"streams": max(8, n // 16), "seconds": ds, "out_bytes": ts, "calls": cs,
                                    "decompressed_gbs": ts / ds / 1e9, "parity": oks},
            "parity": "ok" if okm and oks else "MISMATCH"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("configs", nargs="*", default=["C3", "C4", "C5"])
    ap.add_argument("--path", type=int, default=0, help="0 length-aware split (default), 1 lane path only, 2 group path only")
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--small", action="store_true")
    ap.add_argument("--no-split", action="store_true", help="option split_big = 0: no k_big for large members")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--arena-gib", type=int, default=0, help="option seq_arena_bytes (GiB); 0 = the default")
    args = ap.parse_args()
    import config_runner
    import workloads
    import zsblib
    full = {"C3": dict(distinct=8, tiles=512), "C4": dict(distinct=512, tiles=32), "C5": dict(distinct=2048, tiles=512)}
    small = {"C3": dict(distinct=8, tiles=32), "C4": dict(distinct=512, tiles=2), "C5": dict(distinct=2048, tiles=32)}
    for c in args.configs:
        if c == "ZIP":
            print(json.dumps(zip_bench(args)), flush=True)
            continue
        if c == "STREAMS":
            print(json.dumps(streams_bench(args)), flush=True)
            continue
        w = workloads.BUILDERS[c](**(small if args.small else full)[c])
        eng = zsblib.Zsb(devices=[0])
        eng.set_option("path", args.path)
        if args.no_split:
            eng.set_option("split_big", 0)
        if args.arena_gib:
            eng.set_option("seq_arena_bytes", args.arena_gib << 30)
        if args.lanes:
            eng.set_option("lanes_per_member", args.lanes)
        r = config_runner.run(w, eng, steps=args.steps, warmup=2)
        eng.close()
        r["config"] = c
        r["path"] = args.path
        r["parity"] = "ok" if not r["mismatches"] else "MISMATCH"
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
