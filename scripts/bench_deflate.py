#!/usr/bin/env python
"""Batched deflate (compression) throughput on one B200.

Workload: the C2 plaintexts — 65,536 x 64 KiB word-soup members, raw deflate — resident in HBM; value = input
GB/s through zsb_deflate_batch() (device pointers; the call's own descriptor copies and the host synchronisation
are inside the timed region).  Every compressed member is then inflated again on the GPU and compared with its
plaintext; the total compressed size is compared with zlib level 6 on the distinct plaintexts.  cpu_baseline =
zlib's deflate level 6 (the reference's deflate is upstream zlib's, src/zlib submodule @5a82f71, absent here: the
system zlib 1.3 stands in), one member per host thread.  Prints one JSON line."""
import argparse
import ctypes
import json
import os
import sys
import time
import zlib
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--members", type=int, default=65536)
    ap.add_argument("--distinct", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--wbits", type=int, default=-15)
    args = ap.parse_args()
    import numpy as np
    import torch
    import gen
    import zsblib
    from concurrent.futures import ProcessPoolExecutor
    dev = torch.device("cuda", 0)
    eng = zsblib.Zsb(devices=[0])
    lib = eng.lib
    K, n = args.distinct, args.members
    with ProcessPoolExecutor(min(os.cpu_count() or 1, 32)) as ex:
        plains = list(ex.map(gen.word_soup, [65536] * K, [1000 + i for i in range(K)], chunksize=16))
    blob = np.frombuffer(b"".join(plains), dtype=np.uint8)
    tiles = n // K
    t_in = torch.from_numpy(blob.copy()).to(dev).repeat(tiles)
    cap = lib.zsb_deflate_bound(65536, args.wbits)
    cap = (cap + 15) & ~15
    t_out = torch.zeros(n * cap, dtype=torch.uint8, device=dev)
    arr = (zsblib.DeflateMember * n)()
    a = np.frombuffer(arr, dtype=np.dtype([("inp", "<u8"), ("in_len", "<u8"), ("out", "<u8"), ("out_cap", "<u8"), ("wb", "<i4"), ("lv", "<i4")]))
    idx = np.arange(n, dtype=np.int64)
    a["inp"] = t_in.data_ptr() + idx * 65536
    a["in_len"] = 65536
    a["out"] = t_out.data_ptr() + idx * cap
    a["out_cap"] = cap
    a["wb"] = args.wbits
    a["lv"] = 6
    res = (zsblib.DeflateResult * n)()

    def step():
        rc = lib.zsb_deflate_batch(eng.ctx, n, arr, res)
        if rc:
            raise RuntimeError("zsb_deflate_batch rc=%d %s" % (rc, eng.error()))
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / args.steps * 1e3
    r = np.frombuffer(res, dtype=np.dtype([("ret", "<i4"), ("blocks", "<i4"), ("tin", "<u8"), ("tout", "<u8"), ("chk", "<u4"), ("rsv", "<u4")]))
    assert (r["ret"] == 1).all() and (r["tin"] == 65536).all()
    sum_in, sum_out = n * 65536, int(r["tout"].sum())
    # round trip on the GPU: inflate every member, compare with the plaintext tiles
    t_back = torch.zeros(n * 65536, dtype=torch.uint8, device=dev)
    mem = (zsblib.Member * n)()
    m = np.frombuffer(mem, dtype=np.dtype([("inp", "<u8"), ("in_len", "<u8"), ("out", "<u8"), ("out_cap", "<u8"), ("wb", "<i4"), ("fl", "<i4")]))
    m["inp"] = a["out"]
    m["in_len"] = r["tout"]
    m["out"] = t_back.data_ptr() + idx * 65536
    m["out_cap"] = 65536
    m["wb"] = args.wbits
    ires = (zsblib.Result * n)()
    assert lib.zsb_inflate_batch(eng.ctx, n, mem, ires) == 0, eng.error()
    ir = np.frombuffer(ires, dtype=np.dtype([("ret", "<i4"), ("msg", "<i4"), ("tin", "<u8"), ("tout", "<u8"), ("chk", "<u4"), ("dt", "<i4")]))
    badm = np.nonzero((ir["ret"] != 1) | (ir["tout"] != 65536))[0]
    if len(badm):
        j = int(badm[0])
        comp = t_out[j * cap:j * cap + int(r["tout"][j])].cpu().numpy().tobytes()
        open(os.path.join(ROOT, "gpurun_out", "deflate_bad_member.bin"), "wb").write(comp)
        open(os.path.join(ROOT, "gpurun_out", "deflate_bad_plain.bin"), "wb").write(plains[j % K])
        raise SystemExit("round trip: %d of %d members fail to inflate; first %d: deflate %r inflate %r" % (
            len(badm), n, j, r[j], ir[j]))
    assert bool((t_back == t_in).all()), "round trip differs"
    # ratio against zlib level 6 on the distinct plaintexts
    with ThreadPoolExecutor(os.cpu_count() or 1) as ex:
        t1 = time.perf_counter()
        ref_sizes = list(ex.map(lambda p: len(zlib.compress(p, 6, wbits=args.wbits)), plains))
        cpu_s = time.perf_counter() - t1
    ours_distinct = int(r["tout"][:K].sum())
    peak = 6531.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    out = {"metric": "batched deflate input GB/s", "value": sum_in / ms / 1e6, "unit": "GB/s", "n_gpus": 1, "steps": args.steps,
           "ms_per_step": ms, "dtype": "u8", "data": "synthetic",
           "config": {"workload": "C2 plaintexts: %d x 64 KiB word-soup members, windowBits %d, device-resident" % (n, args.wbits),
                      "distinct_members": K, "verified": "every member inflated again on the GPU and compared with its plaintext"},
           "compressed_bytes": sum_out, "ratio": sum_in / sum_out,
           "size_vs_zlib_level6": ours_distinct / sum(ref_sizes),
           "roofline": {"bound": "hbm", "achieved": (sum_in + sum_out) / ms / 1e6, "peak": peak, "unit": "GB/s",
                        "frac": (sum_in + sum_out) / ms / 1e6 / peak, "kernel": "k_deflate+k_dstitch", "traffic": None},
           "cpu_baseline": {"value": K * 65536 / cpu_s / 1e9, "unit": "GB/s", "cores": os.cpu_count(), "kind": "zlib 1.3 deflate level 6 (stand-in for the absent src/zlib submodule)",
                            "sample": "%d distinct members, one per thread" % K},
           "gpu_launches": int(lib.zsb_last_launches(eng.ctx)) * args.steps}
    print(json.dumps(out))
    eng.close()


if __name__ == "__main__":
    main()
