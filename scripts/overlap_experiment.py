#!/usr/bin/env python
"""Experiment: does k_lz's work fit into the issue slots k_lane leaves idle?

The C2 batch is cut into K parts; part i is decoded by its own context (own scratch) on its own CUDA stream with
k_lane in full-size CTAs (one per SM: k_lz's CTAs fit beside it, the next part's k_lane takes the SM when this
part's leaves) or in small CTAs (option small_ctas: 2-warp CTAs, six per SM), so that k_lz of one part runs beside
k_lane of the next on the same SMs — the overlap a fused kernel with resolver warps would give, without the
fusion.  CAVEAT (found afterwards): the timed steps are queued back to back, so with K > 1 a step also overlaps the
NEXT step; inside one zsb_inflate_batch_device() call the same parts give no gain (profiles/r02_bench_overlap_parts8.json:
30.8 ms).  Prints ms per step for K = 1 (one launch pair, full-size CTAs: the product's default), and K = 2 .. 16."""
import ctypes
import json
import os
import sys

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import numpy as np
    import torch
    import batchlib
    import workloads
    import zsblib
    dev = torch.device("cuda", 0)
    w = workloads.c2(distinct=1024, tiles=64)
    exp = batchlib.expected(w["distinct"])
    n = 1024 * 64
    out = {}
    for K, small in ((1, 0), (2, 0), (4, 0), (8, 0), (16, 0), (2, 1), (4, 1), (8, 1), (16, 1)):
        engs = [zsblib.Zsb(devices=[0]) for _ in range(K)]
        streams = [torch.cuda.Stream(dev) for _ in range(K)]
        parts = []
        for i in range(K):
            members = np.arange(i, n, K, dtype=np.int64)          # every part has the same mix of members
            parts.append(batchlib.DeviceBatch(w, dev, members=members, exp=exp).to_device())
            if small:
                engs[i].set_option("small_ctas", 1)
        def step():
            for i in range(K):
                parts[i].step(engs[i], streams[i])
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        for p in parts:
            p.clear_outputs()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 10
        e0.record(torch.cuda.current_stream())
        for s in streams:
            s.wait_event(e0)
        for _ in range(steps):
            step()
        for s in streams:
            ev = torch.cuda.Event()
            ev.record(s)
            torch.cuda.current_stream().wait_event(ev)
        e1.record(torch.cuda.current_stream())
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        bad = [b for p in parts for b in p.verify()]
        assert not bad, bad[:3]
        out["K=%d%s" % (K, " small CTAs" if small else "")] = {"ms_per_step": ms, "decompressed_gbs": n * 65536 / ms / 1e6, "members_verified": n}
        del parts
        for e in engs:
            e.close()
        torch.cuda.empty_cache()
    print(json.dumps({"experiment": "k_lane / k_lz overlap through K concurrent part-batches with small CTAs (C2, one B200)",
                      "results": out}))


if __name__ == "__main__":
    main()
