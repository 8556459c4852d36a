#!/usr/bin/env python
"""One large member on one B200: a 16 MiB word-soup gzip member (level 6) resident in HBM,
decoded through zsb_inflate_batch_device() with and without the segmented path; MB/s of output, verified against the
plaintext.  Prints one JSON line."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import numpy as np
    import torch
    import batchlib
    import gen
    import zsblib
    dev = torch.device("cuda", 0)
    nbytes = 16 << 20
    plain = b"".join(gen.word_soup(65536, 9000 + i) for i in range(nbytes // 65536))
    out = {}
    for name, comp, wb in (("gzip level 6 (dynamic blocks)", gen.deflate(plain, 6, 31), 31),
                           ("zlib Z_FIXED (fixed blocks)", gen.deflate(plain, 6, 15, 4), 15)):
        w = {"distinct": [(comp, wb, plain)], "tiles": 1, "desc": name}
        b = batchlib.DeviceBatch(w, dev).to_device()
        st = torch.cuda.Stream(dev)
        res = {}
        for label, segment in (("segments", 1), ("serial (k_big)", 0)):
            eng = zsblib.Zsb(devices=[0])
            eng.set_option("segment", segment)
            if not segment:
                eng.set_option("split_big", 2)        # without segments the member gets a warp of its own
            for _ in range(2):
                b.step(eng, st)
            st.synchronize()
            b.clear_outputs()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 3
            e0.record(st)
            for _ in range(n):
                b.step(eng, st)
            e1.record(st)
            st.synchronize()
            ms = e0.elapsed_time(e1) / n
            bad = b.verify()
            assert not bad, bad
            res[label] = {"ms": ms, "output_mb_per_s": nbytes / ms / 1e3}
            eng.close()
        out[name] = {"compressed_bytes": len(comp), "output_bytes": nbytes, **res}
    print(json.dumps({"single_member": out}))


if __name__ == "__main__":
    main()
